"""GPU development tool (not a test): timeline of one flash-backward CTA from a -DSPB_FB_TRACE library
(tools/build_variant.sh trace -DSPB_FA_TRACE -DSPB_FB_TRACE).  CTA (0,0,0) = KV tile 0 of the first group (heaviest)."""
import ctypes as C, math, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("SPB_LIB_PATH", os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "solvingpapers_b200", "_ab", "libspb200_trace.so"))
import torch
from solvingpapers_b200 import _lib as L, ops

NAMES = {2: "prod q_empty ok (hd256 role K)", 16: "mma do_full seen", 17: "mma dP issued", 26: "ew exps done", 27: "ew dp_full seen",
         1: "prod qdo_empty / do_empty ok", 10: "mma qdo_full seen", 11: "mma sdp_empty seen", 12: "mma S,dP issued", 13: "mma pds_full seen", 14: "mma dq_empty seen",
         15: "mma dV,dK,dQ issued", 20: "ew bar", 21: "ew sdp_full seen", 22: "ew S,dP in regs", 23: "ew math done", 24: "ew pds_empty seen",
         25: "ew P,dS stored+arrived", 30: "dr dq_full seen", 31: "dr dQ in regs", 32: "dr staged", 33: "dr reduce issued"}
B, T, Hq, Hkv, hd = (int(os.environ.get(k, d)) for k, d in (('TR_B', 4), ('TR_T', 2048), ('TR_HQ', 32), ('TR_HKV', 8), ('TR_HD', 128)))
causal = os.environ.get("FA_CAUSAL", "1") == "1"
g = torch.Generator(device="cuda").manual_seed(1)
mk = lambda h: torch.randn(B, T, h, hd, generator=g, device="cuda").bfloat16()
q, k, v, do = mk(Hq), mk(Hkv), mk(Hkv), mk(Hq)
o, lse = ops.flash_fwd(q, k, v, causal, 1 / math.sqrt(hd), impl=L.IMPL_TCGEN05)
fn = lambda: ops.flash_bwd(q, k, v, o, do, lse, causal, 1 / math.sqrt(hd), impl=L.IMPL_TCGEN05)
for _ in range(3): fn()
torch.cuda.synchronize()
lib = L.load()
lib.spb_debug_fb_trace_clear()
fn(); torch.cuda.synchronize()
NE = 4096
buf = (C.c_ulonglong * (4 * NE))()
rc = lib.spb_debug_fb_trace_read(buf, 4 * NE)
assert rc == 0, rc
ev = []
for role in range(4):
    for i in range(NE):
        w = buf[role * NE + i]
        if w == 0: break
        ev.append((w >> 8, role, w & 0xFF))
ev.sort()
t0 = ev[0][0]
lo, hi = int(os.environ.get("TR_LO", "0")), int(os.environ.get("TR_HI", "1000000000"))
print(f"{len(ev)} events, span {ev[-1][0]-t0} clk")
last = {}
for t, role, e in ev:
    rel = t - t0
    if lo <= rel <= hi:
        d = rel - last.get(role, rel)
        print(f"{rel:8d}  (+{d:5d})  {' ' * role * 14}r{role} {NAMES.get(e, e)}")
    last[role] = rel
