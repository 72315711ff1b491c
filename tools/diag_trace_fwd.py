"""GPU development tool (not a test): timeline of one flash-forward CTA from the -DSPB_FA_TRACE library
(tools/build_trace_lib.sh).  Prints, for CTA (0,0,0) (heaviest causal tile pair), the events of each role in clock order."""
import ctypes as C, math, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("SPB_LIB_PATH", os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "solvingpapers_b200", "_ab", "libspb200_trace.so"))
import torch
from solvingpapers_b200 import _lib as L, ops

NAMES = {1: "prod k_empty ok", 2: "prod v_empty ok", 10: "mma QK0 issued", 11: "mma QK1 issued", 12: "mma p_full0 seen", 13: "mma p_full1 seen",
         14: "mma PV0 issued", 15: "mma PV1 issued", 16: "mma v_full seen", 17: "mma k_full seen", 20: "sm wait S0", 21: "sm wait S1",
         22: "sm S0 seen", 23: "sm S1 seen", 24: "sm S in regs", 25: "sm local max", 26: "sm exchanged", 27: "sm exp done", 28: "sm P stored", 29: "sm arrived"}
B, T, Hq, Hkv, hd = 4, 2048, 32, 8, 128
causal = os.environ.get("FA_CAUSAL", "1") == "1"
g = torch.Generator(device="cuda").manual_seed(1)
mk = lambda h: torch.randn(B, T, h, hd, generator=g, device="cuda").bfloat16()
q, k, v = mk(Hq), mk(Hkv), mk(Hkv)
fn = lambda: ops.flash_fwd(q, k, v, causal, 1 / math.sqrt(hd), impl=L.IMPL_TCGEN05)
for _ in range(3): fn()
torch.cuda.synchronize()
lib = L.load()
lib.spb_debug_fa_trace_clear()
fn(); torch.cuda.synchronize()
n = 4 * 2048
buf = (C.c_ulonglong * n)()
rc = lib.spb_debug_fa_trace_read(buf, n)
assert rc == 0, rc
ev = []
for role in range(4):
    for i in range(2048):
        w = buf[role * 2048 + i]
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
        print(f"{rel:8d}  (+{d:5d})  {'  ' * role * 6}r{role} {NAMES.get(e, e)}")
    last[role] = rel
