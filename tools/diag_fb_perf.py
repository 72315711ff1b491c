"""GPU timing (not a test): tcgen05 flash forward / backward stand-alone at the C3 and C4 head geometries, CUDA events."""
import math, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from solvingpapers_b200 import _lib as L, ops

def run(B, T, Hq, Hkv, hd, causal=True, iters=20):
    g = torch.Generator(device="cuda").manual_seed(1)
    mk = lambda h: torch.randn(B, T, h, hd, generator=g, device="cuda").bfloat16()
    q, k, v, do = mk(Hq), mk(Hkv), mk(Hkv), mk(Hq)
    sc = 1 / math.sqrt(hd)
    o, lse = ops.flash_fwd(q, k, v, causal, sc, impl=L.IMPL_TCGEN05)
    res = {}
    for name, fn in (("fwd", lambda: ops.flash_fwd(q, k, v, causal, sc, impl=L.IMPL_TCGEN05)),
                     ("bwd", lambda: ops.flash_bwd(q, k, v, o, do, lse, causal, sc, impl=L.IMPL_TCGEN05))):
        for _ in range(3): fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters): fn()
        e1.record(); torch.cuda.synchronize()
        res[name] = e0.elapsed_time(e1) / iters
    fa = 4.0 * B * Hq * T * T * hd * ((T + 1) / (2.0 * T) if causal else 1.0)
    print(f"{os.path.basename(L.LIB_PATH):28s} B{B} T{T} {Hq}/{Hkv}x{hd} causal{int(causal)}: fwd {res['fwd']:.4f} ms ({fa/res['fwd']/1e9:.0f} TF/s)  "
          f"bwd {res['bwd']:.4f} ms ({2.5*fa/res['bwd']/1e9:.0f} TF/s)", flush=True)

if __name__ == "__main__":
    run(4, 2048, 32, 8, 128)
    run(4, 4096, 8, 1, 256)
    run(16, 1024, 12, 12, 64, causal=False)


def run_wide(B, T, Hq, Hkv, hd, causal=True, iters=10):
    g = torch.Generator(device="cuda").manual_seed(1)
    mk = lambda h: torch.randn(B, T, h, hd, generator=g, device="cuda").bfloat16()
    q, k, v, do = mk(Hq), mk(Hkv), mk(Hkv), mk(Hq)
    sc = 1 / math.sqrt(hd)
    o, lse = ops.attn_wide_fwd(q, k, v, causal, sc)
    res = {}
    for name, fn in (("fwd", lambda: ops.attn_wide_fwd(q, k, v, causal, sc)), ("bwd", lambda: ops.attn_wide_bwd(q, k, v, o, do, lse, causal, sc))):
        for _ in range(2): fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters): fn()
        e1.record(); torch.cuda.synchronize()
        res[name] = e0.elapsed_time(e1) / iters
    fa = 4.0 * B * Hq * T * T * hd * ((T + 1) / (2.0 * T) if causal else 1.0)
    print(f"WIDE (GEMM pipeline)         B{B} T{T} {Hq}/{Hkv}x{hd} causal{int(causal)}: fwd {res['fwd']:.4f} ms ({fa/res['fwd']/1e9:.0f} TF/s)  "
          f"bwd {res['bwd']:.4f} ms ({2.5*fa/res['bwd']/1e9:.0f} TF/s)", flush=True)


if __name__ == "__main__" and os.environ.get("WIDE"):
    run_wide(4, 4096, 8, 1, 256)
    run_wide(1, 4096, 2, 1, 768)
