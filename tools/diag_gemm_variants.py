"""GPU timing experiment: one GEMM shape, every operand-major combination and output type."""
import math, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from solvingpapers_b200 import _lib as L, ops
M, N, K = 4096, 6144, 8192
def t(fn, n=20):
    for _ in range(3): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
for tA in (0, 1):
    for tB in (0, 1):
        A = torch.randn((K, M) if tA else (M, K), device="cuda").bfloat16()
        B = torch.randn((N, K) if tB else (K, N), device="cuda").bfloat16()
        for od in (torch.bfloat16, torch.float32):
            out = torch.empty((M, N), device="cuda", dtype=od)
            ms = t(lambda: ops.gemm(A, B, transA=bool(tA), transB=bool(tB), out=out, impl=L.IMPL_TCGEN05))
            print(f"M{M} N{N} K{K} tA{tA} tB{tB} out={str(od)[6:]:8s}: {ms*1e3:7.1f} us {2.0*M*N*K/ms/1e9:7.1f} TF/s")
