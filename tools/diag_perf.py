"""GPU diagnostic (not a test): time the individual kernels of the C3 step in isolation (CUDA events, L2 flushed by
rotating through buffers larger than L2)."""
import ctypes as C
import math
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import solvingpapers_b200 as sp
from solvingpapers_b200 import _lib as L
from solvingpapers_b200 import ops


def timeit(fn, n=20, warm=3):
    for _ in range(warm):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def gemm_case(name, M, N, K, tA, tB, out_dtype=torch.bfloat16):
    A = torch.randn((K, M) if tA else (M, K), device="cuda").bfloat16()
    B = (torch.randn((N, K) if tB else (K, N), device="cuda") / math.sqrt(K)).bfloat16()
    out = torch.empty((M, N), device="cuda", dtype=out_dtype)
    ms = timeit(lambda: ops.gemm(A, B, transA=tA, transB=tB, out=out, impl=L.IMPL_TCGEN05))
    ref = timeit(lambda: torch.matmul(A.t() if tA else A, B.t() if tB else B))
    fl = 2.0 * M * N * K
    print(f"gemm {name:8s} M{M} N{N} K{K} tA{int(tA)} tB{int(tB)} {str(out_dtype)[6:]:8s}: {ms*1e3:7.1f} us {fl/ms/1e9:7.1f} TF/s | cuBLAS {ref*1e3:7.1f} us {fl/ref/1e9:7.1f} TF/s")


if __name__ == "__main__":
    torch.manual_seed(0)
    N, Cm, Nq, Ntot = 8192, 4096, 4096, 6144
    gemm_case("qkv", N, Ntot, Cm, False, False)
    gemm_case("out", N, Cm, Nq, False, False)
    gemm_case("dattn", N, Nq, Cm, False, True)
    gemm_case("dx", N, Cm, Ntot, False, True)
    gemm_case("dwo", Nq, Cm, N, True, False, torch.float32)
    gemm_case("dwqkv", Cm, Ntot, N, True, False, torch.float32)
    gemm_case("dwo-bf", Nq, Cm, N, True, False)            # same operands, bf16 output: isolates the fp32 epilogue / tail split
    gemm_case("dwqkv-bf", Cm, Ntot, N, True, False)
    gemm_case("mn-k", Nq, Cm, N, True, True)               # A MN-major, B K-major
    gemm_case("k-mn", Nq, Cm, N, False, False)             # A K-major, B MN-major (the forward's majors, dW shape)
    if "--gemm-only" in sys.argv:
        sys.exit(0)
    gemm_case("sq8k", 8192, 8192, 8192, False, True)
    # block-level with and without the fused rotary (rope=False -> no rotary at all, shows the epilogue's cost)
    for rope in (True, False):
        cfg = sp.BlockConfig(4, 2048, 4096, 32, 8, 128, dtype=torch.bfloat16, rope=rope)
        blk = sp.AttentionBlock(cfg)
        p = {"wqkv": (torch.randn(Cm, Ntot, device="cuda") / 64).bfloat16(), "wo": (torch.randn(Nq, Cm, device="cuda") / 64).bfloat16(),
             "norm_w": torch.randn(Cm, device="cuda").bfloat16()}
        if rope:
            ang = torch.outer(torch.arange(2048.0), 1.0 / (10000.0 ** (torch.arange(0, 64.0) / 128)))
            p["rope"] = sp.llama_rope_table(torch.cos(ang), torch.sin(ang)).cuda()
        x = torch.randn(4, 2048, Cm, device="cuda").bfloat16()
        dy = torch.randn(4, 2048, Cm, device="cuda").bfloat16()
        g = blk.alloc_grads(p)
        g["dx"] = torch.empty_like(x)
        lib = L.load()
        lib.spb_prof_enable(20)

        def step():
            blk.forward(x, p)
            blk.backward(dy, grads=g)

        for _ in range(3):
            step()
        lib.spb_prof_enable(20)
        ms = timeit(step, n=20, warm=0)
        names = (C.c_char_p * 64)()
        msb = (C.c_float * 64)()
        agg = {}
        for s in range(lib.spb_prof_steps()):
            n = lib.spb_prof_read(s, 64, names, msb)
            for i in range(max(n, 0)):
                agg[names[i].decode()] = agg.get(names[i].decode(), 0.0) + msb[i] / 20
        lib.spb_prof_enable(0)
        print(f"block rope={rope}: {ms:.3f} ms/step ", {k: round(v, 3) for k, v in agg.items() if v > 0.01})
    # norm kernels
    x = torch.randn(N, Cm, device="cuda").bfloat16()
    w = torch.randn(Cm, device="cuda").bfloat16()
    y, st = ops.norm_fwd(x, w)
    ms = timeit(lambda: ops.norm_fwd(x, w))
    print(f"norm_fwd (incl. torch.empty): {ms*1e3:.1f} us -> {2*x.numel()*2/ms/1e6:.0f} GB/s")
