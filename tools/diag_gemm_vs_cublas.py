"""GPU diagnostic (not a test, not a bench line): the six GEMMs of a C3 step, each timed alone through spb_gemm (tcgen05
kernels of this repo) and through torch.matmul (cuBLAS) on the same operands - the library number is the ceiling the repo's
`roofline.peak` stands for, measured on the step's own shapes instead of a square 8192^3.  Rotating operand copies keep the
126 MB L2 from serving a repeat."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from solvingpapers_b200 import _lib as L
from solvingpapers_b200 import ops

N_TOK, C, NTOT, NQ = 8192, 4096, 6144, 4096   # C3: B 4 x T 2048 tokens
# name, M, N, K, transA, transB, out dtype
SHAPES = [
    ("gemm_qkv   x.Wqkv", N_TOK, NTOT, C, False, False, torch.bfloat16),
    ("gemm_out   a.Wo", N_TOK, C, NQ, False, False, torch.bfloat16),
    ("gemm_dattn dy.Wo^T", N_TOK, NQ, C, False, True, torch.bfloat16),
    ("gemm_dwo   a^T.dy (fp32 out)", NQ, C, N_TOK, True, False, torch.float32),
    ("gemm_dx    dqkv.Wqkv^T", N_TOK, C, NTOT, False, True, torch.bfloat16),
    ("gemm_dwqkv xhat^T.dqkv (fp32 out)", C, NTOT, N_TOK, True, False, torch.float32),
]


# short-K shapes of the other configs (SHAPES=c5 / c4): where the epilogue, not the main loop, sets the pace
if os.environ.get("SHAPES") == "c5":   # ViT-B/16: 256 x 197 tokens, d 768
    N_TOK = 256 * 197
    SHAPES = [("c5 gemm_qkv", N_TOK, 2304, 768, False, True, torch.bfloat16), ("c5 gemm_out", N_TOK, 768, 768, False, True, torch.bfloat16),
              ("c5 gemm_dx", N_TOK, 768, 2304, False, False, torch.bfloat16), ("c5 gemm_dwqkv (fp32 out)", 2304, 768, N_TOK, True, False, torch.float32)]
elif os.environ.get("SHAPES") == "c4":  # Gemma-2B-shaped: 16384 tokens, d 2048, 8 q heads + 1 kv head of 256
    N_TOK = 16384
    SHAPES = [("c4 gemm_qkv", N_TOK, 2560, 2048, False, False, torch.bfloat16), ("c4 gemm_out", N_TOK, 2048, 2048, False, False, torch.bfloat16),
              ("c4 gemm_dx", N_TOK, 2048, 2560, False, True, torch.bfloat16), ("c4 gemm_dwqkv (fp32 out)", 2048, 2560, N_TOK, True, False, torch.float32)]


def timed(fn, reps=10):
    for _ in range(3):
        fn(0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(reps):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    torch.manual_seed(0)
    out = []
    for name, M, N, K, tA, tB, od in SHAPES:
        ncopy = 3
        As = [torch.randn((K, M) if tA else (M, K), device="cuda").bfloat16() for _ in range(ncopy)]
        Bs = [(torch.randn((N, K) if tB else (K, N), device="cuda") / K ** 0.5).bfloat16() for _ in range(ncopy)]
        D = torch.empty((M, N), dtype=od, device="cuda")
        Dl = torch.empty((M, N), dtype=torch.bfloat16, device="cuda")   # cuBLAS through torch: bf16 in -> bf16 out

        def ours(i):
            ops.gemm(As[i % ncopy], Bs[i % ncopy], transA=tA, transB=tB, out=D, impl=L.IMPL_TCGEN05)

        def lib(i):
            a = As[i % ncopy].t() if tA else As[i % ncopy]
            b = Bs[i % ncopy].t() if tB else Bs[i % ncopy]
            torch.matmul(a, b, out=Dl)

        t_o, t_l = timed(ours), timed(lib)
        fl = 2.0 * M * N * K / 1e12
        rec = {"gemm": name, "M": M, "N": N, "K": K, "ours_ms": round(t_o, 4), "ours_tflops": round(fl / (t_o * 1e-3), 1),
               "cublas_ms": round(t_l, 4), "cublas_tflops": round(fl / (t_l * 1e-3), 1), "ours_over_cublas": round(t_l / t_o, 3),
               "note": "cuBLAS writes bf16 here; ours writes fp32" if od == torch.float32 else ""}
        out.append(rec)
        print(json.dumps(rec), flush=True)
        del As, Bs, D, Dl
        torch.cuda.empty_cache()
    so, sl = sum(r["ours_ms"] for r in out), sum(r["cublas_ms"] for r in out)
    print(json.dumps({"sum_ours_ms": round(so, 4), "sum_cublas_ms": round(sl, 4), "ours_over_cublas": round(sl / so, 3)}))


if __name__ == "__main__":
    main()
