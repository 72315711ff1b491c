"""GPU timing (not a test): the ViT-B/16 QKV GEMM (50432 x 2304 x 768, torch [out,in] weight) through spb_gemm without / with the
bias epilogue, and torch.addmm (cuBLAS) on the same operands - the shape on which the epilogue, not the main loop, sets the pace."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from solvingpapers_b200 import _lib as L, ops
M,N,K=50432,2304,768
A=[torch.randn(M,K,device="cuda").bfloat16() for _ in range(3)]
B=[(torch.randn(N,K,device="cuda")/K**0.5).bfloat16() for _ in range(3)]
bias=torch.randn(N,device="cuda").bfloat16()
D=torch.empty(M,N,dtype=torch.bfloat16,device="cuda")
def t(fn,reps=20):
    for i in range(3): fn(i)
    torch.cuda.synchronize()
    e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(reps): fn(i)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1)/reps
print("no bias", t(lambda i: ops.gemm(A[i%3],B[i%3],transB=True,out=D,impl=L.IMPL_TCGEN05)))
print("bias   ", t(lambda i: ops.gemm(A[i%3],B[i%3],transB=True,out=D,bias=bias,impl=L.IMPL_TCGEN05)))
print("cublas addmm", t(lambda i: torch.addmm(bias, A[i%3], B[i%3].t(), out=D)))
