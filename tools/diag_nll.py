"""GPU timing (not a test): spb_nll_fwd / spb_nll_bwd bandwidth at N = 8192 rows, V = 50257 (row pitch 50264), bf16 and fp32."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from solvingpapers_b200 import llama as spl

for dtype in (torch.bfloat16, torch.float32):
    N, V, ld = 8192, 50257, 50264
    buf = (torch.randn(N, ld, device="cuda") * 3).to(dtype)
    x = buf[:, :V].detach().requires_grad_(True)
    tg = torch.randint(0, V, (N,), device="cuda")
    e = buf.element_size()
    for name, nbytes in (("fwd", N * V * e), ("fwd+bwd", 3 * N * V * e)):
        def step():
            loss = spl.nll_loss(x, tg)
            if name != "fwd":
                loss.backward()
                x.grad = None
        for _ in range(3):
            step()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            step()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        print(f"nll {name:8s} {str(dtype)[6:]:9s} N{N} V{V}: {ms:.3f} ms  {nbytes/ms/1e6:.0f} GB/s algorithmic")

# whole loss head at the notebook's dims (llama3:350-356: d = 256, V = 50257), N = 8192 tokens, bf16: unfused (logits GEMM with the
# vocabulary padded by the caller + nll kernels + two gradient GEMMs) vs the fused head (spb_loss_head_fwd / bwd)
from solvingpapers_b200 import ops
N, C, V, Vp = 8192, 256, 50257, 50264
h = torch.randn(N, C, device="cuda").bfloat16()
wpad = torch.zeros(C, Vp, device="cuda", dtype=torch.bfloat16)
wpad[:, :V] = (torch.randn(C, V, device="cuda") / 16).bfloat16()
w = wpad[:, :V].contiguous()
tg = torch.randint(0, V, (N,), device="cuda")
def unfused():
    hh, ww = h.clone().requires_grad_(True), wpad.clone().requires_grad_(True)
    logits = ops.matmul(hh, ww)[:, :V]
    spl.nll_loss(logits, tg).backward()
def fused():
    hh, ww = h.clone().requires_grad_(True), w.clone().requires_grad_(True)
    spl._LossHeadFn.apply(hh, ww, tg).backward()
for name, fn in (("unfused", unfused), ("fused", fused)):
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        fn()
    e1.record(); torch.cuda.synchronize()
    print(f"loss head fwd+bwd {name:8s} bf16 N{N} C{C} V{V}: {e0.elapsed_time(e1)/10:.3f} ms")
