"""GPU timing (not a test): the feed-forward sub-block at the C3 shape - N = 4 x 2048 tokens, d 4096, hidden 16384,
bf16 - forward + backward through spb_ffn_fwd / spb_ffn_bwd, next to the same op sequence in torch (cuBLAS GEMMs + eager
element-wise kernels) for scale."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import solvingpapers_b200 as sp
from solvingpapers_b200 import _lib as L


def timeit(fn, n=10, warm=3):
    for _ in range(warm):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


if __name__ == "__main__":
    torch.manual_seed(0)
    N, C, H = 8192, 4096, 16384
    if len(sys.argv) > 1:
        N, C, H = (int(a) for a in sys.argv[1:4])
    bf = torch.bfloat16
    x = torch.randn(N, C, device="cuda").to(bf)
    dy = torch.randn(N, C, device="cuda").to(bf)
    w13 = (torch.randn(C, 2 * H, device="cuda") / C ** 0.5).to(bf)
    w2 = (torch.randn(H, C, device="cuda") / H ** 0.5).to(bf)
    nw = torch.randn(C, device="cuda").to(bf)
    for residual in (True, False, True, False):
        cfg = sp.FfnConfig(N=N, C=C, H=H, dtype=bf, residual=residual)
        ff = sp.FeedForward(cfg)
        p = {"w13": w13, "w2": w2, "norm_w": nw}
        g = ff.alloc_grads(p)
        g["dx"] = torch.empty_like(x)

        def step():
            ff.forward(x, p)
            ff.backward(dy, grads=g)

        ms = timeit(step)
        print(f"spb ffn  N{N} C{C} H{H} residual={int(residual)}: {ms:7.3f} ms  {N/ms/1e3:7.2f} M tokens/s  {cfg.flops_fwd_bwd()/ms/1e9:7.1f} TFLOP/s  "
              f"ws {ff.layout.total/2**20:.0f} MiB")

    # the same op sequence with torch ops (cuBLAS + eager element-wise), no autograd graph overhead: explicit backward
    w1, w3 = w13[:, :H], w13[:, H:]

    def torch_step():
        r = torch.rsqrt(x.float().pow(2).mean(-1, keepdim=True) + 1e-6)
        xh = (x.float() * nw.float() * r).to(bf)
        g13 = xh @ w13
        g1, g3 = g13[:, :H], g13[:, H:]
        a = torch.nn.functional.silu(g3) * g1
        y = x + a @ w2
        da = dy @ w2.t()
        dw2 = (a.t() @ dy).float()
        s = torch.sigmoid(g3.float())
        dg1 = (da.float() * g3.float() * s).to(bf)
        dg3 = (da.float() * g1.float() * s * (1 + g3.float() * (1 - s))).to(bf)
        dg13 = torch.cat([dg1, dg3], dim=1)
        dw13 = (xh.t() @ dg13).float()
        dxh = dg13 @ w13.t()
        return y, dw2, dw13, dxh

    ms = timeit(torch_step, n=5, warm=2)
    print(f"torch ops (cuBLAS + eager element-wise, norm VJP not included): {ms:7.3f} ms  {N/ms/1e3:7.2f} M tokens/s")
