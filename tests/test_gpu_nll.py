"""GPU parity of the loss head's memory-bound half: spb_nll_fwd / spb_nll_bwd against the oracle's
restatement of llama3:959-967 (log_softmax + take_along_axis + mean), incl. the notebook's vocabulary 50257 (not a multiple of
the 128-bit vector width) and a padded row pitch."""
import os

import numpy as np
import pytest
import torch

from oracle import llama_oracle, max_rel_err
from oracle.metrics import TOL_BF16, TOL_F32

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("N,V,pad", [(64, 50257, 0), (64, 50257, 7), (37, 1000, 0), (5, 33, 0), (300, 4096, 0)])
def test_nll_loss_and_gradient(dtype, N, V, pad):
    from solvingpapers_b200 import llama as spl
    rng = np.random.default_rng(N + V)
    lg = torch.from_numpy((rng.standard_normal((N, V)) * 3).astype(np.float32)).to(dtype)
    tg = torch.from_numpy(rng.integers(0, V, size=(N,)))
    ref_in = lg.float().requires_grad_(True)
    ref = llama_oracle.nll_loss(ref_in, tg)
    ref.backward()
    buf = torch.zeros(N, V + pad, dtype=dtype, device="cuda")
    buf[:, :V] = lg.cuda()
    x = buf[:, :V].detach().requires_grad_(True)  # row pitch V + pad
    loss = spl.nll_loss(x, tg.cuda())
    loss.backward()
    torch.cuda.synchronize()
    tol = TOL_BF16 if dtype == torch.bfloat16 else TOL_F32
    assert abs(float(loss) - float(ref)) <= tol * abs(float(ref)), (float(loss), float(ref))
    assert max_rel_err(x.grad, ref_in.grad) < tol


def test_loss_head_end_to_end_fp32():
    """rms_norm -> logits GEMM -> NLL through the library's ops (llama3:929-930, 959-967), forward + gradients of h and the weights."""
    from solvingpapers_b200 import llama as spl, ops
    rng = np.random.default_rng(3)
    B, T, C, V = 2, 32, 64, 1000
    h = torch.from_numpy(rng.standard_normal((B, T, C)).astype(np.float32))
    nw = torch.from_numpy(rng.standard_normal((C,)).astype(np.float32))
    w = torch.from_numpy((rng.standard_normal((C, V)) / 8).astype(np.float32))
    tg = torch.from_numpy(rng.integers(0, V, size=(B, T)))
    hr, nr, wr = (t.clone().requires_grad_(True) for t in (h, nw, w))
    ref = llama_oracle.loss_head(nr, wr, hr, tg)
    ref.backward()
    hd, nd, wd = (t.cuda().requires_grad_(True) for t in (h, nw, w))
    logits = ops.matmul(spl.rms_norm(hd, nd).reshape(-1, C), wd)
    loss = spl.nll_loss(logits, tg.cuda())
    loss.backward()
    torch.cuda.synchronize()
    assert abs(float(loss) - float(ref)) <= TOL_F32 * abs(float(ref))
    for got, want in ((hd.grad, hr.grad), (nd.grad, nr.grad), (wd.grad, wr.grad)):
        assert max_rel_err(got, want) < TOL_F32


def test_nll_against_golden_from_notebook_source(golden_dir):
    """the notebook's own compute_loss lines (llama3:959-967, executed for tests/golden/llama_loss_small.npz)"""
    from solvingpapers_b200 import llama as spl
    g = dict(np.load(os.path.join(golden_dir, "llama_loss_small.npz")))
    loss = spl.nll_loss(torch.from_numpy(g["logits"]).cuda(), torch.from_numpy(g["targets"]).cuda())
    assert abs(float(loss) - float(g["loss64"])) < TOL_F32 * abs(float(g["loss64"]))


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_nll_with_vocabulary_masked_logits(dtype):
    """-inf logits (masked vocabulary entries), including whole leading vector chunks and whole thread strides of -inf:
    loss and gradient stay finite and equal log_softmax's (the per-thread online max used to produce exp(-inf+inf) = NaN)."""
    from solvingpapers_b200 import llama as spl
    rng = np.random.default_rng(5)
    N, V = 9, 9000
    lg = torch.from_numpy((rng.standard_normal((N, V)) * 2).astype(np.float32)).to(dtype)
    lg[0, :4100] = -float("inf")       # more than one full pass of the CTA (256 threads x 8 x 2 = 4096 columns)
    lg[1, :16] = -float("inf")         # a thread's first two vectors
    lg[2, ::2] = -float("inf")
    lg[3, 100:] = -float("inf")        # only the head of the row is live
    tg = torch.from_numpy(rng.integers(0, 100, size=(N,)))
    tg[0] = 5000
    tg[2] = 1
    ref_in = lg.float().clone().requires_grad_(True)
    ref = llama_oracle.nll_loss(ref_in, tg)
    ref.backward()
    x = lg.detach().cuda().requires_grad_(True)
    loss = spl.nll_loss(x, tg.cuda())
    loss.backward()
    torch.cuda.synchronize()
    tol = TOL_BF16 if dtype == torch.bfloat16 else TOL_F32
    assert torch.isfinite(loss) and torch.isfinite(x.grad).all()
    assert abs(float(loss) - float(ref)) <= tol * abs(float(ref))
    assert max_rel_err(x.grad, ref_in.grad) < tol


@pytest.mark.parametrize("N,C,V", [(256, 256, 50257), (300, 64, 1000), (2048, 256, 4096), (64, 128, 257)])
def test_fused_loss_head_bf16_never_writes_logits(N, C, V):
    """spb_loss_head_fwd / bwd (logits GEMM + NLL fused) against the oracle's loss_head (llama3:929-930, 959-967) on
    bf16-rounded inputs: loss, dh (through the final rms_norm: dh and dnorm_f) and the [C, V] weight gradient; V = 50257 is the
    notebook's vocabulary (not a multiple of 8: the library pads a private copy)."""
    from solvingpapers_b200 import llama as spl
    rng = np.random.default_rng(V + N)
    rd = lambda t: t.to(torch.bfloat16).float()
    h = rd(torch.from_numpy(rng.standard_normal((N, C)).astype(np.float32)))
    nw = rd(torch.from_numpy(rng.standard_normal((C,)).astype(np.float32)))
    w = rd(torch.from_numpy((rng.standard_normal((C, V)) / np.sqrt(C)).astype(np.float32)))
    tg = torch.from_numpy(rng.integers(0, V, size=(N,)))
    tg[0], tg[-1] = V - 1, 0
    hr, nr, wr = (t.clone().requires_grad_(True) for t in (h, nw, w))
    ref = llama_oracle.loss_head(nr, wr, hr, tg)
    ref.backward()
    hd, nd, wd = (t.to(torch.bfloat16).cuda().requires_grad_(True) for t in (h, nw, w))
    loss = spl.loss_head(nd, wd, hd, tg.cuda())
    loss.backward()
    torch.cuda.synchronize()
    assert abs(float(loss) - float(ref)) <= TOL_BF16 * abs(float(ref)), (float(loss), float(ref))
    for name, got, want in (("dh", hd.grad, hr.grad), ("dnorm_f", nd.grad, nr.grad), ("dw", wd.grad, wr.grad)):
        assert max_rel_err(got, want) < TOL_BF16, name


def test_fused_loss_head_matches_unfused_kernels():
    """same inputs through ops.matmul + nll_loss (logits materialised) and through the fused head: the loss agrees to fp32 rounding of
    the bf16 logits, the gradients to bf16 rounding"""
    from solvingpapers_b200 import llama as spl, ops
    rng = np.random.default_rng(9)
    N, C, V = 512, 256, 8192
    h = torch.from_numpy(rng.standard_normal((N, C)).astype(np.float32)).bfloat16().cuda()
    w = torch.from_numpy((rng.standard_normal((C, V)) / 16).astype(np.float32)).bfloat16().cuda()
    tg = torch.from_numpy(rng.integers(0, V, size=(N,))).cuda()
    h1, w1 = h.clone().requires_grad_(True), w.clone().requires_grad_(True)
    l1 = spl.nll_loss(ops.matmul(h1, w1), tg)
    l1.backward()
    h2, w2 = h.clone().requires_grad_(True), w.clone().requires_grad_(True)
    l2 = spl._LossHeadFn.apply(h2, w2, tg)
    l2.backward()
    torch.cuda.synchronize()
    assert abs(float(l1) - float(l2)) < 5e-3 * abs(float(l1))
    assert max_rel_err(h2.grad, h1.grad) < TOL_BF16 and max_rel_err(w2.grad, w1.grad) < TOL_BF16
