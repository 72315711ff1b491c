"""GPU parity tests, one kernel family at a time, through the C ABI (solvingpapers_b200.ops -> libspb200.so).
Checker: the torch-CPU oracle / a plain torch fp32-fp64 restatement of the same op on the same seeded inputs.
Metric: max|a-b| / max|b| (oracle.metrics); tolerance 1e-4 (fp32 mode) and 2e-2 (bf16 mode) from BASELINE.json."""
import math

import numpy as np
import pytest
import torch

from oracle import gemma_oracle, gpt_oracle, llama_oracle, max_rel_err
from oracle.metrics import TOL_BF16, TOL_F32

pytestmark = pytest.mark.gpu

DEV = "cuda"


def _ops():
    from solvingpapers_b200 import ops
    return ops


def _L():
    from solvingpapers_b200 import _lib
    return _lib


def rnd(shape, seed, dtype=torch.float32, scale=1.0):
    g = np.random.default_rng(seed)
    t = torch.from_numpy(g.standard_normal(shape).astype(np.float32) * scale)
    return t.to(dtype)


def tol(dtype):
    return TOL_BF16 if dtype == torch.bfloat16 else TOL_F32


# ------------------------------------------------------------------------------------------------ norm
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("N,C", [(64, 256), (37, 64), (16, 4096), (8, 768), (5, 100), (3, 10), (4, 8192)])
@pytest.mark.parametrize("kind", ["rms", "rms_upcast", "layer"])
def test_norm_fwd_bwd(dtype, N, C, kind):
    ops, L = _ops(), _L()
    x = rnd((N, C), 1, dtype)
    w = rnd((C,), 2, dtype)  # N(0,1) weights like llama3:748
    b = rnd((C,), 3, dtype, 0.5)
    dy = rnd((N, C), 4, dtype)
    xr, wr, br, dyr = (t.float().double().requires_grad_(True) for t in (x, w, b, dy))
    if kind == "layer":
        ref = gpt_oracle.layer_norm(xr, wr, br, eps=1e-6)
        y, stats = ops.norm_fwd(x.to(DEV), w.to(DEV), b.to(DEV), kind=L.NORM_LAYER, eps=1e-6)
    elif kind == "rms":
        ref = llama_oracle.rms_norm(xr, wr, 1e-6)
        y, stats = ops.norm_fwd(x.to(DEV), w.to(DEV), None, kind=L.NORM_RMS, eps=1e-6)
    else:
        ref = gemma_oracle.rms_norm(xr, wr, 1e-6)
        y, stats = ops.norm_fwd(x.to(DEV), w.to(DEV), None, kind=L.NORM_RMS, eps=1e-6, upcast=True)
    assert max_rel_err(y, ref) < tol(dtype)
    ref.backward(dyr)
    dx, dw, db = ops.norm_bwd(x.to(DEV), w.to(DEV), stats, dy.to(DEV), kind=L.NORM_LAYER if kind == "layer" else L.NORM_RMS,
                              want_db=(kind == "layer"))
    assert max_rel_err(dx, xr.grad) < tol(dtype)
    assert max_rel_err(dw, wr.grad) < tol(dtype)
    if kind == "layer":
        assert max_rel_err(db, br.grad) < tol(dtype)


def test_norm_empty_and_bad_args():
    ops, L = _ops(), _L()
    x = torch.empty((0, 64), device=DEV)
    y, _ = ops.norm_fwd(x, torch.ones(64, device=DEV))
    assert y.shape == (0, 64)
    with pytest.raises(L.SpbError):
        ops.norm_fwd(torch.ones((2, 64), device=DEV), torch.ones(64, device=DEV), kind=7)


# ------------------------------------------------------------------------------------------------ rope
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("B,T,H,hd", [(2, 16, 6, 64), (1, 9, 3, 16), (2, 128, 4, 128), (1, 5, 2, 6)])
def test_rope_llama_and_transpose(dtype, B, T, H, hd):
    ops = _ops()
    from solvingpapers_b200 import llama_rope_table
    x = rnd((B, T, H, hd), 5, dtype)
    cos, sin = llama_oracle.precompute_freqs_cis(hd, T)
    ref, _ = llama_oracle.apply_rotary_emb(x.float().double(), x.float().double()[:, :, :1], (cos.double(), sin.double()))
    tab = llama_rope_table(cos, sin).to(DEV)
    # strided: the heads live in the first H*hd columns of a wider row (packed q|k|v buffer)
    buf = torch.zeros((B * T, H * hd + 2 * hd), dtype=dtype, device=DEV)
    buf[:, : H * hd] = x.reshape(B * T, H * hd).to(DEV)
    ops.rope_apply_(buf, T, H, hd, tab)
    assert max_rel_err(buf[:, : H * hd].reshape(B, T, H, hd), ref) < tol(dtype)
    assert float(buf[:, H * hd:].abs().max()) == 0.0
    # the transpose is the inverse for a rotation: R^T R x = x
    ops.rope_apply_(buf, T, H, hd, tab, transpose=True)
    assert max_rel_err(buf[:, : H * hd].reshape(B, T, H, hd), x.float()) < (3 * tol(dtype))


def test_rope_gemma_table_matches_notebook_matrix():
    ops = _ops()
    from solvingpapers_b200 import gemma_rope_table
    B, T, D = 2, 16, 64
    x = rnd((B, T, D), 6)
    c, s = gemma_oracle.rotary_coeffs(T, D)
    ref = gemma_oracle.apply_rotary(x.double(), c.double(), s.double())
    buf = x.reshape(B * T, D).to(DEV).clone()
    ops.rope_apply_(buf, T, 1, D, gemma_rope_table(c, s, D // 2).to(DEV))
    assert max_rel_err(buf.reshape(B, T, D), ref) < TOL_F32


# ------------------------------------------------------------------------------------------------ colsum
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_colsum(dtype):
    ops = _ops()
    x = rnd((1000, 300), 7, dtype)
    out = ops.colsum(x.to(DEV)[:, 20:220])
    assert max_rel_err(out, x.double()[:, 20:220].sum(0)) < tol(dtype)


# ------------------------------------------------------------------------------------------------ GEMM
def _gemm_case(dtype, M, N, K, tA, tB, impl, bias=False, out_dtype=None, accumulate=False, seed=0):
    ops = _ops()
    A = rnd((K, M) if tA else (M, K), seed, dtype)
    B = rnd((N, K) if tB else (K, N), seed + 1, dtype, 1.0 / math.sqrt(K))
    bs = rnd((N,), seed + 2, dtype) if bias else None
    out_dtype = out_dtype or dtype
    ref = (A.double().t() if tA else A.double()) @ (B.double().t() if tB else B.double())
    if bias:
        ref = ref + bs.double()
    out = None
    if accumulate:
        init = rnd((M, N), seed + 3, out_dtype)
        ref = ref + init.double()
        out = init.to(DEV)
    d = ops.gemm(A.to(DEV), B.to(DEV), transA=tA, transB=tB, bias=bs.to(DEV) if bias else None, out=out, out_dtype=out_dtype,
                 accumulate=accumulate, impl=impl)
    torch.cuda.synchronize()
    return max_rel_err(d, ref)


@pytest.mark.parametrize("tA", [False, True])
@pytest.mark.parametrize("tB", [False, True])
@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (64, 64, 16), (130, 70, 33), (256, 512, 256), (2048, 256, 256)])
def test_gemm_simt_f32(tA, tB, M, N, K):
    L = _L()
    assert _gemm_case(torch.float32, M, N, K, tA, tB, L.IMPL_SIMT, bias=True) < TOL_F32


def test_gemm_simt_f32_accumulate_and_long_k():
    L = _L()
    assert _gemm_case(torch.float32, 96, 80, 8192, True, False, L.IMPL_SIMT, accumulate=True) < TOL_F32


@pytest.mark.parametrize("tA,tB", [(False, True), (False, False), (True, False), (True, True)])
@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (128, 128, 128), (256, 512, 512), (384, 768, 256), (200, 136, 72),
                                   (1024, 6144, 512),
                                   (256, 96, 128),     # N % 64 == 32: the last TMA store box of a tile is half stale, clipped at N
                                   (136, 288, 64),     # 256 + 32 columns: the second tile ends on its first chunk; rows past M clipped
                                   (8200, 160, 64)])   # many row tiles per CTA: the staging halves are reused across tiles
def test_gemm_tcgen05_bf16(tA, tB, M, N, K):
    """All four operand-major combinations: (K-major, K-major) = dX; (K-major, MN-major) = forward with [in,out]
    weights; (MN, MN) = dW.  Ragged sizes exercise TMA zero fill and the predicated epilogue."""
    L = _L()
    if (tA and M % 8) or ((not tB) and N % 8) or ((not tA) and K % 8) or (tB and K % 8):
        pytest.skip("TMA needs leading dimensions that are multiples of 8 elements")
    assert _gemm_case(torch.bfloat16, M, N, K, tA, tB, L.IMPL_TCGEN05) < TOL_BF16 / 4


def test_gemm_tcgen05_epilogues():
    L = _L()
    assert _gemm_case(torch.bfloat16, 256, 512, 256, False, False, L.IMPL_TCGEN05, bias=True) < TOL_BF16 / 4
    assert _gemm_case(torch.bfloat16, 256, 512, 256, False, True, L.IMPL_TCGEN05, accumulate=True) < TOL_BF16 / 4
    assert _gemm_case(torch.bfloat16, 256, 512, 256, True, False, L.IMPL_TCGEN05, out_dtype=torch.float32) < 1e-3
    assert _gemm_case(torch.bfloat16, 200, 200, 200, True, False, L.IMPL_TCGEN05, out_dtype=torch.float32, accumulate=True) < 1e-3
    # epilogue through shared memory + TMA stores (plain, non-accumulating tiles): ragged edges in both output types, with bias
    assert _gemm_case(torch.bfloat16, 200, 232, 72, False, False, L.IMPL_TCGEN05, bias=True) < TOL_BF16 / 4
    assert _gemm_case(torch.bfloat16, 328, 72, 136, True, False, L.IMPL_TCGEN05, out_dtype=torch.float32) < 1e-3
    assert _gemm_case(torch.bfloat16, 4096, 2304, 768, False, True, L.IMPL_TCGEN05, bias=True) < TOL_BF16 / 4   # ViT QKV shape: CTA-pair kernel


@pytest.mark.parametrize("out_dtype", [torch.bfloat16, torch.float32])
def test_gemm_tcgen05_output_view_keeps_its_neighbours(out_dtype):
    """D is a column range of a wider buffer (how the block writes q | k | v and the packed dW): the TMA-store epilogue's 64- /
    32-column boxes overhang the range at its right edge and must be clipped there, not at the row pitch."""
    ops, L = _ops(), _L()
    M, N, K, LD = 264, 96, 128, 160
    A, B = rnd((M, K), 3, torch.bfloat16), rnd((K, N), 4, torch.bfloat16, 1.0 / math.sqrt(K))
    big = torch.full((M + 8, LD), 7.0, dtype=out_dtype, device=DEV)
    out = big[:M, 32:32 + N]
    ops.gemm(A.to(DEV), B.to(DEV), out=out, impl=L.IMPL_TCGEN05)
    torch.cuda.synchronize()
    ref = A.double() @ B.double()
    assert max_rel_err(out, ref) < (TOL_BF16 / 4 if out_dtype == torch.bfloat16 else 1e-3)
    assert torch.all(big[:M, :32] == 7.0) and torch.all(big[:M, 32 + N:] == 7.0) and torch.all(big[M:] == 7.0)


def test_gemm_tcgen05_split_k_weight_gradient_shape():
    """d=256 models: dW is 256x768 with K = tokens; few tiles -> split-K with fp32 atomics (fp32 output)."""
    L = _L()
    assert _gemm_case(torch.bfloat16, 256, 768, 16384, True, False, L.IMPL_TCGEN05, out_dtype=torch.float32) < 1e-3


def test_gemm_tcgen05_many_tiles_persistent():
    L = _L()
    assert _gemm_case(torch.bfloat16, 4096, 2048, 1024, False, False, L.IMPL_TCGEN05) < TOL_BF16 / 4


def test_gemm_rejects_bad_arguments():
    ops, L = _ops(), _L()
    a = torch.ones((8, 8), device=DEV)
    with pytest.raises(L.SpbError):
        ops.gemm(a, torch.ones((9, 8), device=DEV))
    with pytest.raises(L.SpbError):  # tcgen05 path takes bf16 only - no silent conversion
        ops.gemm(a, a, impl=L.IMPL_TCGEN05)
    with pytest.raises(L.SpbError):  # TMA alignment: ld must be a multiple of 8
        ops.gemm(torch.ones((8, 12), device=DEV, dtype=torch.bfloat16), torch.ones((12, 8), device=DEV, dtype=torch.bfloat16),
                 impl=L.IMPL_TCGEN05)


# ------------------------------------------------------------------------------------------------ attention
def _attn_ref(q, k, v, causal, scale):
    """Oracle math (llama3:820-828) on [B,T,H,hd] tensors with autograd, in fp64."""
    B, T, Hq, hd = q.shape
    n_rep = Hq // k.shape[2]
    kk = llama_oracle.repeat_kv(k, n_rep).permute(0, 2, 1, 3)
    vv = llama_oracle.repeat_kv(v, n_rep).permute(0, 2, 1, 3)
    s = torch.matmul(q.permute(0, 2, 1, 3), kk.transpose(-1, -2)) * scale
    if causal:
        s = s + llama_oracle.make_mask(T, s.dtype)
    p = torch.softmax(s, dim=-1)
    lse = torch.logsumexp(s, dim=-1)
    return torch.matmul(p, vv).permute(0, 2, 1, 3), lse


def _attn_case(dtype, B, T, Hq, Hkv, hd, causal, impl, seed=11):
    ops = _ops()
    q, k, v, do = (rnd((B, T, h, hd), seed + i, dtype) for i, h in enumerate((Hq, Hkv, Hkv, Hq)))
    scale = 1.0 / math.sqrt(hd)
    qr, kr, vr = (t.float().double().requires_grad_(True) for t in (q, k, v))
    oref, lseref = _attn_ref(qr, kr, vr, causal, scale)
    oref.backward(do.float().double())
    qd, kd, vd, dod = (t.to(DEV) for t in (q, k, v, do))
    o, lse = ops.flash_fwd(qd, kd, vd, causal, scale, impl=impl)
    dq, dk, dv = ops.flash_bwd(qd, kd, vd, o, dod, lse, causal, scale, impl=impl)
    torch.cuda.synchronize()
    return {"o": max_rel_err(o, oref), "lse": max_rel_err(lse, lseref), "dq": max_rel_err(dq, qr.grad),
            "dk": max_rel_err(dk, kr.grad), "dv": max_rel_err(dv, vr.grad)}


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("B,T,Hq,Hkv,hd,causal", [
    (2, 128, 4, 2, 64, True),    # C1 shape (llama notebook)
    (1, 197, 3, 3, 64, False),   # ViT-B/16 ragged, bidirectional
    (2, 17, 4, 4, 16, False),    # ViT notebook literal
    (1, 256, 1, 1, 256, True),   # GPT literal: one head of 256
    (1, 64, 2, 1, 768, True),    # Gemma literal: 2 q heads share one 768-wide kv
    (1, 1, 2, 2, 32, True),      # single token
])
def test_attention_simt(dtype, B, T, Hq, Hkv, hd, causal):
    L = _L()
    errs = _attn_case(dtype, B, T, Hq, Hkv, hd, causal, L.IMPL_SIMT)
    for k, e in errs.items():
        assert e < tol(dtype), (k, errs)


@pytest.mark.parametrize("B,T,Hq,Hkv,hd,causal", [
    (2, 128, 4, 2, 64, True),     # C1 shape in bf16
    (1, 197, 3, 3, 64, False),    # ViT-B/16: ragged T, bidirectional
    (2, 256, 4, 1, 128, True),    # MQA, two KV tiles, diagonal masking
    (1, 384, 8, 2, 128, True),    # Llama-8B head geometry (4 q heads per kv head)
    (1, 200, 2, 2, 128, False),   # ragged, hd 128
    (1, 1, 2, 2, 64, True),       # single token
    (1, 130, 2, 1, 64, True),     # just past one tile
    (2, 256, 1, 1, 256, True),    # GPT notebook geometry: one head of 256 (single-tile forward, hd-split backward)
    (1, 200, 4, 1, 256, True),    # Gemma-2B-shaped MQA, ragged
    (1, 128, 2, 2, 256, False),   # hd 256 bidirectional
    (40, 1024, 1, 1, 256, True),  # hd 256 with >= 4 x SMs role-CTAs: one CTA per role (the small cases above run the cluster-of-two form)
    (1, 64, 1, 1, 256, True),     # hd 256, one 64-row query tile: the second CTA of a role cluster has no iteration at all
])
def test_attention_tcgen05_flash(B, T, Hq, Hkv, hd, causal):
    """tcgen05 flash forward + backward (bf16) against the oracle math; bf16-mode tolerance 2e-2."""
    L = _L()
    errs = _attn_case(torch.bfloat16, B, T, Hq, Hkv, hd, causal, L.IMPL_TCGEN05)
    for k, e in errs.items():
        assert e < TOL_BF16, (k, errs)


def test_attention_tcgen05_is_deterministic_forward_and_rejects_unsupported():
    ops, L = _ops(), _L()
    q = rnd((1, 256, 2, 128), 3, torch.bfloat16).to(DEV)
    o1, l1 = ops.flash_fwd(q, q, q, True, 0.1, impl=L.IMPL_TCGEN05)
    o2, l2 = ops.flash_fwd(q, q, q, True, 0.1, impl=L.IMPL_TCGEN05)
    assert torch.equal(o1, o2) and torch.equal(l1, l2)
    with pytest.raises(L.SpbError):  # fp32 has no tcgen05 attention: error, not a silent slow path
        ops.flash_fwd(q.float(), q.float(), q.float(), True, 0.1, impl=L.IMPL_TCGEN05)
    q768 = rnd((1, 64, 1, 768), 4, torch.bfloat16).to(DEV)
    with pytest.raises(L.SpbError):  # hd 768 (Gemma literal) has no tcgen05 tiling: explicit request is an error
        ops.flash_fwd(q768, q768, q768, True, 0.1, impl=L.IMPL_TCGEN05)


# ------------------------------------------------------------------------------------------------ wide heads (Gemma literal, hd 768)
@pytest.mark.parametrize("B,T,Hq,Hkv,hd,causal", [
    (1, 64, 2, 1, 768, True),      # Gemma literal geometry: 2 q heads share one 768-wide k/v
    (2, 200, 2, 1, 768, True),     # ragged T
    (1, 2300, 2, 1, 768, True),    # more than one 2048-row block: causal key ranges differ per block
    (1, 130, 2, 2, 512, False),    # bidirectional, MHA
    (1, 96, 4, 2, 320, True),      # a width that is not a power of two
])
def test_attention_wide_heads_tcgen05_gemm_pipeline(B, T, Hq, Hkv, hd, causal):
    """spb_attn_wide_fwd / bwd (bf16, hd > 256: every contraction on the tcgen05 GEMM) against the oracle math in fp64."""
    ops = _ops()
    q, k, v, do = (rnd((B, T, h, hd), 31 + i, torch.bfloat16) for i, h in enumerate((Hq, Hkv, Hkv, Hq)))
    scale = 1.0 / math.sqrt(hd)
    qr, kr, vr = (t.float().double().requires_grad_(True) for t in (q, k, v))
    oref, lseref = _attn_ref(qr, kr, vr, causal, scale)
    oref.backward(do.float().double())
    qd, kd, vd, dod = (t.to(DEV) for t in (q, k, v, do))
    o, lse = ops.attn_wide_fwd(qd, kd, vd, causal, scale)
    dq, dk, dv = ops.attn_wide_bwd(qd, kd, vd, o, dod, lse, causal, scale)
    torch.cuda.synchronize()
    errs = {"o": max_rel_err(o, oref), "lse": max_rel_err(lse, lseref), "dq": max_rel_err(dq, qr.grad), "dk": max_rel_err(dk, kr.grad),
            "dv": max_rel_err(dv, vr.grad)}
    for name, e in errs.items():
        assert e < TOL_BF16, (name, errs)
