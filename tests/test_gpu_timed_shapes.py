"""GPU: value checks at the sizes bench.py actually times - the small parity cases never wrap the
K/V/Q ring phases as often as T=2048/4096 does, never reach K=4096/8192 on the CTA-pair GEMM and never run its K-split
tail wave (TMA reduce-add into pre-zeroed tiles), which computes dWo and dWqkv in every C3 step.

Checkers: the oracle's attention math in fp64 on the host (flash), an fp64 product (GEMM; the full matrix against
torch's fp64 matmul on the GPU plus sampled rows against a host fp64 product), and the oracle's block in fp32 on the host
(whole C3 block at B=1 - what bench.py's CPU legs run per step)."""
import math

import numpy as np
import pytest
import torch

from oracle import llama_oracle, max_rel_err
from oracle.metrics import TOL_BF16

pytestmark = pytest.mark.gpu
DEV = "cuda"


def rnd(shape, seed, dtype=torch.bfloat16, scale=1.0):
    g = np.random.default_rng(seed)
    return torch.from_numpy(g.standard_normal(shape).astype(np.float32) * scale).to(dtype)


# ------------------------------------------------------------------------------------------------ flash at C3 / C4 geometry
def _attn_ref64(q, k, v, do, causal, scale):
    """llama3:820-828 in fp64 with autograd, one q head at a time (bounded host memory: T^2 doubles per head)."""
    B, T, Hq, hd = q.shape
    Hkv = k.shape[2]
    n_rep = Hq // Hkv
    q64, k64, v64 = (t.double() for t in (q, k, v))
    o = torch.empty(B, T, Hq, hd, dtype=torch.float64)
    lse = torch.empty(B, Hq, T, dtype=torch.float64)
    dq, dk, dv = torch.zeros_like(q64), torch.zeros_like(k64), torch.zeros_like(v64)
    mask = llama_oracle.make_mask(T, torch.float64)[0, 0] if causal else None
    for b in range(B):
        for h in range(Hq):
            g = h // n_rep  # repeat_kv's consecutive grouping (llama3:626-627)
            qh = q64[b, :, h].clone().requires_grad_(True)
            kh = k64[b, :, g].clone().requires_grad_(True)
            vh = v64[b, :, g].clone().requires_grad_(True)
            s = qh @ kh.t() * scale
            if mask is not None:
                s = s + mask
            p = torch.softmax(s, dim=-1)
            oh = p @ vh
            oh.backward(do[b, :, h].double())
            o[b, :, h] = oh.detach()
            lse[b, h] = torch.logsumexp(s.detach(), dim=-1)
            dq[b, :, h] = qh.grad
            dk[b, :, g] += kh.grad
            dv[b, :, g] += vh.grad
    return o, lse, dq, dk, dv


@pytest.mark.parametrize("B,T,Hq,Hkv,hd", [
    (1, 2048, 8, 2, 128),   # C3 head geometry and length: 16 KV tiles, 4 q heads per kv head
    (1, 4096, 4, 1, 256),   # C4 (Gemma-2B-shaped MQA): 32 KV tiles, hd-split backward
    (1, 2048, 2, 2, 64),    # hd 64 at the same length (4-stage Q/dO ring)
])
def test_flash_at_timed_lengths(B, T, Hq, Hkv, hd):
    from solvingpapers_b200 import ops, _lib as L
    q, k, v, do = (rnd((B, T, h, hd), 40 + i) for i, h in enumerate((Hq, Hkv, Hkv, Hq)))
    scale = 1.0 / math.sqrt(hd)
    oref, lseref, dqref, dkref, dvref = _attn_ref64(q.float(), k.float(), v.float(), do.float(), True, scale)
    qd, kd, vd, dod = (t.to(DEV) for t in (q, k, v, do))
    o, lse = ops.flash_fwd(qd, kd, vd, True, scale, impl=L.IMPL_TCGEN05)
    dq, dk, dv = ops.flash_bwd(qd, kd, vd, o, dod, lse, True, scale, impl=L.IMPL_TCGEN05)
    torch.cuda.synchronize()
    errs = {"o": max_rel_err(o, oref), "lse": max_rel_err(lse, lseref), "dq": max_rel_err(dq, dqref), "dk": max_rel_err(dk, dkref),
            "dv": max_rel_err(dv, dvref)}
    for name, e in errs.items():
        assert e < TOL_BF16, (name, errs)
    # per-tile check: a wrong tile late in the sequence must not hide behind a large global maximum
    for t0 in range(0, T, 512):
        sl = slice(t0, t0 + 512)
        assert max_rel_err(o[:, sl], oref[:, sl]) < TOL_BF16 and max_rel_err(dq[:, sl], dqref[:, sl]) < TOL_BF16, t0
        assert max_rel_err(dk[:, sl], dkref[:, sl]) < TOL_BF16 and max_rel_err(dv[:, sl], dvref[:, sl]) < TOL_BF16, t0


# ------------------------------------------------------------------------------------------------ CTA-pair GEMM at long K
def _gemm_long_k(M, N, K, tA, tB, out_dtype, bias=False, accumulate=False, seed=0):
    from solvingpapers_b200 import ops, _lib as L
    A = rnd((K, M) if tA else (M, K), seed).to(DEV)
    Bm = rnd((N, K) if tB else (K, N), seed + 1, scale=1.0 / math.sqrt(K)).to(DEV)
    bs = rnd((N,), seed + 2).to(DEV) if bias else None
    opA = A.double().t() if tA else A.double()
    opB = Bm.double().t() if tB else Bm.double()
    ref = opA @ opB  # torch fp64 on the GPU = the full-matrix checker
    if bias:
        ref = ref + bs.double()
    out = None
    if accumulate:
        out = rnd((M, N), seed + 3, out_dtype).to(DEV)
        ref = ref + out.double()
    d = ops.gemm(A, Bm, transA=tA, transB=tB, bias=bs, out=out, out_dtype=out_dtype, accumulate=accumulate, impl=L.IMPL_TCGEN05)
    torch.cuda.synchronize()
    # independent of torch's GEMM: 96 sampled rows (every 256-row tile of M is hit) against a host fp64 product
    rows = torch.from_numpy(np.random.default_rng(seed + 9).choice(M, size=min(96, M), replace=False)).sort().values
    rows = torch.unique(torch.cat([rows, torch.arange(0, M, 256), torch.arange(255, M, 256)]))
    host = (opA[rows].cpu() @ opB.cpu())
    if bias:
        host = host + bs.double().cpu()
    if not accumulate:
        assert max_rel_err(d[rows], host) < (3e-5 if out_dtype == torch.float32 else 4e-3)
    return max_rel_err(d, ref), float((d.double() - ref).abs().max() / ref.abs().max())


@pytest.mark.parametrize("tA,tB", [(False, True), (False, False), (True, False), (True, True)])
@pytest.mark.parametrize("K", [4096, 8192])
def test_gemm_pair_kernel_long_k_all_majors(tA, tB, K):
    """2048 x 4096 output = 128 tiles of 256 x 256 >= 74: gemm_tc2_kernel; K = 4096 and 8192 wrap the 6-stage ring 10-21
    times per tile.  bf16 output: the error is the output rounding (2^-9 of an element, elements <= max) - a dropped
    64-wide k-block would be ~sqrt(64/K) = 9-12 % of an element."""
    err, _ = _gemm_long_k(2048, 4096, K, tA, tB, torch.bfloat16, seed=K // 64 + 2 * tA + tB)
    assert err < 4e-3, err


@pytest.mark.parametrize("M,N,K,tA,tB", [
    (4096, 4096, 8192, True, False),   # the real dWo of a C3 step: 256 tiles = 3 waves of 74 + 34 tail tiles split 2-way along K
    (4096, 6144, 8192, True, False),   # the real dWqkv: 384 tiles = 5 waves + 14 tail tiles split 5-way
    (6144, 4096, 8192, True, False),   # torch-layout ([out,in]) weights: dW^T orientation
    (4096, 4096, 4096, False, True),   # fp32 output on K-major operands
    (2304, 4096, 2048, True, False),   # ragged tile rows (2304 = 9 x 256) + tail
])
def test_gemm_pair_kernel_fp32_output_k_split_tail(M, N, K, tA, tB):
    """fp32 weight-gradient GEMMs whose tile count leaves a tail wave: full waves store, tail tiles are zeroed by
    gemm_tc_zero_tail_kernel and filled by cp.reduce.async.bulk.tensor adds of the K parts.  fp32 accumulation of K
    products of exact bf16 inputs: tolerance 2e-5 * sqrt(K / 4096) of max|ref| (measured error is ~1e-6)."""
    err, _ = _gemm_long_k(M, N, K, tA, tB, torch.float32, seed=M // 256 + N // 256)
    assert err < 2e-5 * math.sqrt(K / 4096), err


def test_gemm_pair_kernel_epilogues_at_size():
    """bias / accumulate epilogues on the pair kernel (the small epilogue tests stay below its 74-tile threshold)."""
    e_bias, _ = _gemm_long_k(2048, 4096, 1024, False, False, torch.bfloat16, bias=True, seed=5)
    e_acc, _ = _gemm_long_k(2048, 4096, 1024, False, True, torch.bfloat16, accumulate=True, seed=6)
    e_acc32, _ = _gemm_long_k(2048, 4096, 1024, True, False, torch.float32, accumulate=True, seed=7)
    e_b32, _ = _gemm_long_k(4096, 2048, 512, False, False, torch.float32, bias=True, seed=8)
    assert e_bias < 4e-3 and e_acc < 4e-3 and e_acc32 < 2e-5 and e_b32 < 2e-5, (e_bias, e_acc, e_acc32, e_b32)


# ------------------------------------------------------------------------------------------------ the whole C3 block at B = 1
@pytest.mark.parametrize("residual", [False, True])
def test_c3_block_values_against_oracle(residual):
    """d4096, 32 q / 8 kv heads x 128, T2048, bf16, B=1: every output and gradient of spb_block_fwd/bwd against the
    oracle's fp32 host run on the bf16-rounded inputs (llama3:809-829,881,885,995).  B=1 keeps the host run to seconds;
    the per-batch kernels (flash grid y, GEMM row tiles) see the same code paths as B_local=4."""
    import solvingpapers_b200 as sp
    from solvingpapers_b200 import _lib as L
    B, T, C, H, Hkv = 1, 2048, 4096, 32, 8
    hd = C // H
    rng = np.random.default_rng(2)
    rd = lambda t: t.to(torch.bfloat16).float()
    bp = llama_oracle.init_block_params(rng, C, H, Hkv)
    a = {k: rd(v) for k, v in bp["attention"].items()}
    pr = {"attention": a, "attention_norm": rd(bp["attention_norm"])}
    x = rd(torch.from_numpy(rng.standard_normal((B, T, C)).astype(np.float32)))
    dy = rd(torch.from_numpy(rng.standard_normal((B, T, C)).astype(np.float32)))
    torch.set_num_threads(max(1, torch.get_num_threads()))
    ref = llama_oracle.block_fwd_bwd(pr, x, dy, H, Hkv, causal=True, residual=residual)
    blk = sp.AttentionBlock(sp.BlockConfig(B, T, C, H, Hkv, hd, dtype=torch.bfloat16, norm_kind=L.NORM_RMS, rope=True, causal=True,
                                           residual=residual))
    cos, sin = llama_oracle.precompute_freqs_cis(hd, T)
    dev = lambda t: t.to(torch.bfloat16).to(DEV)
    p = {"wqkv": dev(torch.cat([a["wq"], a["wk"], a["wv"]], dim=1)), "wo": dev(a["wo"]), "norm_w": dev(pr["attention_norm"]),
         "rope": sp.llama_rope_table(cos, sin).to(DEV)}
    y = blk.forward(dev(x), p)
    k, v = blk.kv_cache()
    g = blk.backward(dev(dy))
    torch.cuda.synchronize()
    nq, nkv = H * hd, Hkv * hd
    got = {"y": y, "k": k, "v": v, "dx": g["dx"], "dwq": g["dwqkv"][:, :nq], "dwk": g["dwqkv"][:, nq:nq + nkv],
           "dwv": g["dwqkv"][:, nq + nkv:], "dwo": g["dwo"], "dnorm": g["dnorm_w"]}
    errs = {n: max_rel_err(t, ref[n]) for n, t in got.items()}
    assert all(e < TOL_BF16 for e in errs.values()), errs
