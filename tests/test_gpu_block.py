"""GPU parity of the whole block (forward + backward) through spb_block_fwd / spb_block_bwd against
  (a) the golden fixtures made from the reference's own code (tests/golden/*.npz), and
  (b) the torch-CPU oracle on seeded synthetic inputs drawn from the notebooks' distributions.
fp32 mode: 1e-4; bf16 mode: inputs rounded to bf16, oracle in fp32 on the rounded values, 2e-2 (BASELINE.json)."""
import math
import os

import numpy as np
import pytest
import torch

from oracle import gemma_oracle, gpt_oracle, llama_oracle, max_rel_err, vit_oracle
from oracle.metrics import TOL_BF16, TOL_F32

pytestmark = pytest.mark.gpu
DEV = "cuda"
T_ = torch.from_numpy


def tol(dtype):
    return TOL_BF16 if dtype == torch.bfloat16 else TOL_F32


def _sp():
    import solvingpapers_b200 as sp
    from solvingpapers_b200 import _lib
    return sp, _lib


def _check(got, ref, dtype, names):
    errs = {n: max_rel_err(got[n], ref[n]) for n in names}
    bad = {n: e for n, e in errs.items() if not (e < tol(dtype))}
    assert not bad, f"over tolerance {tol(dtype)}: {bad}; all: {errs}"
    return errs


def run_llama(B, T, C, H, Hkv, dtype, causal=True, seed=0, attn_impl=0, gemm_impl=0, packed=False, params=None, x=None, dy=None,
              residual=False):
    sp, L = _sp()
    rng = np.random.default_rng(seed)
    hd = C // H
    if params is None:
        params = llama_oracle.init_block_params(rng, C, H, Hkv)
        x = T_(rng.standard_normal((B, T, C)).astype(np.float32))
        dy = T_(rng.standard_normal((B, T, C)).astype(np.float32))
    # bf16 mode: round once on the host, oracle sees the rounded values
    rd = lambda t: t.to(dtype).float()
    a = {k: rd(v) for k, v in params["attention"].items()}
    pr = {"attention": a, "attention_norm": rd(params["attention_norm"])}
    xr, dyr = rd(x), rd(dy)
    ref = llama_oracle.block_fwd_bwd(pr, xr, dyr, H, Hkv, causal=causal, dtype=torch.float32, residual=residual)
    cfg = sp.BlockConfig(B=B, T=T, C=C, Hq=H, Hkv=Hkv, hd=hd, dtype=dtype, norm_kind=L.NORM_RMS, norm_eps=1e-6, rope=True,
                         causal=causal, weight_layout=L.W_IN_OUT, attn_impl=attn_impl, gemm_impl=gemm_impl, residual=residual)
    blk = sp.AttentionBlock(cfg)
    cos, sin = llama_oracle.precompute_freqs_cis(hd, T)
    dev = lambda t: t.to(dtype).to(DEV)
    p = {"wo": dev(a["wo"]), "norm_w": dev(pr["attention_norm"]), "rope": sp.llama_rope_table(cos, sin).to(DEV)}
    if packed:
        p["wqkv"] = dev(torch.cat([a["wq"], a["wk"], a["wv"]], dim=1))
    else:
        p.update(wq=dev(a["wq"]), wk=dev(a["wk"]), wv=dev(a["wv"]))
    y = blk.forward(dev(xr), p)
    k, v = blk.kv_cache()
    g = blk.backward(dev(dyr))
    torch.cuda.synchronize()
    got = {"y": y, "k": k, "v": v, "dx": g["dx"], "dwo": g["dwo"], "dnorm": g["dnorm_w"]}
    if packed:
        nq, nkv = H * hd, Hkv * hd
        got.update(dwq=g["dwqkv"][:, :nq], dwk=g["dwqkv"][:, nq:nq + nkv], dwv=g["dwqkv"][:, nq + nkv:])
    else:
        got.update(dwq=g["dwq"], dwk=g["dwk"], dwv=g["dwv"])
    return got, ref


NAMES = ("y", "k", "v", "dx", "dwq", "dwk", "dwv", "dwo", "dnorm")


@pytest.mark.parametrize("tag", ["gqa_causal", "mha_nomask", "mqa_causal"])
def test_llama_block_fp32_against_golden_from_notebook_source(golden_dir, tag):
    g = dict(np.load(os.path.join(golden_dir, f"llama_{tag}.npz")))
    B, Tn, C, H, Hkv, causal = (int(v) for v in g["meta"])
    params = {"attention": {k: T_(g[k]) for k in ("wq", "wk", "wv", "wo")}, "attention_norm": T_(g["norm_w"])}
    got, ref = run_llama(B, Tn, C, H, Hkv, torch.float32, causal=bool(causal), params=params, x=T_(g["x"]), dy=T_(g["dy"]))
    # forward: against the literal notebook source run (fp32) and its fp64 run
    for n in ("y", "k", "v"):
        assert max_rel_err(got[n], T_(g[n])) < TOL_F32, n
        assert max_rel_err(got[n], T_(g[n + "64"])) < TOL_F32, n
    # gradients: directional derivatives from finite differences of the notebook source
    names = {"x": "dx", "norm_w": "dnorm", "wq": "dwq", "wk": "dwk", "wv": "dwv", "wo": "dwo"}
    for name, gname in names.items():
        grad = got[gname].detach().double().cpu().numpy()
        scale = float(np.abs(grad).max()) * math.sqrt(grad.size)
        for u, fd in zip(g["fd_dir_" + name], g["fd_val_" + name]):
            assert abs(float(np.sum(grad * u)) - fd) < 1e-4 * max(scale, 1e-6), (name, float(np.sum(grad * u)), fd)
    _check(got, ref, torch.float32, NAMES)


def test_c1_llama_notebook_default_dims_fp32():
    """BASELINE config 1: B16 T128 d256 4Q/2KV x 64, fp32, causal (llama3/LLaMA-jax.ipynb:349-358)."""
    got, ref = run_llama(16, 128, 256, 4, 2, torch.float32, seed=0)
    _check(got, ref, torch.float32, NAMES)


@pytest.mark.parametrize("B,T,C,H,Hkv", [(2, 256, 512, 8, 2), (1, 200, 256, 4, 4), (2, 128, 256, 2, 1)])
@pytest.mark.parametrize("packed", [False, True])
def test_llama_block_bf16(B, T, C, H, Hkv, packed):
    got, ref = run_llama(B, T, C, H, Hkv, torch.bfloat16, seed=1, packed=packed)
    _check(got, ref, torch.bfloat16, NAMES)


@pytest.mark.parametrize("dtype,B,T,C,H,Hkv", [(torch.bfloat16, 2, 256, 512, 8, 2), (torch.bfloat16, 1, 200, 256, 4, 4),
                                               (torch.float32, 2, 128, 256, 4, 2)])
def test_llama_block_with_fused_residual(dtype, B, T, C, H, Hkv):
    """h = x + attention(...) (llama3:885) in the out-projection epilogue, dx += dy in the norm VJP's store."""
    got, ref = run_llama(B, T, C, H, Hkv, dtype, packed=True, residual=True)
    _check(got, ref, dtype, NAMES)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_block_without_norm_with_fused_residual(dtype):
    """residual with norm_kind NONE: the dx GEMM's epilogue adds dy (no norm VJP kernel runs)."""
    sp, L = _sp()
    B, T, C, H, Hkv = 2, 128, 256, 4, 2
    hd = C // H
    rng = np.random.default_rng(11)
    params = llama_oracle.init_block_params(rng, C, H, Hkv)
    rd = lambda t: t.to(dtype).float()
    a = {k: rd(v).requires_grad_(True) for k, v in params["attention"].items()}
    x = rd(T_(rng.standard_normal((B, T, C)).astype(np.float32))).requires_grad_(True)
    dy = rd(T_(rng.standard_normal((B, T, C)).astype(np.float32)))
    fc = llama_oracle.precompute_freqs_cis(hd, T)
    y_ref, _ = llama_oracle.attention(a, x, llama_oracle.make_mask(T), fc, H, Hkv)
    y_ref = x + y_ref
    y_ref.backward(dy)
    cfg = sp.BlockConfig(B=B, T=T, C=C, Hq=H, Hkv=Hkv, hd=hd, dtype=dtype, norm_kind=L.NORM_NONE, rope=True, causal=True,
                         weight_layout=L.W_IN_OUT, residual=True)
    blk = sp.AttentionBlock(cfg)
    dev = lambda t: t.detach().to(dtype).to(DEV)
    p = {"wq": dev(a["wq"]), "wk": dev(a["wk"]), "wv": dev(a["wv"]), "wo": dev(a["wo"]), "rope": sp.llama_rope_table(*fc).to(DEV)}
    y = blk.forward(dev(x), p)
    g = blk.backward(dev(dy))
    torch.cuda.synchronize()
    got = {"y": y, "dx": g["dx"], "dwq": g["dwq"], "dwk": g["dwk"], "dwv": g["dwv"], "dwo": g["dwo"]}
    ref = {"y": y_ref.detach(), "dx": x.grad, "dwq": a["wq"].grad, "dwk": a["wk"].grad, "dwv": a["wv"].grad, "dwo": a["wo"].grad}
    _check(got, ref, dtype, tuple(got))


def test_llama_block_bf16_noncausal_mask_none():
    got, ref = run_llama(2, 96, 256, 4, 2, torch.bfloat16, causal=False, seed=2)
    _check(got, ref, torch.bfloat16, NAMES)


def test_llama_block_fp32_with_bf16_capable_impls_forced_simt():
    sp, L = _sp()
    got, ref = run_llama(2, 64, 128, 4, 2, torch.bfloat16, seed=3, attn_impl=L.IMPL_SIMT, gemm_impl=L.IMPL_SIMT)
    _check(got, ref, torch.bfloat16, NAMES)


# ------------------------------------------------------------------------------------------------ GPT (C2)
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("B,T,C,H", [(4, 256, 256, 1), (2, 64, 128, 4)])
def test_gpt_block(dtype, B, T, C, H):
    sp, L = _sp()
    rng = np.random.default_rng(5)
    params = gpt_oracle.init_block_params(rng, C, nonzero_bias=True)
    x = T_(rng.standard_normal((B, T, C)).astype(np.float32))
    dy = T_(rng.standard_normal((B, T, C)).astype(np.float32))
    rd = lambda t: t.to(dtype).float()
    pr = {k: rd(v) for k, v in params.items()}
    ref = gpt_oracle.block_fwd_bwd(pr, rd(x), rd(dy), H, dtype=torch.float32)
    cfg = sp.BlockConfig(B=B, T=T, C=C, Hq=H, Hkv=H, hd=C // H, dtype=dtype, norm_kind=L.NORM_LAYER, norm_eps=1e-6, rope=False,
                         causal=True, weight_layout=L.W_IN_OUT, out_bias=True)
    blk = sp.AttentionBlock(cfg)
    dev = lambda t: t.to(dtype).to(DEV)
    p = {"wqkv": dev(pr["qkv_kernel"]), "wo": dev(pr["proj_kernel"]), "bo": dev(pr["proj_bias"]), "norm_w": dev(pr["ln_scale"]),
         "norm_b": dev(pr["ln_bias"])}
    y = blk.forward(dev(rd(x)), p)
    g = blk.backward(dev(rd(dy)))
    got = {"y": y, "dx": g["dx"], "dqkv_kernel": g["dwqkv"], "dproj_kernel": g["dwo"], "dproj_bias": g["dbo"],
           "dln_scale": g["dnorm_w"], "dln_bias": g["dnorm_b"]}
    _check(got, ref, dtype, tuple(got.keys()))


@pytest.mark.parametrize("tag", ["mha", "one_head"])
def test_gpt_block_fp32_against_golden_from_notebook_source(golden_dir, tag):
    """a11 against fixtures made by exec'ing the notebook's own CausalSelfAttention / DecoderBlock cells (gpt c9, c11):
    y = attn(ln1(x)) and x + attn(ln1(x)) (gpt:420) in fp32 mode, gradients against fp64 finite differences of that source."""
    sp, L = _sp()
    g = dict(np.load(os.path.join(golden_dir, f"gpt_{tag}.npz")))
    B, Tn, C, H = (int(v) for v in g["meta"])
    for residual in (False, True):
        cfg = sp.BlockConfig(B=B, T=Tn, C=C, Hq=H, Hkv=H, hd=C // H, dtype=torch.float32, norm_kind=L.NORM_LAYER, norm_eps=1e-6,
                             rope=False, causal=True, weight_layout=L.W_IN_OUT, out_bias=True, residual=residual)
        blk = sp.AttentionBlock(cfg)
        p = {"wqkv": T_(g["qkv_kernel"]).to(DEV), "wo": T_(g["proj_kernel"]).to(DEV), "bo": T_(g["proj_bias"]).to(DEV),
             "norm_w": T_(g["ln1_scale"]).to(DEV), "norm_b": T_(g["ln1_bias"]).to(DEV)}
        y = blk.forward(T_(g["x"]).to(DEV), p)
        want = T_(g["attn64"]) + (T_(g["x"]).double() if residual else 0.0)
        assert max_rel_err(y, want) < TOL_F32, residual
        gr = blk.backward(T_(g["dy"]).to(DEV))
        torch.cuda.synchronize()
        names = {"x": "dx", "ln1_scale": "dnorm_w", "ln1_bias": "dnorm_b", "qkv_kernel": "dwqkv", "proj_kernel": "dwo", "proj_bias": "dbo"}
        for name, gname in names.items():
            grad = gr[gname].double().cpu().numpy()
            for u, fd in zip(g["fd_dir_attn_" + name], g["fd_val_attn_" + name]):
                an = float(np.sum(grad * u))
                if residual and name == "x":
                    an -= float(np.sum(g["dy"].astype(np.float64) * u))   # d(x + f(x)) = dy + f'(x)^T dy
                scale = float(np.linalg.norm(grad))  # directional derivative of a unit direction: compare on the gradient's scale
                assert abs(an - fd) <= TOL_F32 * max(scale, abs(fd)), (name, an, fd, residual)


# ------------------------------------------------------------------------------------------------ Gemma (C4 literal shape, small)
def test_gemma_block_fp32_against_golden_from_notebook_modules(golden_dir):
    sp, L = _sp()
    from solvingpapers_b200 import gemma
    g = dict(np.load(os.path.join(golden_dir, "gemma_small.npz")))
    nq = int(g["n_q"])
    B, Tn, D = g["x"].shape
    cfg = sp.BlockConfig(B=B, T=Tn, C=D, Hq=nq, Hkv=1, hd=D, dtype=torch.float32, norm_kind=L.NORM_RMS, norm_eps=1e-6,
                         norm_upcast=True, rope=True, causal=True, scale=1.0 / math.sqrt(D), weight_layout=L.W_OUT_IN)
    blk = sp.AttentionBlock(cfg)
    wqkv = torch.cat([T_(g[f"wq{i}"]) for i in range(nq)] + [T_(g["wk"]), T_(g["wv"])], dim=0).to(DEV)
    c, s = gemma.rotary_coeffs(Tn, D)
    p = {"wqkv": wqkv, "wo": T_(g["wo"]).to(DEV), "norm_w": T_(g["norm_w"]).to(DEV), "rope": sp.gemma_rope_table(c, s, D // 2)}
    y = blk.forward(T_(g["x"]).to(DEV), p)
    gr = blk.backward(T_(g["dy"]).to(DEV))
    assert max_rel_err(y, T_(g["y"])) < TOL_F32
    assert max_rel_err(gr["dx"], T_(g["dx"])) < TOL_F32
    dw = gr["dwqkv"]
    for i in range(nq):
        assert max_rel_err(dw[i * D:(i + 1) * D], T_(g[f"dwq{i}"])) < TOL_F32
    assert max_rel_err(dw[nq * D:(nq + 1) * D], T_(g["dwk"])) < TOL_F32
    assert max_rel_err(dw[(nq + 1) * D:], T_(g["dwv"])) < TOL_F32
    assert max_rel_err(gr["dwo"], T_(g["dwo"])) < TOL_F32
    assert max_rel_err(gr["dnorm_w"], T_(g["dnorm"])) < TOL_F32


def test_gemma_block_bf16_literal_width_768():
    """C4 literal: d768, 2 q heads x 768 sharing one kv head, closed-form rotary; T small enough for the CPU oracle."""
    sp, L = _sp()
    from solvingpapers_b200 import gemma
    rng = np.random.default_rng(8)
    B, Tn, D, nq = 1, 128, 768, 2
    dtype = torch.bfloat16
    params = gemma_oracle.init_block_params(rng, D, nq)
    x = T_(rng.standard_normal((B, Tn, D)).astype(np.float32))
    dy = T_(rng.standard_normal((B, Tn, D)).astype(np.float32))
    rd = lambda t: t.to(dtype).float()
    pr = {"wq": [rd(w) for w in params["wq"]], "wk": rd(params["wk"]), "wv": rd(params["wv"]), "wo": rd(params["wo"]),
          "norm_w": rd(params["norm_w"])}
    ref = gemma_oracle.block_fwd_bwd(pr, rd(x), rd(dy), dtype=torch.float32)
    cfg = sp.BlockConfig(B=B, T=Tn, C=D, Hq=nq, Hkv=1, hd=D, dtype=dtype, norm_kind=L.NORM_RMS, norm_eps=1e-6, norm_upcast=True,
                         rope=True, causal=True, scale=1.0 / math.sqrt(D), weight_layout=L.W_OUT_IN)
    blk = sp.AttentionBlock(cfg)
    dev = lambda t: t.to(dtype).to(DEV)
    c, s = gemma.rotary_coeffs(Tn, D)
    p = {"wqkv": dev(torch.cat(pr["wq"] + [pr["wk"], pr["wv"]], dim=0)), "wo": dev(pr["wo"]), "norm_w": dev(pr["norm_w"]),
         "rope": sp.gemma_rope_table(c, s, D // 2)}
    y = blk.forward(dev(rd(x)), p)
    gr = blk.backward(dev(rd(dy)))
    dw = gr["dwqkv"]
    got = {"y": y, "dx": gr["dx"], "dwk": dw[nq * D:(nq + 1) * D], "dwv": dw[(nq + 1) * D:], "dwo": gr["dwo"], "dnorm": gr["dnorm_w"]}
    _check(got, ref, dtype, tuple(got.keys()))
    for i in range(nq):
        assert max_rel_err(dw[i * D:(i + 1) * D], ref["dwq"][i]) < TOL_BF16


# ------------------------------------------------------------------------------------------------ ViT (C5)
def _vit_run(p_cpu, x, dy, H, dtype):
    sp, L = _sp()
    B, Tn, E = x.shape
    cfg = sp.BlockConfig(B=B, T=Tn, C=E, Hq=H, Hkv=H, hd=E // H, dtype=dtype, norm_kind=L.NORM_LAYER, norm_eps=1e-5, rope=False,
                         causal=False, scale=(E // H) ** -0.5, weight_layout=L.W_OUT_IN, qkv_bias=True, out_bias=True)
    blk = sp.AttentionBlock(cfg)
    dev = lambda t: t.to(dtype).to(DEV)
    p = {"wqkv": dev(p_cpu["in_proj_weight"]), "bqkv": dev(p_cpu["in_proj_bias"]), "wo": dev(p_cpu["out_proj_weight"]),
         "bo": dev(p_cpu["out_proj_bias"]), "norm_w": dev(p_cpu["ln_weight"]), "norm_b": dev(p_cpu["ln_bias"])}
    y = blk.forward(dev(x), p)
    g = blk.backward(dev(dy))
    return {"y": y, "dx": g["dx"], "din_proj_weight": g["dwqkv"], "din_proj_bias": g["dbqkv"], "dout_proj_weight": g["dwo"],
            "dout_proj_bias": g["dbo"], "dln_weight": g["dnorm_w"], "dln_bias": g["dnorm_b"]}


def test_vit_block_fp32_against_golden_from_torch_modules(golden_dir):
    g = dict(np.load(os.path.join(golden_dir, "vit_nb.npz")))
    B, Tn, E, H = (int(v) for v in g["meta"])
    keys = ("in_proj_weight", "in_proj_bias", "out_proj_weight", "out_proj_bias", "ln_weight", "ln_bias")
    got = _vit_run({k: T_(g[k]) for k in keys}, T_(g["x"]), T_(g["dy"]), H, torch.float32)
    ref = {k: T_(g[k]) for k in got}
    _check(got, ref, torch.float32, tuple(got.keys()))


def test_vit_b16_shape_bf16_ragged_197_tokens():
    """C5 shape per sample (T=197, d768, 12 heads x 64, bidirectional, biases), small batch for the CPU oracle."""
    rng = np.random.default_rng(9)
    B, Tn, E, H = 2, 197, 768, 12
    dtype = torch.bfloat16
    params = vit_oracle.init_block_params(rng, E, nonzero_bias=True)
    x = T_(rng.standard_normal((B, Tn, E)).astype(np.float32))
    dy = T_(rng.standard_normal((B, Tn, E)).astype(np.float32))
    rd = lambda t: t.to(dtype).float()
    pr = {k: rd(v) for k, v in params.items()}
    ref = vit_oracle.block_fwd_bwd(pr, rd(x), rd(dy), H, dtype=torch.float32)
    got = _vit_run(pr, rd(x), rd(dy), H, dtype)
    _check(got, ref, dtype, tuple(got.keys()))


# ------------------------------------------------------------------------------------------------ properties at full size
def test_c3_shape_properties_bf16():
    """Llama-8B shape (d4096, 32Q/8KV x 128, T2048) at B=1 - too big for the CPU oracle in seconds, so check
    size-independent properties: (1) causality: y[:, :t] does not depend on x[:, t:];  (2) linearity of the VJP in dy;
    (3) batch independence: y of sample 0 is unchanged by sample 1 (the property data parallelism rests on)."""
    sp, L = _sp()
    B, T, C, H, Hkv, hd = 2, 2048, 4096, 32, 8, 128
    dtype = torch.bfloat16
    gen = torch.Generator(device="cpu").manual_seed(0)
    cfg = sp.BlockConfig(B=B, T=T, C=C, Hq=H, Hkv=Hkv, hd=hd, dtype=dtype)
    blk = sp.AttentionBlock(cfg)
    cos, sin = llama_oracle.precompute_freqs_cis(hd, T)
    w = lambda *s: (torch.randn(*s, generator=gen) / math.sqrt(s[0])).to(dtype).to(DEV)
    p = {"wqkv": w(C, (H + 2 * Hkv) * hd), "wo": w(H * hd, C), "norm_w": torch.randn(C, generator=gen).to(dtype).to(DEV),
         "rope": sp.llama_rope_table(cos, sin).to(DEV)}
    x = torch.randn(B, T, C, generator=gen).to(dtype).to(DEV)
    y0 = blk.forward(x, p).clone()
    assert torch.isfinite(y0.float()).all()
    x2 = x.clone()
    x2[:, 1024:] = torch.randn(B, T - 1024, C, generator=gen).to(dtype).to(DEV)
    x2[1] = torch.randn(T, C, generator=gen).to(dtype).to(DEV)
    y2 = blk.forward(x2, p).clone()
    assert torch.equal(y0[0, :1024], y2[0, :1024])          # causal + batch independent, bit-exact
    assert not torch.equal(y0[0, 1024:], y2[0, 1024:])
    dy = torch.randn(B, T, C, generator=gen).to(dtype).to(DEV)
    blk.forward(x, p)
    g1 = {k: v.clone() for k, v in blk.backward(dy).items() if v is not None}
    blk.forward(x, p)
    g2 = {k: v.clone() for k, v in blk.backward((2 * dy.float()).to(dtype)).items() if v is not None}
    for k in g1:
        assert max_rel_err(g2[k].float(), 2 * g1[k].float()) < 5e-3, k   # VJP is linear in dy


def test_llama_block_fp32_ffma_gemm_path_still_matches():
    """fp32 blocks take the tensor-core split GEMM by default; gemm_impl = SIMT keeps the FFMA GEMM (and its tighter error)."""
    sp, L = _sp()
    rng = np.random.default_rng(21)
    B, T, C, H, Hkv = 2, 96, 256, 4, 2
    hd = C // H
    bp = llama_oracle.init_block_params(rng, C, H, Hkv)
    x = T_(rng.standard_normal((B, T, C)).astype(np.float32))
    dy = T_(rng.standard_normal((B, T, C)).astype(np.float32))
    ref = llama_oracle.block_fwd_bwd(bp, x, dy, H, Hkv, causal=True, dtype=torch.float64)
    cos, sin = llama_oracle.precompute_freqs_cis(hd, T)
    a = bp["attention"]
    p = {"wqkv": torch.cat([a["wq"], a["wk"], a["wv"]], dim=1).to(DEV), "wo": a["wo"].to(DEV), "norm_w": bp["attention_norm"].to(DEV),
         "rope": sp.llama_rope_table(cos, sin).to(DEV)}
    errs = {}
    for name, impl in (("split", L.IMPL_AUTO), ("ffma", L.IMPL_SIMT)):
        blk = sp.AttentionBlock(sp.BlockConfig(B, T, C, H, Hkv, hd, dtype=torch.float32, norm_kind=L.NORM_RMS, rope=True, causal=True, gemm_impl=impl))
        y = blk.forward(x.to(DEV), p)
        g = blk.backward(dy.to(DEV))
        torch.cuda.synchronize()
        errs[name] = max(max_rel_err(y, ref["y"]), max_rel_err(g["dx"], ref["dx"]), max_rel_err(g["dwo"], ref["dwo"]),
                         max_rel_err(g["dwqkv"][:, :H * hd], ref["dwq"]), max_rel_err(g["dnorm_w"], ref["dnorm"]))
    assert errs["split"] < TOL_F32 and errs["ffma"] < TOL_F32 / 10, errs


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_captured_step_replays_the_same_numbers_on_new_inputs(dtype):
    """AttentionBlock.capture_step: the recorded forward+backward, replayed after x / dy were refilled in place, gives what
    the eager calls give on those inputs (C1 dims)."""
    sp, L = _sp()
    B, T, C, H, Hkv = 16, 128, 256, 4, 2
    hd = C // H
    g = torch.Generator(device="cpu").manual_seed(5)
    r = lambda *s: torch.randn(*s, generator=g)
    cfg = sp.BlockConfig(B=B, T=T, C=C, Hq=H, Hkv=Hkv, hd=hd, dtype=dtype, norm_kind=L.NORM_RMS, norm_eps=1e-6, rope=True,
                         causal=True, weight_layout=L.W_IN_OUT, residual=True)
    cos, sin = llama_oracle.precompute_freqs_cis(hd, T)
    p = {"wqkv": (r(C, (H + 2 * Hkv) * hd) * 0.05).to(dtype).to(DEV), "wo": (r(H * hd, C) * 0.05).to(dtype).to(DEV),
         "norm_w": torch.ones(C, dtype=dtype, device=DEV), "rope": sp.llama_rope_table(cos, sin).to(DEV)}
    x, dy = r(B, T, C).to(dtype).to(DEV), r(B, T, C).to(dtype).to(DEV)
    blk = sp.AttentionBlock(cfg)
    cap = blk.capture_step(x, dy, p)
    x2, dy2 = r(B, T, C).to(dtype).to(DEV), r(B, T, C).to(dtype).to(DEV)
    x.copy_(x2)
    dy.copy_(dy2)
    y, gr = cap.replay()
    torch.cuda.synchronize()
    got = {"y": y.clone(), **{k: v.clone() for k, v in gr.items()}}
    eager = sp.AttentionBlock(cfg)
    y_e = eager.forward(x2, p)
    g_e = eager.backward(dy2)
    torch.cuda.synchronize()
    ref = {"y": y_e, **g_e}
    for k in ref:
        # same kernels on the same inputs; only the order of the fp32 reduce-adds (dQ) may differ between two runs
        assert max_rel_err(got[k], ref[k]) < (1e-5 if dtype == torch.float32 else 1e-2), k
    assert torch.isfinite(got["dx"].float()).all()
