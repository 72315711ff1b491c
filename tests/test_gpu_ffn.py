"""GPU parity of the gated feed-forward sub-block through spb_ffn_fwd / spb_ffn_bwd against
  (a) the golden fixture made from the notebook's own feed_forward cell (tests/golden/llama_ffn_small.npz) and
  (b) the torch-CPU oracle (oracle/llama_oracle.py: feed_forward llama3:854-855, ffn_block :886,:890) on seeded inputs.
fp32 mode 1e-4; bf16 mode: inputs rounded to bf16, oracle in fp32 on the rounded values, 2e-2 (BASELINE.json)."""
import os

import numpy as np
import pytest
import torch

from oracle import gemma_oracle, gpt_oracle, llama_oracle, max_rel_err
from oracle.metrics import TOL_BF16, TOL_F32

pytestmark = pytest.mark.gpu
DEV = "cuda"
T_ = torch.from_numpy


def tol(dtype):
    return TOL_BF16 if dtype == torch.bfloat16 else TOL_F32


def run_ffn(N_shape, C, H, dtype, seed=0, packed=True, residual=True, norm=True, params=None, h=None, dy=None):
    import solvingpapers_b200 as sp
    from solvingpapers_b200 import _lib as L
    rng = np.random.default_rng(seed)
    if params is None:
        params = llama_oracle.init_ffn_params(rng, C, H)
        h = T_(rng.standard_normal((*N_shape, C)).astype(np.float32))
        dy = T_(rng.standard_normal((*N_shape, C)).astype(np.float32))
    rd = lambda t: t.to(dtype).float()
    pr = {"ffn": {k: rd(v) for k, v in params["ffn"].items()}, "ffn_norm": rd(params["ffn_norm"])}
    hr, dyr = rd(h), rd(dy)
    ref = llama_oracle.ffn_fwd_bwd(pr, hr, dyr, dtype=torch.float32, residual=residual, norm=norm)
    n_tok = int(np.prod(N_shape))
    cfg = sp.FfnConfig(N=n_tok, C=C, H=H, dtype=dtype, norm_kind=L.NORM_RMS if norm else L.NORM_NONE, norm_eps=1e-6, residual=residual)
    ff = sp.FeedForward(cfg)
    dev = lambda t: t.to(dtype).to(DEV).contiguous()
    f = pr["ffn"]
    p = {"w2": dev(f["w2"])}
    if packed:
        p["w13"] = dev(torch.cat([f["w1"], f["w3"]], dim=1))
    else:
        p.update(w1=dev(f["w1"]), w3=dev(f["w3"]))
    if norm:
        p["norm_w"] = dev(pr["ffn_norm"])
    y = ff.forward(dev(hr).reshape(n_tok, C), p)
    g = ff.backward(dev(dyr).reshape(n_tok, C))
    torch.cuda.synchronize()
    got = {"y": y.reshape(*N_shape, C), "dx": g["dx"].reshape(*N_shape, C), "dw2": g["dw2"]}
    if packed:
        got.update(dw1=g["dw13"][:, :H], dw3=g["dw13"][:, H:])
    else:
        got.update(dw1=g["dw1"], dw3=g["dw3"])
    if norm:
        got["dnorm"] = g["dnorm_w"]
    return got, ref


def _check(got, ref, dtype):
    errs = {n: max_rel_err(v, ref[n]) for n, v in got.items()}
    bad = {n: e for n, e in errs.items() if not (e < tol(dtype))}
    assert not bad, f"over tolerance {tol(dtype)}: {bad}; all: {errs}"


def test_ffn_fp32_against_golden_from_notebook_source(golden_dir):
    g = dict(np.load(os.path.join(golden_dir, "llama_ffn_small.npz")))
    B, Tn, C, H = (int(v) for v in g["meta"])
    params = {"ffn": {k: T_(g[k]) for k in ("w1", "w2", "w3")}, "ffn_norm": T_(g["norm_w"])}
    got, ref = run_ffn((B, Tn), C, H, torch.float32, params=params, h=T_(g["h"]), dy=T_(g["dy"]), packed=False)
    _check(got, ref, torch.float32)
    # and directly against what the notebook's cell produced (fp32 run and fp64 run of the literal source)
    assert max_rel_err(got["y"], T_(g["out"])) < TOL_F32
    assert max_rel_err(got["y"], T_(g["out64"])) < TOL_F32
    for name, gname in {"h": "dx", "norm_w": "dnorm", "w1": "dw1", "w2": "dw2", "w3": "dw3"}.items():
        grad = got[gname].double().cpu().numpy()
        for u, fd in zip(g["fd_dir_" + name], g["fd_val_" + name]):
            an = float(np.sum(grad * u))
            assert abs(an - fd) <= 1e-4 * max(1.0, abs(fd)), (name, an, fd)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("packed", [False, True])
@pytest.mark.parametrize("residual", [False, True])
def test_ffn_block_notebook_dims(dtype, packed, residual):
    """C1 dims: d=256, hidden 1024, B16 x T128 tokens."""
    got, ref = run_ffn((16, 128), 256, 1024, dtype, seed=1, packed=packed, residual=residual)
    _check(got, ref, dtype)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_feed_forward_alone_without_norm(dtype):
    """plain feed_forward(params, x) (llama3:854-855), no ffn_norm, with and without the residual in the epilogue."""
    for residual in (False, True):
        got, ref = run_ffn((3, 50), 128, 384, dtype, seed=2, packed=False, residual=residual, norm=False)
        _check(got, ref, dtype)


def test_ffn_ragged_shapes_bf16():
    """token count and hidden width that do not tile 128/256: N = 197*3, H = 1000 (multiple of 8 for TMA)."""
    got, ref = run_ffn((3, 197), 192, 1000, torch.bfloat16, seed=3, packed=True, residual=True)
    _check(got, ref, torch.bfloat16)


def test_ffn_autograd_mirror_matches_oracle():
    """solvingpapers_b200.llama.ffn_block as a differentiable torch op (the drop-in call surface)."""
    from solvingpapers_b200 import llama as spl
    rng = np.random.default_rng(4)
    C, H, B, Tn = 128, 512, 2, 64
    params = llama_oracle.init_ffn_params(rng, C, H)
    h = T_(rng.standard_normal((B, Tn, C)).astype(np.float32))
    dy = T_(rng.standard_normal((B, Tn, C)).astype(np.float32))
    ref = llama_oracle.ffn_fwd_bwd(params, h, dy)
    bp = {"ffn": {k: v.to(DEV).requires_grad_(True) for k, v in params["ffn"].items()}, "ffn_norm": params["ffn_norm"].to(DEV).requires_grad_(True)}
    hd = h.to(DEV).requires_grad_(True)
    out = spl.ffn_block(bp, hd)
    out.backward(dy.to(DEV))
    got = {"y": out.detach(), "dx": hd.grad, "dw1": bp["ffn"]["w1"].grad, "dw2": bp["ffn"]["w2"].grad, "dw3": bp["ffn"]["w3"].grad,
           "dnorm": bp["ffn_norm"].grad}
    _check(got, ref, torch.float32)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_whole_transformer_block_mirror_matches_oracle(dtype):
    """llama.transformer_block (llama3:880-891): attention block + feed-forward sub-block, both residual adds inside the kernels,
    forward and backward through torch autograd against the oracle's transformer_block."""
    from solvingpapers_b200 import llama as spl
    rng = np.random.default_rng(6)
    B, Tn, C, H, Hkv = 2, 128, 256, 4, 2
    params = {**llama_oracle.init_block_params(rng, C, H, Hkv), **llama_oracle.init_ffn_params(rng, C)}
    x = T_(rng.standard_normal((B, Tn, C)).astype(np.float32))
    dy = T_(rng.standard_normal((B, Tn, C)).astype(np.float32))
    rd = lambda t: t.to(dtype).float()
    flat = lambda p, f: {k: (flat(v, f) if isinstance(v, dict) else f(v)) for k, v in p.items()}
    pr = flat(params, lambda t: rd(t).clone().requires_grad_(True))
    xr = rd(x).clone().requires_grad_(True)
    fc = llama_oracle.precompute_freqs_cis(C // H, Tn)
    out_ref, _ = llama_oracle.transformer_block(pr, xr, llama_oracle.make_mask(Tn), fc, H, Hkv)
    out_ref.backward(rd(dy))
    pd = flat(params, lambda t: t.detach().to(dtype).to(DEV).requires_grad_(True))
    xd = x.detach().to(dtype).to(DEV).requires_grad_(True)
    fcd = spl.precompute_freqs_cis(C // H, Tn)
    out, (k, v) = spl.transformer_block(pd, xd, torch.ones(1), fcd, H, Hkv)
    out.backward(dy.to(dtype).to(DEV))
    torch.cuda.synchronize()
    got = {"out": out.detach(), "dx": xd.grad}
    ref = {"out": out_ref.detach(), "dx": xr.grad}
    for grp in ("attention", "ffn"):
        for name in params[grp]:
            got[f"d{grp}.{name}"] = pd[grp][name].grad
            ref[f"d{grp}.{name}"] = pr[grp][name].grad
    for name in ("attention_norm", "ffn_norm"):
        got["d" + name] = pd[name].grad
        ref["d" + name] = pr[name].grad
    _check(got, ref, dtype)


def test_gemma_geglu_fp32_against_golden_from_notebook_module(golden_dir):
    """GeGLU (gemma:269-286): kernels against the outputs AND gradients of the notebook's own torch module."""
    from solvingpapers_b200 import gemma as spg
    g = dict(np.load(os.path.join(golden_dir, "gemma_geglu_small.npz")))
    p = {f"linear_layer{i}": T_(g[f"l{i}"]).to(DEV).requires_grad_(True) for i in (1, 2, 3)}
    x = T_(g["x"]).to(DEV).requires_grad_(True)
    y = spg.GeGLU(p)(x)
    y.backward(T_(g["dy"]).to(DEV))
    got = {"y": y.detach(), "dx": x.grad, "dl1": p["linear_layer1"].grad, "dl2": p["linear_layer2"].grad, "dl3": p["linear_layer3"].grad}
    _check(got, {k: T_(g[k]) for k in got}, torch.float32)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_gemma_geglu_literal_width_768(dtype):
    from solvingpapers_b200 import gemma as spg
    rng = np.random.default_rng(7)
    D, B, Tn = 768, 2, 96
    params = gemma_oracle.init_geglu_params(rng, D)
    x = T_(rng.standard_normal((B, Tn, D)).astype(np.float32))
    dy = T_(rng.standard_normal((B, Tn, D)).astype(np.float32))
    rd = lambda t: t.to(dtype).float()
    ref = gemma_oracle.geglu_fwd_bwd({k: rd(v) for k, v in params.items()}, rd(x), rd(dy))
    p = {k: v.to(dtype).to(DEV).requires_grad_(True) for k, v in params.items()}
    xd = x.to(dtype).to(DEV).requires_grad_(True)
    y = spg.GeGLU(p)(xd)
    y.backward(dy.to(dtype).to(DEV))
    got = {"y": y.detach(), "dx": xd.grad, "dl1": p["linear_layer1"].grad, "dl2": p["linear_layer2"].grad, "dl3": p["linear_layer3"].grad}
    _check(got, ref, dtype)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_gpt_mlp_with_biases(dtype):
    """GELU MLP (gpt:376-389): flax parameter tree, non-zero biases, tanh GELU; C2 width 256."""
    from solvingpapers_b200 import gpt as spgpt
    rng = np.random.default_rng(8)
    C, B, Tn = 256, 4, 100
    params = gpt_oracle.init_mlp_params(rng, C)
    x = T_(rng.standard_normal((B, Tn, C)).astype(np.float32))
    dy = T_(rng.standard_normal((B, Tn, C)).astype(np.float32))
    rd = lambda t: t.to(dtype).float()
    ref = gpt_oracle.mlp_fwd_bwd({k: rd(v) for k, v in params.items()}, rd(x), rd(dy))
    d = lambda t: t.to(dtype).to(DEV).requires_grad_(True)
    tree = {"params": {"fc1": {"kernel": d(params["fc1_kernel"]), "bias": d(params["fc1_bias"])},
                       "fc2": {"kernel": d(params["fc2_kernel"]), "bias": d(params["fc2_bias"])}}}
    xd = d(x)
    y = spgpt.MLP(emb_dim=C, dropout_rate=0.1).apply(tree, xd, deterministic=True)
    y.backward(dy.to(dtype).to(DEV))
    t = tree["params"]
    got = {"y": y.detach(), "dx": xd.grad, "dfc1_kernel": t["fc1"]["kernel"].grad, "dfc1_bias": t["fc1"]["bias"].grad,
           "dfc2_kernel": t["fc2"]["kernel"].grad, "dfc2_bias": t["fc2"]["bias"].grad}
    _check(got, ref, dtype)


def test_ffn_rejects_unknown_activation_and_bad_shapes():
    import ctypes as C
    from solvingpapers_b200 import _lib as L
    lib = L.load()
    d = L.FfnDesc(128, 64, 256, L.F32, L.NORM_RMS, 1e-6, 0, 7, L.W_IN_OUT, 0, 0)
    ws = L.FfnWs()
    assert lib.spb_ffn_ws_layout(C.byref(d), C.byref(ws)) == -2  # SPB_E_UNSUPPORTED
    assert b"activation" in lib.spb_last_error()
    d = L.FfnDesc(0, 64, 256, L.F32, L.NORM_RMS, 1e-6, 0, 0, L.W_IN_OUT, 0, 0)
    assert lib.spb_ffn_ws_layout(C.byref(d), C.byref(ws)) == -1


def test_gpt_mlp_fp32_against_golden_from_notebook_source(golden_dir):
    """GELU-MLP third of f2 against fixtures from the notebook's own MLP / DecoderBlock cells (gpt c10, c11):
    ``x + mlp(ln2(x))`` (gpt:421) through spb_ffn_fwd/bwd with the LayerNorm in front and the residual in the w2 epilogue."""
    import os
    import solvingpapers_b200 as sp
    from solvingpapers_b200 import _lib as L
    g = dict(np.load(os.path.join(golden_dir, "gpt_mha.npz")))
    B, Tn, C, H = (int(v) for v in g["meta"])
    T_ = torch.from_numpy
    cfg = sp.FfnConfig(N=B * Tn, C=C, H=4 * C, dtype=torch.float32, norm_kind=L.NORM_LAYER, norm_eps=1e-6, act=L.ACT_GELU_TANH, residual=True)
    ff = sp.FeedForward(cfg)
    p = {"w1": T_(g["fc1_kernel"]).cuda(), "b1": T_(g["fc1_bias"]).cuda(), "w2": T_(g["fc2_kernel"]).cuda(), "b2": T_(g["fc2_bias"]).cuda(),
         "norm_w": T_(g["ln2_scale"]).cuda(), "norm_b": T_(g["ln2_bias"]).cuda()}
    x = T_(g["x"]).cuda().reshape(B * Tn, C)
    y = ff.forward(x, p)
    want = T_(g["mlp64"]) + T_(g["x"]).double()
    assert max_rel_err(y.reshape(B, Tn, C), want) < TOL_F32
    gr = ff.backward(T_(g["dy"]).cuda().reshape(B * Tn, C))
    torch.cuda.synchronize()
    names = {"x": "dx", "ln2_scale": "dnorm_w", "ln2_bias": "dnorm_b", "fc1_kernel": "dw1", "fc1_bias": "db1", "fc2_kernel": "dw2", "fc2_bias": "db2"}
    for name, gname in names.items():
        grad = gr[gname].double().cpu().numpy().reshape(g["fd_dir_mlp_" + name][0].shape)
        for u, fd in zip(g["fd_dir_mlp_" + name], g["fd_val_mlp_" + name]):
            an = float(np.sum(grad * u))
            if name == "x":
                an -= float(np.sum(g["dy"].astype(np.float64) * u))
            assert abs(an - fd) <= TOL_F32 * max(float(np.linalg.norm(grad)), abs(fd)), (name, an, fd)
