"""GPU: the drop-in mirrors keep the notebooks' call surface (names, argument order, return structure, error
behaviour) and are differentiable; results checked against the oracle."""
import math

import numpy as np
import pytest
import torch

from oracle import gpt_oracle, llama_oracle, max_rel_err, vit_oracle
from oracle.metrics import TOL_BF16, TOL_F32

pytestmark = pytest.mark.gpu
DEV = "cuda"
T_ = torch.from_numpy


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_llama_functions_read_like_the_notebook(dtype):
    from solvingpapers_b200 import llama
    tol = TOL_BF16 if dtype == torch.bfloat16 else TOL_F32
    rng = np.random.default_rng(0)
    B, T, C, H, Hkv = 2, 64, 128, 4, 2
    hd = C // H
    rd = lambda t: t.to(dtype).float()
    bp = llama_oracle.init_block_params(rng, C, H, Hkv)
    bp = {"attention": {k: rd(v) for k, v in bp["attention"].items()}, "attention_norm": rd(bp["attention_norm"])}
    x = rd(T_(rng.standard_normal((B, T, C)).astype(np.float32)))
    dy = rd(T_(rng.standard_normal((B, T, C)).astype(np.float32)))
    ref = llama_oracle.block_fwd_bwd(bp, x, dy, H, Hkv, causal=True)
    # --- the notebook's own call sequence (llama3:881): attention(params['attention'], rms_norm(x, w), mask, freqs_cis, H, Hkv)
    params = {"attention": {k: v.to(dtype).to(DEV).requires_grad_(True) for k, v in bp["attention"].items()},
              "attention_norm": bp["attention_norm"].to(dtype).to(DEV).requires_grad_(True)}
    xd = x.to(dtype).to(DEV).requires_grad_(True)
    freqs_cis = llama.precompute_freqs_cis(hd, T)
    mask = torch.zeros(1, 1, T, T, device=DEV)  # contents ignored: "mask is not None" means causal
    out, (k, v) = llama.attention(params["attention"], llama.rms_norm(xd, params["attention_norm"]), mask, freqs_cis, H, Hkv)
    assert out.shape == (B, T, C) and k.shape == (B, T, Hkv, hd) and v.shape == (B, T, Hkv, hd)
    out.backward(dy.to(dtype).to(DEV))
    assert max_rel_err(out, ref["y"]) < tol and max_rel_err(k, ref["k"]) < tol and max_rel_err(v, ref["v"]) < tol
    assert max_rel_err(xd.grad, ref["dx"]) < tol
    for n in ("wq", "wk", "wv", "wo"):
        assert max_rel_err(params["attention"][n].grad, ref["d" + n]) < tol, n
    assert max_rel_err(params["attention_norm"].grad, ref["dnorm"]) < tol
    # --- fused block call gives the same numbers
    xd2 = x.to(dtype).to(DEV).requires_grad_(True)
    out2, _ = llama.attention_block({"attention": {k_: v_.detach() for k_, v_ in params["attention"].items()},
                                     "attention_norm": params["attention_norm"].detach()}, xd2, mask, freqs_cis, H, Hkv)
    assert max_rel_err(out2, ref["y"]) < tol
    # --- apply_rotary_emb and repeat_kv keep their signatures
    q4 = T_(rng.standard_normal((B, T, H, hd)).astype(np.float32))
    k4 = T_(rng.standard_normal((B, T, Hkv, hd)).astype(np.float32))
    rq, rk = llama.apply_rotary_emb(q4.to(dtype).to(DEV), k4.to(dtype).to(DEV), freqs_cis)
    oq, ok = llama_oracle.apply_rotary_emb(rd(q4), rd(k4), llama_oracle.precompute_freqs_cis(hd, T))
    assert max_rel_err(rq, oq) < tol and max_rel_err(rk, ok) < tol
    assert llama.repeat_kv(rk, 2).shape == (B, T, 2 * Hkv, hd)


def test_llama_attention_mask_none_is_bidirectional():
    from solvingpapers_b200 import llama
    rng = np.random.default_rng(1)
    B, T, C, H, Hkv = 1, 32, 64, 2, 2
    bp = llama_oracle.init_block_params(rng, C, H, Hkv)
    x = T_(rng.standard_normal((B, T, C)).astype(np.float32))
    fc = llama_oracle.precompute_freqs_cis(C // H, T)
    ref, _ = llama_oracle.attention(bp["attention"], x, None, fc, H, Hkv)
    p = {k: v.to(DEV) for k, v in bp["attention"].items()}
    out, _ = llama.attention(p, x.to(DEV), None, llama.precompute_freqs_cis(C // H, T), H, Hkv)
    assert max_rel_err(out, ref) < TOL_F32


def test_gpt_module_call_surface():
    from solvingpapers_b200 import gpt
    rng = np.random.default_rng(2)
    B, T, C, H = 2, 64, 128, 2
    p = gpt_oracle.init_block_params(rng, C, nonzero_bias=True)
    x = T_(rng.standard_normal((B, T, C)).astype(np.float32))
    ref = gpt_oracle.attention_block(p, x, H)
    attn = gpt.CausalSelfAttention(C, H, 0.1)
    params = {"ln1": {"scale": p["ln_scale"].to(DEV), "bias": p["ln_bias"].to(DEV)},
              "attn": {"qkv": {"kernel": p["qkv_kernel"].to(DEV)}, "proj": {"kernel": p["proj_kernel"].to(DEV), "bias": p["proj_bias"].to(DEV)}}}
    y1 = attn(params["attn"], gpt.layer_norm(params["ln1"], x.to(DEV)), deterministic=True)
    y2 = attn.pre_ln(params, x.to(DEV))
    assert max_rel_err(y1, ref) < TOL_F32 and max_rel_err(y2, ref) < TOL_F32


def test_vit_module_call_surface(golden_dir):
    import os
    from solvingpapers_b200 import vit
    g = dict(np.load(os.path.join(golden_dir, "vit_nb.npz")))
    B, Tn, E, H = (int(v) for v in g["meta"])
    mha = vit.MultiheadAttention(E, H, batch_first=True)
    params = {"in_proj_weight": T_(g["in_proj_weight"]).to(DEV), "in_proj_bias": T_(g["in_proj_bias"]).to(DEV),
              "out_proj.weight": T_(g["out_proj_weight"]).to(DEV), "out_proj.bias": T_(g["out_proj_bias"]).to(DEV)}
    x = T_(g["x"]).to(DEV)
    h = vit.layer_norm(T_(g["ln_weight"]).to(DEV), T_(g["ln_bias"]).to(DEV), x)
    y = mha(params, h, h, h)[0]  # ViT.ipynb:221
    assert max_rel_err(y, T_(g["y"])) < TOL_F32


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_two_same_shape_layers_train_through_transformer_block(dtype):
    """model_forward (llama3/LLaMA-jax.ipynb:925-930) calls transformer_block once per layer with identical shapes, then
    the loss head; every layer's gradients and new_cache entry must survive the other layers' forwards."""
    from solvingpapers_b200 import llama, ops
    from solvingpapers_b200.autograd import BLOCK_POOL
    tol = TOL_BF16 if dtype == torch.bfloat16 else TOL_F32
    rng = np.random.default_rng(11)
    B, T, C, H, Hkv, V, NL = 2, 64, 128, 4, 2, 256, 3
    hd = C // H
    rd = lambda t: t.to(dtype).float()

    def rd_tree(t):
        return {k: rd_tree(v) for k, v in t.items()} if isinstance(t, dict) else rd(t)

    layers = []
    for _ in range(NL):
        p = llama_oracle.init_block_params(rng, C, H, Hkv)
        p.update(llama_oracle.init_ffn_params(rng, C))
        layers.append(rd_tree(p))
    norm_f = rd(T_(rng.standard_normal(C).astype(np.float32)))
    w_out = rd(T_((rng.standard_normal((C, V)) / math.sqrt(C)).astype(np.float32)))
    x = rd(T_(rng.standard_normal((B, T, C)).astype(np.float32)))
    targets = T_(rng.integers(0, V, size=(B, T)))

    # ---- oracle: the notebook's op sequence with autograd
    def leaf(t):
        return {k: leaf(v) for k, v in t.items()} if isinstance(t, dict) else t.clone().requires_grad_(True)

    o_layers, o_nf, o_wo, o_x = [leaf(p) for p in layers], leaf(norm_f), leaf(w_out), leaf(x)
    fc = llama_oracle.precompute_freqs_cis(hd, T)
    mask = llama_oracle.make_mask(T)
    h, o_caches = o_x, []
    for p in o_layers:
        h, c = llama_oracle.transformer_block(p, h, mask, fc, H, Hkv)
        o_caches.append(c)
    o_loss = llama_oracle.loss_head(o_nf, o_wo, h, targets)
    o_loss.backward()

    # ---- the drop-in on the GPU, same call sequence
    def dev(t):
        return {k: dev(v) for k, v in t.items()} if isinstance(t, dict) else t.to(dtype).to(DEV).requires_grad_(True)

    g_layers, g_nf, g_wo, g_x = [dev(p) for p in layers], dev(norm_f), dev(w_out), dev(x)
    freqs_cis = llama.precompute_freqs_cis(hd, T)
    gmask = torch.zeros(1, 1, T, T, device=DEV)
    created0 = BLOCK_POOL.created
    for step in range(2):  # the second step must reuse the first step's workspaces, not allocate new ones
        for t in [g_nf, g_wo, g_x] + [w for p in g_layers for w in (list(p["attention"].values()) + list(p["ffn"].values()) + [p["attention_norm"], p["ffn_norm"]])]:
            t.grad = None
        h, caches = g_x, []
        for p in g_layers:
            h, c = llama.transformer_block(p, h, gmask, freqs_cis, H, Hkv)
            caches.append(c)
        logits = ops.matmul(llama.rms_norm(h, g_nf).reshape(B * T, C), g_wo)
        loss = llama.nll_loss(logits, targets.to(DEV))
        loss.backward()
    assert BLOCK_POOL.created - created0 == NL, "one workspace per simultaneously-live layer, reused across steps"
    assert abs(loss.item() - o_loss.item()) < tol * max(1.0, abs(o_loss.item()))
    gtol = 3 * tol  # three chained layers + the loss head accumulate rounding
    assert max_rel_err(g_x.grad, o_x.grad) < gtol
    for li, (gp, op, gc, oc) in enumerate(zip(g_layers, o_layers, caches, o_caches)):
        assert max_rel_err(gc[0], oc[0]) < gtol and max_rel_err(gc[1], oc[1]) < gtol, f"layer {li} cache"
        for n in ("wq", "wk", "wv", "wo"):
            assert max_rel_err(gp["attention"][n].grad, op["attention"][n].grad) < gtol, (li, n)
        for n in ("w1", "w2", "w3"):
            assert max_rel_err(gp["ffn"][n].grad, op["ffn"][n].grad) < gtol, (li, n)
        assert max_rel_err(gp["attention_norm"].grad, op["attention_norm"].grad) < gtol, li
        assert max_rel_err(gp["ffn_norm"].grad, op["ffn_norm"].grad) < gtol, li
    # distinct layers own distinct caches
    assert caches[0][0].data_ptr() != caches[1][0].data_ptr()
