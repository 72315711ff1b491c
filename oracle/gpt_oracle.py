"""Oracle (TEST INFRASTRUCTURE, never on the product path): torch-CPU restatement of
the GPT pre-LN causal attention block of ``gpt/gpt-jax.ipynb``.

Pinned by tests/golden/gpt_*.npz: fixtures made by executing the notebook's own cells c9 (CausalSelfAttention),
c10 (MLP) and c11 (DecoderBlock) with ``jnp`` bound to numpy and ``nn`` bound to a numpy stand-in for flax.linen
(tests/golden/make_golden.py: flax/jax cannot be installed here).  What stays restated rather than executed is
flax's own third-party arithmetic, from its published behaviour (version never printed by the notebook):
  * ``nn.LayerNorm()``: eps 1e-6, biased variance over the last axis, learnable
    scale (init ones) and bias (init zeros): (x-mean)*rsqrt(var+eps)*scale+bias.
  * ``nn.Dense(n, use_bias)``: x @ kernel[in,out] (+ bias).
  * ``nn.gelu``: approximate=True (tanh form).
"""
import math
import numpy as np
import torch


def layer_norm(x, scale, bias, eps=1e-6):
    """flax nn.LayerNorm as used at gpt/gpt-jax.ipynb:414 (ln1)."""
    mean = x.mean(dim=-1, keepdim=True)
    var = ((x - mean) ** 2).mean(dim=-1, keepdim=True)
    return (x - mean) * torch.rsqrt(var + eps) * scale + bias


def causal_self_attention(params, x, num_heads):
    """gpt/gpt-jax.ipynb:333-357 with deterministic=True (dropout off, the
    ``eval_step`` path).  params: qkv_kernel [C,3C] (no bias, :329), proj_kernel
    [C,C] + proj_bias [C] (:330).  q|k|v = split(qkv, 3, -1) (:336).  Scale is a
    division by sqrt(hd) (:344); masked scores are REPLACED by -1e4 (:347-348)."""
    B, T, C = x.shape
    hd = C // num_heads
    qkv = torch.matmul(x, params["qkv_kernel"])
    q, k, v = torch.split(qkv, C, dim=-1)
    q = q.reshape(B, T, num_heads, hd).permute(0, 2, 1, 3)
    k = k.reshape(B, T, num_heads, hd).permute(0, 2, 1, 3)
    v = v.reshape(B, T, num_heads, hd).permute(0, 2, 1, 3)
    w = torch.matmul(q, k.permute(0, 1, 3, 2)) / math.sqrt(hd)
    causal = torch.tril(torch.ones((T, T), dtype=x.dtype)).reshape(1, 1, T, T)
    w = torch.where(causal == 0, torch.tensor(-1e4, dtype=x.dtype), w)
    w = torch.softmax(w, dim=-1)
    out = torch.matmul(w, v)
    out = out.permute(0, 2, 1, 3).reshape(B, T, -1)
    return torch.matmul(out, params["proj_kernel"]) + params["proj_bias"]


def attention_block(params, x, num_heads):
    """gpt/gpt-jax.ipynb:420: ``self.attn(self.ln1(x))`` (residual add outside)."""
    return causal_self_attention(params, layer_norm(x, params["ln_scale"], params["ln_bias"]), num_heads)


def init_block_params(rng, dim, dtype=torch.float32, nonzero_bias=False):
    """flax defaults: Dense kernel lecun_normal (N(0,1)/sqrt(fan_in) up to the
    truncation), biases zeros, LN scale ones / bias zeros.  ``nonzero_bias`` gives the
    one non-zero-bias parity case SURVEY 8(d) asks for."""
    def n(shape, scale):
        return torch.from_numpy(rng.standard_normal(shape) * scale).to(dtype)

    p = {
        "qkv_kernel": n((dim, 3 * dim), 1.0 / math.sqrt(dim)),
        "proj_kernel": n((dim, dim), 1.0 / math.sqrt(dim)),
        "proj_bias": torch.zeros(dim, dtype=dtype),
        "ln_scale": torch.ones(dim, dtype=dtype),
        "ln_bias": torch.zeros(dim, dtype=dtype),
    }
    if nonzero_bias:
        p["proj_bias"] = n((dim,), 0.5)
        p["ln_scale"] = 1.0 + n((dim,), 0.3)
        p["ln_bias"] = n((dim,), 0.3)
    return p


def block_fwd_bwd(params, x, dy, num_heads, dtype=torch.float32):
    p = {k: v.detach().to(dtype).requires_grad_(True) for k, v in params.items()}
    xx = x.detach().to(dtype).requires_grad_(True)
    y = attention_block(p, xx, num_heads)
    y.backward(dy.to(dtype))
    return {
        "y": y.detach(), "dx": xx.grad,
        "dqkv_kernel": p["qkv_kernel"].grad, "dproj_kernel": p["proj_kernel"].grad,
        "dproj_bias": p["proj_bias"].grad, "dln_scale": p["ln_scale"].grad, "dln_bias": p["ln_bias"].grad,
    }


# ---- the GELU MLP --------------------------------------------------------------------
def mlp(params, x):
    """gpt/gpt-jax.ipynb:385-389 with deterministic=True (dropout off): ``fc2(nn.gelu(fc1(x)))``; flax ``nn.Dense`` = x @ kernel
    + bias with kernel [in, out] (:381-382, hidden = 4 * emb_dim); flax ``nn.gelu`` defaults to approximate=True (tanh form)."""
    hdn = torch.nn.functional.gelu(torch.matmul(x, params["fc1_kernel"]) + params["fc1_bias"], approximate="tanh")
    return torch.matmul(hdn, params["fc2_kernel"]) + params["fc2_bias"]


def init_mlp_params(rng, dim, dtype=torch.float32, nonzero_bias=True):
    """flax Dense: lecun_normal kernel, zero bias; non-zero biases here so that the bias path is exercised."""
    Hh = 4 * dim
    w = lambda i, o: torch.from_numpy(rng.standard_normal((i, o)) / math.sqrt(i)).to(dtype)
    b = lambda n: torch.from_numpy(rng.standard_normal((n,)) * 0.1 if nonzero_bias else np.zeros((n,))).to(dtype)
    return {"fc1_kernel": w(dim, Hh), "fc1_bias": b(Hh), "fc2_kernel": w(Hh, dim), "fc2_bias": b(dim)}


def mlp_fwd_bwd(params, x, dy, dtype=torch.float32):
    p = {k: v.detach().to(dtype).requires_grad_(True) for k, v in params.items()}
    xx = x.detach().to(dtype).requires_grad_(True)
    y = mlp(p, xx)
    y.backward(dy.to(dtype))
    return {"y": y.detach(), "dx": xx.grad, **{"d" + k: v.grad for k, v in p.items()}}
