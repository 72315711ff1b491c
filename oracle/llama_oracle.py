"""Oracle (TEST INFRASTRUCTURE, never on the product path): torch-CPU restatement of
the Llama-3 attention block of the reference notebook ``llama3/LLaMA-jax.ipynb``.

Every function cites the raw JSON line of the notebook it follows (SURVEY.md
citation convention).  Op ORDER is kept exactly as written in the notebook:
dot -> reshape -> RoPE -> repeat -> transpose -> matmul -> /sqrt -> +mask -> softmax
-> matmul -> transpose/reshape -> dot.  The backward is torch autograd of this
forward (the notebook has no hand-written VJP: ``jax.value_and_grad`` at :995).

Pinned by tests/golden/llama_*.npz (made by executing the notebook's own cell
source under numpy, see tests/golden/make_golden.py).
"""
import math
import torch


def rms_norm(x, weight, eps=1e-6):
    """llama3/LLaMA-jax.ipynb:536-538 - eps INSIDE the sqrt, x*weight first."""
    var = torch.mean(torch.square(x), dim=-1, keepdim=True)
    return x * weight * torch.reciprocal(torch.sqrt(var + eps))


def precompute_freqs_cis(dim, end, theta=10000.0, dtype=torch.float32):
    """llama3/LLaMA-jax.ipynb:563-567 - NOTE exponent is j/dim (not 2j/dim):
    ``arange(0, dim//2)/dim``.  Returned as (cos, sin) pair [end, dim//2] instead
    of complex64 (same numbers: exp(i*a) = cos a + i sin a)."""
    j = torch.arange(0, dim // 2, dtype=torch.float32)
    freqs = 1.0 / (theta ** (j / dim))
    t = torch.arange(end, dtype=torch.float32)
    ang = torch.outer(t, freqs)  # fp32 like the notebook
    return torch.cos(ang).to(dtype), torch.sin(ang).to(dtype)


def apply_rotary_emb(xq, xk, freqs_cis):
    """llama3/LLaMA-jax.ipynb:592-601 - ADJACENT pairs (2j, 2j+1) form the complex
    number; table broadcast over batch and heads ([1,T,1,hd/2], :596)."""
    cos, sin = freqs_cis

    def rot(x):
        xr = x.reshape(*x.shape[:-1], -1, 2)
        a, b = xr[..., 0], xr[..., 1]
        c = cos.reshape(1, cos.shape[0], 1, cos.shape[1]).to(x.dtype)
        s = sin.reshape(1, sin.shape[0], 1, sin.shape[1]).to(x.dtype)
        re = a * c - b * s
        im = a * s + b * c
        return torch.stack([re, im], dim=-1).reshape(x.shape)

    return rot(xq), rot(xk)


def repeat_kv(x, n_rep):
    """llama3/LLaMA-jax.ipynb:626-627 - jnp.repeat on axis 2: CONSECUTIVE grouping
    (kv head g serves q heads g*n_rep .. g*n_rep+n_rep-1)."""
    return x if n_rep == 1 else torch.repeat_interleave(x, n_rep, dim=2)


def make_mask(max_seq_len, dtype=torch.float32):
    """llama3/LLaMA-jax.ipynb:920-923 - additive mask, -1e9 on the strict upper
    triangle, cast to the activation dtype, shape [1,1,T,T]."""
    m = torch.tril(torch.ones((max_seq_len, max_seq_len), dtype=torch.float64))
    m = torch.where(m == 0, torch.tensor(-1e9, dtype=torch.float64), torch.tensor(0.0, dtype=torch.float64))
    return m.to(dtype)[None, None, :, :]


def attention(params, x, mask, freqs_cis, n_heads, n_kv_heads, cache=None, position=0):
    """llama3/LLaMA-jax.ipynb:809-829.  Weights are [in, out] (:684-687)."""
    B, T, C = x.shape
    head_dim = C // n_heads
    q = torch.matmul(x, params["wq"]).reshape(B, T, n_heads, head_dim)
    k = torch.matmul(x, params["wk"]).reshape(B, T, n_kv_heads, head_dim)
    v = torch.matmul(x, params["wv"]).reshape(B, T, n_kv_heads, head_dim)
    cos, sin = freqs_cis
    q, k = apply_rotary_emb(q, k, (cos[position:position + T], sin[position:position + T]))
    if cache is not None:
        k = torch.cat([cache[0], k], dim=1)
        v = torch.cat([cache[1], v], dim=1)
    new_cache = (k, v)
    k = repeat_kv(k, n_heads // n_kv_heads)
    v = repeat_kv(v, n_heads // n_kv_heads)
    q, k, v = (t.permute(0, 2, 1, 3) for t in (q, k, v))
    scores = torch.matmul(q, k.permute(0, 1, 3, 2)) / math.sqrt(head_dim)
    if mask is not None:
        scores = scores + mask[:, :, :T, :T]
    scores = torch.softmax(scores, dim=-1)
    output = torch.matmul(scores, v)
    output = output.permute(0, 2, 1, 3).reshape(B, T, -1)
    return torch.matmul(output, params["wo"]), new_cache


def attention_block(block_params, x, mask, freqs_cis, n_heads, n_kv_heads):
    """The hot path as the notebook calls it, llama3/LLaMA-jax.ipynb:881:
    ``attention(params['attention'], rms_norm(x, params['attention_norm']), ...)``.
    The residual add (:885) is outside the block."""
    return attention(block_params["attention"], rms_norm(x, block_params["attention_norm"]),
                     mask, freqs_cis, n_heads, n_kv_heads)


def init_block_params(rng, dim, n_heads, n_kv_heads, dtype=torch.float32):
    """Input distributions of llama3/LLaMA-jax.ipynb:652-654,680-688,748: projection
    weights N(0,1)/sqrt(fan_in); attention_norm weight N(0,1) (scale=1.0, NOT ones).
    ``rng`` is a numpy Generator (JAX's threefry stream cannot be reproduced)."""
    hd = dim // n_heads

    def w(shape, scale=None):
        scale = 1.0 / math.sqrt(shape[0]) if scale is None else scale
        return torch.from_numpy(rng.standard_normal(shape) * scale).to(dtype)

    return {
        "attention": {
            "wq": w((dim, n_heads * hd)),
            "wk": w((dim, n_kv_heads * hd)),
            "wv": w((dim, n_kv_heads * hd)),
            "wo": w((n_heads * hd, dim)),
        },
        "attention_norm": w((dim,), scale=1.0),
    }


def block_fwd_bwd(block_params, x, dy, n_heads, n_kv_heads, causal=True, dtype=torch.float32,
                  theta=10000.0, residual=False):
    """Forward + autograd backward of the block in ``dtype`` (fp32 or fp64).
    Returns dict with y, k, v, dx, dwq, dwk, dwv, dwo, dnorm.
    residual=True follows the caller, transformer_block llama3:881-885: ``h = x + attention(p, rms_norm(x, w), ...)``."""
    B, T, C = x.shape
    hd = C // n_heads
    p = {
        "attention": {k: v.detach().to(dtype).requires_grad_(True) for k, v in block_params["attention"].items()},
        "attention_norm": block_params["attention_norm"].detach().to(dtype).requires_grad_(True),
    }
    xx = x.detach().to(dtype).requires_grad_(True)
    fc = precompute_freqs_cis(hd, T, theta, dtype=dtype)
    mask = make_mask(T, dtype) if causal else None
    y, (k, v) = attention_block(p, xx, mask, fc, n_heads, n_kv_heads)
    if residual:
        y = xx + y  # llama3:885
    y.backward(dy.to(dtype))
    a = p["attention"]
    return {
        "y": y.detach(), "k": k.detach(), "v": v.detach(), "dx": xx.grad,
        "dwq": a["wq"].grad, "dwk": a["wk"].grad, "dwv": a["wv"].grad, "dwo": a["wo"].grad,
        "dnorm": p["attention_norm"].grad,
    }


# ---- the gated feed-forward sub-block -----------------------------------------------
def feed_forward(params, x):
    """llama3/LLaMA-jax.ipynb:854-855 - ``jnp.dot(jax.nn.silu(jnp.dot(x, w3)) * jnp.dot(x, w1), w2)``; weights [in, out]
    (init_ffn_weights :713-718: w1, w3 [dim, 4*dim], w2 [4*dim, dim])."""
    return torch.matmul(torch.nn.functional.silu(torch.matmul(x, params["w3"])) * torch.matmul(x, params["w1"]), params["w2"])


def ffn_block(block_params, h, residual=True):
    """Second half of transformer_block, llama3/LLaMA-jax.ipynb:886,890:
    ``ffn_output = feed_forward(params['ffn'], rms_norm(h, params['ffn_norm']))``; ``out = h + ffn_output``."""
    out = feed_forward(block_params["ffn"], rms_norm(h, block_params["ffn_norm"]))
    return h + out if residual else out


def init_ffn_params(rng, dim, hidden=None, dtype=torch.float32):
    """llama3/LLaMA-jax.ipynb:713-718 (N(0,1)/sqrt(fan_in), hidden = 4*dim) and :749 (ffn_norm ~ N(0,1))."""
    hidden = 4 * dim if hidden is None else hidden

    def w(shape, scale=None):
        scale = 1.0 / math.sqrt(shape[0]) if scale is None else scale
        return torch.from_numpy(rng.standard_normal(shape) * scale).to(dtype)

    return {"ffn": {"w1": w((dim, hidden)), "w2": w((hidden, dim)), "w3": w((dim, hidden))}, "ffn_norm": w((dim,), scale=1.0)}


def ffn_fwd_bwd(block_params, h, dy, dtype=torch.float32, residual=True, norm=True):
    """Forward + autograd backward of the feed-forward sub-block.  Returns dict with y, dx, dw1, dw2, dw3, dnorm."""
    f = {k: v.detach().to(dtype).requires_grad_(True) for k, v in block_params["ffn"].items()}
    nw = block_params["ffn_norm"].detach().to(dtype).requires_grad_(True)
    hh = h.detach().to(dtype).requires_grad_(True)
    if norm:
        y = ffn_block({"ffn": f, "ffn_norm": nw}, hh, residual=residual)
    else:
        y = feed_forward(f, hh)
        if residual:
            y = hh + y
    y.backward(dy.to(dtype))
    return {"y": y.detach(), "dx": hh.grad, "dw1": f["w1"].grad, "dw2": f["w2"].grad, "dw3": f["w3"].grad,
            "dnorm": nw.grad if norm else None}


def transformer_block(params, x, mask, freqs_cis, n_heads, n_kv_heads):
    """llama3/LLaMA-jax.ipynb:880-891 with training=False (the only way the notebook calls it, :927)."""
    attn_output, new_cache = attention(params["attention"], rms_norm(x, params["attention_norm"]), mask, freqs_cis, n_heads, n_kv_heads)
    h = x + attn_output
    ffn_output = feed_forward(params["ffn"], rms_norm(h, params["ffn_norm"]))
    return h + ffn_output, new_cache


# ---- the loss head -------------------------------------------------------------------
def nll_loss(logits, targets):
    """llama3/LLaMA-jax.ipynb:959-967: logits.reshape(-1, V); targets.reshape(-1);
    ``loss = -mean(take_along_axis(log_softmax(logits), targets[:, None], axis=1))``."""
    V = logits.shape[-1]
    lg = logits.reshape(-1, V)
    tg = targets.reshape(-1).long()
    lp = torch.log_softmax(lg, dim=-1)
    return -torch.mean(torch.gather(lp, 1, tg[:, None]))


def loss_head(norm_f, output_w, h, targets):
    """llama3/LLaMA-jax.ipynb:929-930 + :959-967: ``logits = dot(rms_norm(h, norm_f), output)`` then the NLL above."""
    return nll_loss(torch.matmul(rms_norm(h, norm_f), output_w), targets)
