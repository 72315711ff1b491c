"""Oracle (TEST INFRASTRUCTURE, never on the product path): torch-CPU restatement of
the ViT-notebook attention block ``vision transformer/ViT.ipynb:217-221``:
``nn.LayerNorm`` (torch eps 1e-5) -> ``nn.MultiheadAttention(E, H, batch_first=True)
(x, x, x)[0]``.  The arithmetic of MultiheadAttention lives inside PyTorch
(version unpinned in the reference; this box's torch 2.11 is the pin).

PINNED by tests/golden/vit_*.npz (the real torch modules run on CPU).
"""
import math
import torch


def layer_norm(x, weight, bias, eps=1e-5):
    """torch nn.LayerNorm(E) as at ViT.ipynb:205 - biased variance, eps 1e-5."""
    mean = x.mean(dim=-1, keepdim=True)
    var = ((x - mean) ** 2).mean(dim=-1, keepdim=True)
    return (x - mean) / torch.sqrt(var + eps) * weight + bias


def mha(params, x, num_heads):
    """nn.MultiheadAttention self-attention, ViT.ipynb:208,221.  Packed
    in_proj_weight [3E,E] (+ in_proj_bias [3E]) -> q,k,v; q scaled by hd^-0.5 BEFORE
    q.k^T; no mask; out_proj weight [E,E] + bias.  All torch [out,in]."""
    B, T, E = x.shape
    hd = E // num_heads
    qkv = torch.matmul(x, params["in_proj_weight"].t()) + params["in_proj_bias"]
    q, k, v = qkv.split(E, dim=-1)
    q = q.reshape(B, T, num_heads, hd).permute(0, 2, 1, 3) * (hd ** -0.5)
    k = k.reshape(B, T, num_heads, hd).permute(0, 2, 1, 3)
    v = v.reshape(B, T, num_heads, hd).permute(0, 2, 1, 3)
    w = torch.softmax(torch.matmul(q, k.transpose(-2, -1)), dim=-1)
    o = torch.matmul(w, v).permute(0, 2, 1, 3).reshape(B, T, E)
    return torch.matmul(o, params["out_proj_weight"].t()) + params["out_proj_bias"]


def attention_block(params, x, num_heads):
    return mha(params, layer_norm(x, params["ln_weight"], params["ln_bias"]), num_heads)


def init_block_params(rng, E, dtype=torch.float32, nonzero_bias=True):
    def n(shape, scale):
        return torch.from_numpy(rng.standard_normal(shape) * scale).to(dtype)

    p = {
        "in_proj_weight": n((3 * E, E), 1.0 / math.sqrt(E)),
        "in_proj_bias": torch.zeros(3 * E, dtype=dtype),
        "out_proj_weight": n((E, E), 1.0 / math.sqrt(E)),
        "out_proj_bias": torch.zeros(E, dtype=dtype),
        "ln_weight": torch.ones(E, dtype=dtype),
        "ln_bias": torch.zeros(E, dtype=dtype),
    }
    if nonzero_bias:
        p["in_proj_bias"] = n((3 * E,), 0.3)
        p["out_proj_bias"] = n((E,), 0.3)
        p["ln_weight"] = 1.0 + n((E,), 0.3)
        p["ln_bias"] = n((E,), 0.3)
    return p


def block_fwd_bwd(params, x, dy, num_heads, dtype=torch.float32):
    p = {k: v.detach().to(dtype).requires_grad_(True) for k, v in params.items()}
    xx = x.detach().to(dtype).requires_grad_(True)
    y = attention_block(p, xx, num_heads)
    y.backward(dy.to(dtype))
    out = {"y": y.detach(), "dx": xx.grad}
    out.update({"d" + k: v.grad for k, v in p.items()})
    return out
