"""Oracle (TEST INFRASTRUCTURE, never on the product path): torch-CPU restatement of
the Gemma-notebook attention block ``gemma/gemma.ipynb`` in closed form (the
notebook's dense [T,D,D] "rotary" matrix is never built).

PINNED by tests/golden/gemma_*.npz, produced by executing the notebook's own
cells c0,c1,c6,c7,c8 on CPU (tests/golden/make_golden.py).
"""
import math
import torch


def rms_norm(x, weight, eps=1e-6):
    """gemma/gemma.ipynb:139-150 - upcast to fp32, normalise, cast BACK to x's
    dtype, then multiply by the weight (init ones, :143)."""
    xf = x.float() if x.dtype != torch.float64 else x
    out = (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)).type_as(x)
    return out * weight


def rotary_coeffs(T, D, dtype=torch.float32):
    """gemma/gemma.ipynb:182-200 in closed form.  NOT a rotation: per pair
    (2j,2j+1) the block is [[c, c], [-s, s]] with ONE angle per position,
    a_t = t * 10000^(-2(t-1)/D), identical for every pair.
    Returns (cos[T], sin[T]) computed in fp32 exactly like the notebook."""
    pos = torch.arange(T).unsqueeze(1)
    theta = 10000 ** (-2 * (pos - 1) / D)
    ang = pos * theta
    return torch.cos(ang).squeeze(1).to(dtype), torch.sin(ang).squeeze(1).to(dtype)


def apply_rotary(x, cos, sin):
    """o_even = c*(x_even + x_odd);  o_odd = s*(x_odd - x_even)   (M_t @ x_t,
    gemma/gemma.ipynb:241-242), applied over the FULL width D."""
    B, T, D = x.shape
    xr = x.reshape(B, T, D // 2, 2)
    a, b = xr[..., 0], xr[..., 1]
    c = cos.reshape(1, T, 1).to(x.dtype)
    s = sin.reshape(1, T, 1).to(x.dtype)
    return torch.stack([c * a + c * b, -s * a + s * b], dim=-1).reshape(B, T, D)


def mqa(params, x):
    """gemma/gemma.ipynb:251-259 in eval mode (dropouts are identity).  Torch
    weights are [out, in].  params: wq (list of n_q [D,D]), wk [D,D], wv [D,D],
    wo [D, n_q*D].  Mask -inf then scale 1/sqrt(D) (:244-245)."""
    B, T, D = x.shape
    cos, sin = rotary_coeffs(T, D, torch.float32)
    key = torch.matmul(x, params["wk"].t())
    val = torch.matmul(x, params["wv"].t())
    rk = apply_rotary(key, cos, sin)
    outs = []
    masked = torch.tril(torch.ones((T, T)))
    for wq in params["wq"]:
        q = torch.matmul(x, wq.t())
        rq = apply_rotary(q, cos, sin)
        w = torch.matmul(rq, rk.transpose(-2, -1))
        w = w.masked_fill(masked == 0, float("-inf"))
        w = w / math.sqrt(D)
        w = torch.softmax(w, dim=-1)
        outs.append(torch.matmul(w, val))
    cat = torch.cat(outs, dim=-1)
    return torch.matmul(cat, params["wo"].t())


def attention_block(params, x):
    """gemma/gemma.ipynb:335: ``self.mqa(self.norm1(x))`` (residual outside)."""
    return mqa(params, rms_norm(x, params["norm_w"]))


def init_block_params(rng, D, n_q=2, dtype=torch.float32):
    """torch nn.Linear default init is U(-1/sqrt(in), 1/sqrt(in)); RMSNorm weight
    ones (gemma/gemma.ipynb:143)."""
    def u(shape, fan_in):
        b = 1.0 / math.sqrt(fan_in)
        return torch.from_numpy(rng.uniform(-b, b, size=shape)).to(dtype)

    return {
        "wq": [u((D, D), D) for _ in range(n_q)],
        "wk": u((D, D), D), "wv": u((D, D), D),
        "wo": u((D, n_q * D), n_q * D),
        "norm_w": torch.ones(D, dtype=dtype),
    }


def block_fwd_bwd(params, x, dy, dtype=torch.float32):
    p = {
        "wq": [w.detach().to(dtype).requires_grad_(True) for w in params["wq"]],
        "wk": params["wk"].detach().to(dtype).requires_grad_(True),
        "wv": params["wv"].detach().to(dtype).requires_grad_(True),
        "wo": params["wo"].detach().to(dtype).requires_grad_(True),
        "norm_w": params["norm_w"].detach().to(dtype).requires_grad_(True),
    }
    xx = x.detach().to(dtype).requires_grad_(True)
    y = attention_block(p, xx)
    y.backward(dy.to(dtype))
    return {
        "y": y.detach(), "dx": xx.grad, "dwq": [w.grad for w in p["wq"]],
        "dwk": p["wk"].grad, "dwv": p["wv"].grad, "dwo": p["wo"].grad, "dnorm": p["norm_w"].grad,
    }


# ---- GeGLU feed-forward ------------------------------------------------------------
def geglu(params, x):
    """gemma/gemma.ipynb:281-286 - ``linear_layer3(gelu(linear_layer1(x)) * linear_layer2(x))``; torch ``nn.Linear(bias=False)``
    weights are [out, in] (:277-279, hidden = 4 * embeddings_dims :276); ``nn.functional.gelu`` default = exact erf form."""
    gelu_res = torch.nn.functional.gelu(torch.matmul(x, params["linear_layer1"].t()))
    x_v = torch.matmul(x, params["linear_layer2"].t())
    return torch.matmul(gelu_res * x_v, params["linear_layer3"].t())


def init_geglu_params(rng, D, dtype=torch.float32):
    """torch nn.Linear default init is U(-1/sqrt(in), 1/sqrt(in)); any fixed distribution serves parity."""
    Hh = 4 * D

    def u(out_f, in_f):
        b = 1.0 / math.sqrt(in_f)
        return torch.from_numpy(rng.uniform(-b, b, size=(out_f, in_f))).to(dtype)

    return {"linear_layer1": u(Hh, D), "linear_layer2": u(Hh, D), "linear_layer3": u(D, Hh)}


def geglu_fwd_bwd(params, x, dy, dtype=torch.float32):
    p = {k: v.detach().to(dtype).requires_grad_(True) for k, v in params.items()}
    xx = x.detach().to(dtype).requires_grad_(True)
    y = geglu(p, xx)
    y.backward(dy.to(dtype))
    return {"y": y.detach(), "dx": xx.grad, "dl1": p["linear_layer1"].grad, "dl2": p["linear_layer2"].grad, "dl3": p["linear_layer3"].grad}
