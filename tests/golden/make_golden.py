#!/usr/bin/env python
"""Generate the golden fixtures in tests/golden/ FROM THE REFERENCE ITSELF.

Run in the build container only (needs /root/reference; the GPU box has no copy):

    python tests/golden/make_golden.py

What is executed:
  * llama : the notebook's own cell SOURCE TEXT (llama3/LLaMA-jax.ipynb cells c15
            rms_norm, c16 precompute_freqs_cis, c17 apply_rotary_emb, c18 repeat_kv,
            c24 attention) exec'd with ``jnp`` bound to numpy (every jnp call the
            cells make - mean, square, reciprocal, sqrt, arange, outer, complex64,
            exp, reshape, stack, real, imag, repeat, dot, matmul, concatenate - has
            the same name and semantics in numpy) and ``jax.nn.softmax`` bound to the
            max-subtracted softmax.  JAX itself is not installable here.  Forward
            values are stored from the literal fp32/complex64 run; gradients are
            central finite differences of the same literal source run in fp64
            (``complex64`` bound to complex128 for that run only, so the difference
            quotient is not drowned by fp32 rounding).
  * gemma : the notebook's own torch cells c0 (minus the unused ``tokenizers``
            import), c1, c6, c7, c8 with ``args.device='cpu'``, modules in eval().
            Gradients by torch autograd THROUGH THE NOTEBOOK'S MODULES.
  * vit   : ``nn.LayerNorm`` + ``nn.MultiheadAttention(E,H,batch_first=True)`` called
            exactly like ViT.ipynb:217-221; gradients by autograd.
  * gpt   : the notebook's own module cells (gpt/gpt-jax.ipynb c9 CausalSelfAttention,
            c10 MLP, c11 DecoderBlock) exec'd with ``jnp`` bound to numpy and ``nn``
            bound to a small numpy stand-in for the flax.linen names the cells use
            (Module with annotated fields + setup(), Dense, LayerNorm, Dropout, softmax,
            gelu) - so the split order, head reshapes/transposes, /sqrt(head_dim),
            ``where(mask == 0, -1e4, .)`` and the pre-LN residual structure that run are the
            notebook's own lines; flax's third-party semantics (Dense = x @ kernel + bias,
            LayerNorm eps 1e-6 / biased variance, gelu approximate=True) are what the
            stand-in restates.  Gradients: finite differences of that source in fp64.
"""
import json
import math
import os
import sys
import types

import numpy as np
import torch

REF = "/root/reference"
OUT = os.path.dirname(os.path.abspath(__file__))


def cell_src(nb_path, idx):
    nb = json.load(open(os.path.join(REF, nb_path)))
    return "".join(nb["cells"][idx]["source"])


# --------------------------------------------------------------------------- llama
def _softmax(x, axis=-1):
    m = np.max(x, axis=axis, keepdims=True)
    e = np.exp(x - m)
    return e / np.sum(e, axis=axis, keepdims=True)


def llama_namespace(high_precision):
    jnp = types.ModuleType("jnp_as_numpy")
    for name in dir(np):
        if not name.startswith("__"):
            setattr(jnp, name, getattr(np, name))
    if high_precision:
        jnp.complex64 = lambda a: np.asarray(a, dtype=np.complex128)
        jnp.float32 = np.float64
    jax = types.SimpleNamespace(nn=types.SimpleNamespace(softmax=_softmax))
    ns = {"jnp": jnp, "jax": jax, "math": math}
    for c in (15, 16, 17, 18, 24):
        exec(cell_src("llama3/LLaMA-jax.ipynb", c), ns)
    return ns


def make_llama(tag, B, T, C, H, Hkv, causal, seed):
    rng = np.random.default_rng(seed)
    hd = C // H
    f32 = np.float32
    x = rng.standard_normal((B, T, C)).astype(f32)
    dy = rng.standard_normal((B, T, C)).astype(f32)
    p = {
        "wq": (rng.standard_normal((C, H * hd)) / math.sqrt(C)).astype(f32),
        "wk": (rng.standard_normal((C, Hkv * hd)) / math.sqrt(C)).astype(f32),
        "wv": (rng.standard_normal((C, Hkv * hd)) / math.sqrt(C)).astype(f32),
        "wo": (rng.standard_normal((H * hd, C)) / math.sqrt(H * hd)).astype(f32),
    }
    nw = rng.standard_normal((C,)).astype(f32)  # N(0,1), llama3:748

    def run(ns, x_, p_, nw_, dt):
        fc = ns["precompute_freqs_cis"](hd, T)
        if causal:  # llama3/LLaMA-jax.ipynb:920-923
            mask = np.tril(np.ones((T, T)))
            mask = np.where(mask == 0, -1e9, 0.0).astype(dt)[None, None, :, :]
        else:
            mask = None
        y, (k, v) = ns["attention"](p_, ns["rms_norm"](x_, nw_), mask, fc, H, Hkv)
        return y, k, v

    ns32 = llama_namespace(False)
    y, k, v = run(ns32, x, p, nw, f32)
    assert y.dtype == np.float32, y.dtype
    # freqs table of the literal code (fp32 complex64), for the exponent quirk test
    fc = ns32["precompute_freqs_cis"](hd, T)

    ns64 = llama_namespace(True)
    x64 = x.astype(np.float64)
    p64 = {k_: v_.astype(np.float64) for k_, v_ in p.items()}
    nw64 = nw.astype(np.float64)
    dy64 = dy.astype(np.float64)
    y64, k64, v64 = run(ns64, x64, p64, nw64, np.float64)

    def loss(x_, p_, nw_):
        return float(np.sum(run(ns64, x_, p_, nw_, np.float64)[0] * dy64))

    h = 1e-5
    out = dict(x=x, dy=dy, norm_w=nw, y=y, k=k, v=v, y64=y64, k64=k64, v64=v64,
               fc_re=np.real(fc).astype(f32), fc_im=np.imag(fc).astype(f32),
               meta=np.array([B, T, C, H, Hkv, int(causal)], dtype=np.int64), **p)
    # directional finite-difference derivatives, 3 random directions per tensor
    for name in ("x", "norm_w", "wq", "wk", "wv", "wo"):
        base = x64 if name == "x" else nw64 if name == "norm_w" else p64[name]
        dirs, vals = [], []
        for _ in range(3):
            u = rng.standard_normal(base.shape)
            u /= np.linalg.norm(u)

            def at(eps):
                xx, pp, nn = x64, dict(p64), nw64
                if name == "x":
                    xx = x64 + eps * u
                elif name == "norm_w":
                    nn = nw64 + eps * u
                else:
                    pp[name] = p64[name] + eps * u
                return loss(xx, pp, nn)

            # 4th-order central difference
            d = (-at(2 * h) + 8 * at(h) - 8 * at(-h) + at(-2 * h)) / (12 * h)
            dirs.append(u)
            vals.append(d)
        out["fd_dir_" + name] = np.stack(dirs)
        out["fd_val_" + name] = np.array(vals)
    # full FD gradient of the norm weight (C entries)
    g = np.zeros(C)
    for i in range(C):
        e = np.zeros(C)
        e[i] = 1.0
        g[i] = (loss(x64, p64, nw64 + h * e) - loss(x64, p64, nw64 - h * e)) / (2 * h)
    out["fd_grad_norm_w"] = g
    np.savez_compressed(os.path.join(OUT, f"llama_{tag}.npz"), **out)
    print("llama", tag, "y", y.shape, "max|y|", np.abs(y).max())


def make_llama_decode(tag, B, Tc, Tq, C, H, Hkv, use_mask, seed):
    """the notebook's own attention cell (c24) run through its cache branch (llama3:815-819): a prefill of Tc
    tokens with cache=None gives (k, v); then Tq new tokens with cache=(k, v), position=Tc.  use_mask passes the model's
    causal mask (only valid for Tq = 1: the cell adds mask[:, :, :T, :T] to scores [B, H, T, Tc+T])."""
    rng = np.random.default_rng(seed)
    hd = C // H
    f32 = np.float32
    Tm = Tc + Tq
    x0 = rng.standard_normal((B, Tc, C)).astype(f32)
    x1 = rng.standard_normal((B, Tq, C)).astype(f32)
    p = {"wq": (rng.standard_normal((C, H * hd)) / math.sqrt(C)).astype(f32), "wk": (rng.standard_normal((C, Hkv * hd)) / math.sqrt(C)).astype(f32),
         "wv": (rng.standard_normal((C, Hkv * hd)) / math.sqrt(C)).astype(f32), "wo": (rng.standard_normal((H * hd, C)) / math.sqrt(H * hd)).astype(f32)}
    nw = rng.standard_normal((C,)).astype(f32)
    out = dict(x0=x0, x1=x1, norm_w=nw, meta=np.array([B, Tc, Tq, C, H, Hkv, int(use_mask)], dtype=np.int64), **p)
    for high, sfx in ((False, ""), (True, "64")):
        ns = llama_namespace(high)
        dt = np.float64 if high else f32
        cast = lambda a: a.astype(dt)
        pp = {k_: cast(v_) for k_, v_ in p.items()}
        fc = ns["precompute_freqs_cis"](hd, Tm)
        mask = np.where(np.tril(np.ones((Tm, Tm))) == 0, -1e9, 0.0).astype(dt)[None, None, :, :]  # llama3:920-923
        _, cache = ns["attention"](pp, ns["rms_norm"](cast(x0), cast(nw)), mask, fc, H, Hkv)          # prefill, cache=None
        y, (k, v) = ns["attention"](pp, ns["rms_norm"](cast(x1), cast(nw)), mask if use_mask else None, fc, H, Hkv, cache=cache, position=Tc)
        assert k.shape == (B, Tm, Hkv, hd) and y.shape == (B, Tq, C)
        out["cache_k" + sfx], out["cache_v" + sfx] = cache
        out["y" + sfx], out["k" + sfx], out["v" + sfx] = y, k, v
    np.savez_compressed(os.path.join(OUT, f"llama_decode_{tag}.npz"), **out)
    print("llama decode", tag, "y", out["y"].shape, "max|y|", np.abs(out["y"]).max())


def make_llama_ffn(tag, B, T, C, seed):
    """the notebook's own feed_forward cell (c25) and the two lines of transformer_block that use it
    (llama3/LLaMA-jax.ipynb:886,890: ffn_output = feed_forward(params['ffn'], rms_norm(h, params['ffn_norm'])); out = h + ffn_output),
    executed with jnp bound to numpy; fp32 outputs, fp64 outputs, finite-difference directional derivatives."""
    rng = np.random.default_rng(seed)
    f32 = np.float32
    Hh = 4 * C  # init_ffn_weights, llama3:716-718
    h_in = rng.standard_normal((B, T, C)).astype(f32)
    dy = rng.standard_normal((B, T, C)).astype(f32)
    p = {"w1": (rng.standard_normal((C, Hh)) / math.sqrt(C)).astype(f32), "w2": (rng.standard_normal((Hh, C)) / math.sqrt(Hh)).astype(f32),
         "w3": (rng.standard_normal((C, Hh)) / math.sqrt(C)).astype(f32)}
    nw = rng.standard_normal((C,)).astype(f32)

    def namespace(high):
        ns = llama_namespace(high)
        ns["jax"].nn.silu = lambda z: z / (1.0 + np.exp(-z))
        exec(cell_src("llama3/LLaMA-jax.ipynb", 25), ns)
        return ns

    def run(ns, h_, p_, nw_):
        ffn_output = ns["feed_forward"](p_, ns["rms_norm"](h_, nw_))  # :886
        return h_ + ffn_output, ffn_output                             # :890

    ns32, ns64 = namespace(False), namespace(True)
    out32, ffn32 = run(ns32, h_in, p, nw)
    assert out32.dtype == np.float32
    h64, dy64, nw64 = h_in.astype(np.float64), dy.astype(np.float64), nw.astype(np.float64)
    p64 = {k_: v_.astype(np.float64) for k_, v_ in p.items()}
    out64, ffn64 = run(ns64, h64, p64, nw64)

    def loss(h_, p_, nw_):
        return float(np.sum(run(ns64, h_, p_, nw_)[0] * dy64))

    hstep = 1e-5
    out = dict(h=h_in, dy=dy, norm_w=nw, out=out32, ffn=ffn32, out64=out64, ffn64=ffn64, meta=np.array([B, T, C, Hh], dtype=np.int64), **p)
    for name in ("h", "norm_w", "w1", "w2", "w3"):
        base = h64 if name == "h" else nw64 if name == "norm_w" else p64[name]
        dirs, vals = [], []
        for _ in range(3):
            u = rng.standard_normal(base.shape)
            u /= np.linalg.norm(u)

            def at(eps):
                hh, pp, nn = h64, dict(p64), nw64
                if name == "h":
                    hh = h64 + eps * u
                elif name == "norm_w":
                    nn = nw64 + eps * u
                else:
                    pp[name] = p64[name] + eps * u
                return loss(hh, pp, nn)

            vals.append((-at(2 * hstep) + 8 * at(hstep) - 8 * at(-hstep) + at(-2 * hstep)) / (12 * hstep))
            dirs.append(u)
        out["fd_dir_" + name] = np.stack(dirs)
        out["fd_val_" + name] = np.array(vals)
    np.savez_compressed(os.path.join(OUT, f"llama_ffn_{tag}.npz"), **out)
    print("llama ffn", tag, "out", out32.shape, "max|out|", np.abs(out32).max())


def make_llama_loss(tag, B, T, V, seed):
    """the notebook's own compute_loss cell (c28, llama3:956-968) executed with jnp bound to numpy;
    model_forward is replaced by a stub that returns the given logits, so exactly the reshape / log_softmax / take_along_axis /
    mean lines of the notebook run."""
    rng = np.random.default_rng(seed)
    logits = (rng.standard_normal((B, T, V)) * 3).astype(np.float32)
    targets = rng.integers(0, V, size=(B, T))
    out = dict(logits=logits, targets=targets, meta=np.array([B, T, V], dtype=np.int64))
    for high, key in ((False, "loss"), (True, "loss64")):
        ns = llama_namespace(high)

        def log_softmax(x, axis=-1):
            m = np.max(x, axis=axis, keepdims=True)
            return x - m - np.log(np.sum(np.exp(x - m), axis=axis, keepdims=True))

        ns["jax"].nn.log_softmax = log_softmax
        lg = logits.astype(np.float64) if high else logits
        ns["model_forward"] = lambda params, inputs, config, lg=lg: (lg, None)
        ns["config"] = types.SimpleNamespace(vocab_size=V)
        exec(cell_src("llama3/LLaMA-jax.ipynb", 28), ns)
        out[key] = np.asarray(ns["compute_loss"](None, (None, targets)))
    np.savez_compressed(os.path.join(OUT, f"llama_loss_{tag}.npz"), **out)
    print("llama loss", tag, float(out["loss"]), float(out["loss64"]))


# --------------------------------------------------------------------------- gemma
def make_gemma(tag, B, T, D, seed):
    ns = {}
    src0 = cell_src("gemma/gemma.ipynb", 0).replace("from tokenizers import Tokenizer\n", "")
    exec(src0, ns)
    exec(cell_src("gemma/gemma.ipynb", 1), ns)
    args = ns["args"]
    args.device = "cpu"
    args.embeddings_dims = D
    args.block_size = T
    for c in (6, 7, 8):
        exec(cell_src("gemma/gemma.ipynb", c), ns)
    torch.manual_seed(seed)
    norm = ns["Normalization"](embeddings_dims=D)
    mqa = ns["MQA"](embeddings_dims=D, block_size=T, no_of_kv_heads=args.no_kv_heads,
                    no_of_heads=args.no_of_heads)
    # RotaryEmbeddings default-constructs with the module-level block_size at class
    # definition time; force the instance values we want.
    mqa.rotary_matrix.block_size = T
    mqa.rotary_matrix.embeddings_dims = D
    norm.eval()
    mqa.eval()
    n_params = sum(p.numel() for p in mqa.parameters())
    rng = np.random.default_rng(seed)
    x = torch.from_numpy(rng.standard_normal((B, T, D)).astype(np.float32)).requires_grad_(True)
    dy = torch.from_numpy(rng.standard_normal((B, T, D)).astype(np.float32))
    with torch.no_grad():
        norm.rmsnorm_layer.weight.copy_(torch.from_numpy(1.0 + 0.3 * rng.standard_normal(D).astype(np.float32)))
    y = mqa(norm(x))
    y.backward(dy)
    matrix = mqa.rotary_matrix(T)[:T]
    out = dict(
        x=x.detach().numpy(), dy=dy.numpy(), y=y.detach().numpy(), dx=x.grad.numpy(),
        norm_w=norm.rmsnorm_layer.weight.detach().numpy(), dnorm=norm.rmsnorm_layer.weight.grad.numpy(),
        wk=mqa.key.weight.detach().numpy(), dwk=mqa.key.weight.grad.numpy(),
        wv=mqa.value.weight.detach().numpy(), dwv=mqa.value.weight.grad.numpy(),
        wo=mqa.linear_layer.weight.detach().numpy(), dwo=mqa.linear_layer.weight.grad.numpy(),
        n_q=np.array(len(mqa.multi_query)), n_params=np.array(n_params),
        # pair (0,1) block of the literal rotary matrix per position: [[M00,M01],[M10,M11]]
        rot_block=matrix[:, 0:2, 0:2].detach().numpy(),
        rot_block_last=matrix[:, D - 2:D, D - 2:D].detach().numpy(),
    )
    for i, q in enumerate(mqa.multi_query):
        out[f"wq{i}"] = q.weight.detach().numpy()
        out[f"dwq{i}"] = q.weight.grad.numpy()
    np.savez_compressed(os.path.join(OUT, f"gemma_{tag}.npz"), **out)
    print("gemma", tag, "y", tuple(y.shape), "MQA params", n_params)


def make_gemma_geglu(tag, B, T, D, seed):
    """the notebook's own GeGLU module (gemma/gemma.ipynb cell 9, :269-286) run forward + backward on CPU."""
    ns = {}
    exec(cell_src("gemma/gemma.ipynb", 0).replace("from tokenizers import Tokenizer\n", ""), ns)
    exec(cell_src("gemma/gemma.ipynb", 1), ns)
    args = ns["args"]
    args.device = "cpu"
    args.embeddings_dims = D
    args.block_size = T
    exec(cell_src("gemma/gemma.ipynb", 9), ns)
    torch.manual_seed(seed)
    m = ns["GeGLU"](block_size=T, embeddings_dims=D)
    rng = np.random.default_rng(seed)
    x = torch.from_numpy(rng.standard_normal((B, T, D)).astype(np.float32)).requires_grad_(True)
    dy = torch.from_numpy(rng.standard_normal((B, T, D)).astype(np.float32))
    y = m(x)
    y.backward(dy)
    out = dict(x=x.detach().numpy(), dy=dy.numpy(), y=y.detach().numpy(), dx=x.grad.numpy(), meta=np.array([B, T, D], dtype=np.int64))
    for i, lin in enumerate((m.linear_layer1, m.linear_layer2, m.linear_layer3), 1):
        out[f"l{i}"] = lin.weight.detach().numpy()
        out[f"dl{i}"] = lin.weight.grad.numpy()
    np.savez_compressed(os.path.join(OUT, f"gemma_geglu_{tag}.npz"), **out)
    print("gemma geglu", tag, "y", tuple(y.shape))


def gemma_param_count_pin():
    """MQA at notebook defaults has 3,538,944 parameters."""
    ns = {}
    exec(cell_src("gemma/gemma.ipynb", 0).replace("from tokenizers import Tokenizer\n", ""), ns)
    exec(cell_src("gemma/gemma.ipynb", 1), ns)
    ns["args"].device = "cpu"
    for c in (6, 7, 8):
        exec(cell_src("gemma/gemma.ipynb", c), ns)
    m = ns["MQA"]()
    n = sum(p.numel() for p in m.parameters())
    assert n == 3538944, n
    return n


# --------------------------------------------------------------------------- vit
def make_vit(tag, B, T, E, H, seed):
    torch.manual_seed(seed)
    ln = torch.nn.LayerNorm(E)
    att = torch.nn.MultiheadAttention(E, H, batch_first=True)
    rng = np.random.default_rng(seed)
    with torch.no_grad():
        ln.weight.copy_(torch.from_numpy(1.0 + 0.3 * rng.standard_normal(E).astype(np.float32)))
        ln.bias.copy_(torch.from_numpy(0.3 * rng.standard_normal(E).astype(np.float32)))
        att.in_proj_bias.copy_(torch.from_numpy(0.3 * rng.standard_normal(3 * E).astype(np.float32)))
        att.out_proj.bias.copy_(torch.from_numpy(0.3 * rng.standard_normal(E).astype(np.float32)))
    x = torch.from_numpy(rng.standard_normal((B, T, E)).astype(np.float32)).requires_grad_(True)
    dy = torch.from_numpy(rng.standard_normal((B, T, E)).astype(np.float32))
    h = ln(x)
    y = att(h, h, h)[0]  # ViT.ipynb:221
    y.backward(dy)
    out = dict(
        x=x.detach().numpy(), dy=dy.numpy(), y=y.detach().numpy(), dx=x.grad.numpy(),
        ln_weight=ln.weight.detach().numpy(), ln_bias=ln.bias.detach().numpy(),
        dln_weight=ln.weight.grad.numpy(), dln_bias=ln.bias.grad.numpy(),
        in_proj_weight=att.in_proj_weight.detach().numpy(), din_proj_weight=att.in_proj_weight.grad.numpy(),
        in_proj_bias=att.in_proj_bias.detach().numpy(), din_proj_bias=att.in_proj_bias.grad.numpy(),
        out_proj_weight=att.out_proj.weight.detach().numpy(), dout_proj_weight=att.out_proj.weight.grad.numpy(),
        out_proj_bias=att.out_proj.bias.detach().numpy(), dout_proj_bias=att.out_proj.bias.grad.numpy(),
        meta=np.array([B, T, E, H], dtype=np.int64),
    )
    np.savez_compressed(os.path.join(OUT, f"vit_{tag}.npz"), **out)
    print("vit", tag, "y", tuple(y.shape))


# --------------------------------------------------------------------------- gpt
def flax_standin(dtype):
    """numpy stand-in for the handful of flax.linen names gpt/gpt-jax.ipynb cells c9-c11 use.  Parameters are plain
    attributes (``kernel``, ``bias``, ``scale``) set by the caller after construction."""
    nn = types.ModuleType("flax_linen_standin")

    class Module:
        def __init__(self, *args, **kw):
            fields = list(getattr(type(self), "__annotations__", {}))
            for name, val in zip(fields, args):
                setattr(self, name, val)
            for name, val in kw.items():
                setattr(self, name, val)
            self.setup()

        def setup(self):
            pass

    class Dense(Module):  # flax.linen.Dense: y = x @ kernel[in, out] (+ bias), use_bias defaults to True
        def __init__(self, features, use_bias=True):
            self.features, self.use_bias, self.kernel, self.bias = features, use_bias, None, None

        def __call__(self, x):
            assert self.kernel.shape[-1] == self.features
            y = np.matmul(x, self.kernel)
            return y + self.bias if self.use_bias else y

    class LayerNorm(Module):  # flax.linen.LayerNorm defaults: epsilon 1e-6, biased variance over the last axis, scale + bias
        def __init__(self, epsilon=1e-6):
            self.epsilon, self.scale, self.bias = epsilon, None, None

        def __call__(self, x):
            mean = np.mean(x, axis=-1, keepdims=True)
            var = np.mean(np.square(x), axis=-1, keepdims=True) - np.square(mean)  # flax: E[x^2] - E[x]^2 (use_fast_variance)
            var = np.maximum(var, 0.0)
            return ((x - mean) * (1.0 / np.sqrt(var + self.epsilon)) * self.scale + self.bias).astype(dtype)

    class Dropout(Module):
        def __init__(self, rate):
            self.rate = rate

        def __call__(self, x, deterministic=True):
            assert deterministic, "dropout has no reproducible RNG stream here; the bench/parity path is deterministic=True"
            return x

    def gelu(x, approximate=True):  # flax.linen.gelu default approximate=True (tanh form)
        assert approximate
        return 0.5 * x * (1.0 + np.tanh(math.sqrt(2.0 / math.pi) * (x + 0.044715 * x ** 3)))

    nn.Module, nn.Dense, nn.LayerNorm, nn.Dropout, nn.softmax, nn.gelu = Module, Dense, LayerNorm, Dropout, _softmax, gelu
    return nn


def gpt_namespace(dtype):
    jnp = types.ModuleType("jnp_as_numpy")
    for name in dir(np):
        if not name.startswith("__"):
            setattr(jnp, name, getattr(np, name))
    # jax: jnp.sqrt of a Python int is a (weakly typed) float32 scalar, so `x / jnp.sqrt(head_dim)` keeps x's dtype; numpy's
    # np.float64 scalar would promote an fp32 x to fp64
    jnp.sqrt = lambda v: dtype(np.sqrt(v)) if isinstance(v, (int, float)) else np.sqrt(v)
    ns = {"jnp": jnp, "nn": flax_standin(dtype), "np": np}
    for c in (9, 10, 11):
        exec(cell_src("gpt/gpt-jax.ipynb", c), ns)
    return ns


def make_gpt(tag, B, T, C, H, seed):
    """Fixtures for a11 (CausalSelfAttention behind ln1, with and without the DecoderBlock residual) and the GELU-MLP third
    of f2 (MLP behind ln2 + residual), from the notebook's own DecoderBlock.__call__ lines (gpt/gpt-jax.ipynb:419-422)."""
    rng = np.random.default_rng(seed)
    f32 = np.float32
    x = rng.standard_normal((B, T, C)).astype(f32)
    dy = rng.standard_normal((B, T, C)).astype(f32)
    p = {"ln1_scale": (1.0 + 0.3 * rng.standard_normal(C)).astype(f32), "ln1_bias": (0.3 * rng.standard_normal(C)).astype(f32),
         "qkv_kernel": (rng.standard_normal((C, 3 * C)) / math.sqrt(C)).astype(f32),
         "proj_kernel": (rng.standard_normal((C, C)) / math.sqrt(C)).astype(f32), "proj_bias": (0.5 * rng.standard_normal(C)).astype(f32),
         "ln2_scale": (1.0 + 0.3 * rng.standard_normal(C)).astype(f32), "ln2_bias": (0.3 * rng.standard_normal(C)).astype(f32),
         "fc1_kernel": (rng.standard_normal((C, 4 * C)) / math.sqrt(C)).astype(f32), "fc1_bias": (0.1 * rng.standard_normal(4 * C)).astype(f32),
         "fc2_kernel": (rng.standard_normal((4 * C, C)) / math.sqrt(4 * C)).astype(f32), "fc2_bias": (0.1 * rng.standard_normal(C)).astype(f32)}

    def build(dtype, pp):
        ns = gpt_namespace(dtype)
        blk = ns["DecoderBlock"](C, H, 0.1)
        blk.ln1.scale, blk.ln1.bias = pp["ln1_scale"], pp["ln1_bias"]
        blk.attn.qkv.kernel = pp["qkv_kernel"]
        blk.attn.proj.kernel, blk.attn.proj.bias = pp["proj_kernel"], pp["proj_bias"]
        blk.ln2.scale, blk.ln2.bias = pp["ln2_scale"], pp["ln2_bias"]
        blk.mlp.fc1.kernel, blk.mlp.fc1.bias = pp["fc1_kernel"], pp["fc1_bias"]
        blk.mlp.fc2.kernel, blk.mlp.fc2.bias = pp["fc2_kernel"], pp["fc2_bias"]
        return blk

    def run(blk, x_):
        attn = blk.attn(blk.ln1(x_), deterministic=True)          # the call inside gpt:420
        mlp = blk.mlp(blk.ln2(x_), deterministic=True)            # the call inside gpt:421 (on the same x_, stand-alone)
        full = blk(x_, deterministic=True)                        # gpt:419-422 both halves chained
        return attn, mlp, full

    attn32, mlp32, full32 = run(build(f32, p), x)
    assert attn32.dtype == np.float32 and mlp32.dtype == np.float32 and full32.dtype == np.float32, (attn32.dtype, mlp32.dtype, full32.dtype)
    p64 = {k: v.astype(np.float64) for k, v in p.items()}
    x64, dy64 = x.astype(np.float64), dy.astype(np.float64)
    attn64, mlp64, full64 = run(build(np.float64, p64), x64)
    out = dict(x=x, dy=dy, attn=attn32, mlp=mlp32, block=full32, attn64=attn64, mlp64=mlp64, block64=full64,
               meta=np.array([B, T, C, H], dtype=np.int64), **p)
    hstep = 1e-5
    # directional derivatives of <attn(ln1(x)), dy>, <mlp(ln2(x)), dy> and <block(x), dy> in fp64
    for which, idx, names in (("attn", 0, ("x", "ln1_scale", "ln1_bias", "qkv_kernel", "proj_kernel", "proj_bias")),
                              ("mlp", 1, ("x", "ln2_scale", "ln2_bias", "fc1_kernel", "fc1_bias", "fc2_kernel", "fc2_bias")),
                              ("block", 2, ("x", "qkv_kernel", "fc1_kernel"))):
        def loss(x_, pp):
            return float(np.sum(run(build(np.float64, pp), x_)[idx] * dy64))

        for name in names:
            base = x64 if name == "x" else p64[name]
            dirs, vals = [], []
            for _ in range(2):
                u = rng.standard_normal(base.shape)
                u /= np.linalg.norm(u)

                def at(eps):
                    if name == "x":
                        return loss(x64 + eps * u, p64)
                    pp = dict(p64)
                    pp[name] = p64[name] + eps * u
                    return loss(x64, pp)

                vals.append((-at(2 * hstep) + 8 * at(hstep) - 8 * at(-hstep) + at(-2 * hstep)) / (12 * hstep))
                dirs.append(u)
            out[f"fd_dir_{which}_{name}"] = np.stack(dirs)
            out[f"fd_val_{which}_{name}"] = np.array(vals)
    np.savez_compressed(os.path.join(OUT, f"gpt_{tag}.npz"), **out)
    print("gpt", tag, "attn", attn32.shape, "max|attn|", np.abs(attn32).max(), "max|block|", np.abs(full32).max())


if __name__ == "__main__":
    if not os.path.isdir(REF):
        sys.exit("needs /root/reference (build container only)")
    make_llama("gqa_causal", B=2, T=16, C=64, H=4, Hkv=2, causal=True, seed=0)
    make_llama("mha_nomask", B=1, T=9, C=32, H=2, Hkv=2, causal=False, seed=1)
    make_llama("mqa_causal", B=2, T=12, C=64, H=4, Hkv=1, causal=True, seed=2)
    make_llama_decode("one_token", B=2, Tc=21, Tq=1, C=128, H=2, Hkv=1, use_mask=True, seed=10)   # hd 64 like the notebook
    make_llama_decode("chunk_nomask", B=1, Tc=9, Tq=3, C=128, H=2, Hkv=2, use_mask=False, seed=11)
    make_llama_ffn("small", B=2, T=12, C=32, seed=5)
    make_llama_loss("small", B=2, T=16, V=257, seed=7)
    make_gemma("small", B=2, T=16, D=64, seed=3)
    make_gemma_geglu("small", B=2, T=12, D=32, seed=6)
    print("gemma MQA default param count", gemma_param_count_pin())
    make_vit("nb", B=3, T=17, E=64, H=4, seed=4)  # the notebook's literal shape
    make_gpt("mha", B=2, T=12, C=32, H=2, seed=8)
    make_gpt("one_head", B=1, T=16, C=32, H=1, seed=9)  # the notebook's own num_heads = 1 (gpt:296)
