"""Drop-in for the hot-path functions of llama3/LLaMA-jax.ipynb: SAME names, argument order, defaults and error
behaviour as the notebook's cells (c15 rms_norm :536, c16 precompute_freqs_cis :563, c17 apply_rotary_emb :592,
c18 repeat_kv :626, c24 attention :809) - arrays are torch CUDA tensors because JAX is not installable in this
image (the jax.ffi face of the same library is ffi/spb_jax_ffi.cc + ffi/spb_jax.py).  Differentiable through
torch.autograd the way the notebook's functions are through jax.value_and_grad (:995).

Contract carried over from tracing: under jit the mask is a tracer whose values cannot be read, so
``mask is not None`` means CAUSAL and ``None`` means bidirectional; the tensor itself is ignored.
``cache`` / ``position`` select the KV-cache decode step (llama3:815-819, spb_block_decode); combinations the
notebook's own code cannot compute (a mask with several new tokens, position without a cache) raise.
"""
import math

import torch

from . import _lib as L
from .autograd import BlockFn, NormFn, RopeFn
from .block import BlockConfig, llama_rope_table


def rms_norm(x, weight, eps=1e-6):
    """llama3/LLaMA-jax.ipynb:536-538."""
    return NormFn.apply(L.NORM_RMS, float(eps), False, x, weight, None)


def precompute_freqs_cis(dim: int, end: int, theta: float = 10000.0, device="cuda"):
    """llama3/LLaMA-jax.ipynb:563-567 (host-side table; fp32 math in the notebook's op order, exponent j/dim).
    Returns complex64 [end, dim//2]."""
    freqs = 1.0 / (theta ** (torch.arange(0, dim // 2, dtype=torch.float32) / dim))
    t = torch.arange(end, dtype=torch.float32)
    freqs = torch.outer(t, freqs)
    return torch.polar(torch.ones_like(freqs), freqs).to(device)


def _table(freqs_cis):
    return llama_rope_table(freqs_cis.real, freqs_cis.imag)


def apply_rotary_emb(xq, xk, freqs_cis):
    """llama3/LLaMA-jax.ipynb:592-601; xq [B,T,Hq,hd], xk [B,T,Hkv,hd], freqs_cis [T, hd/2] complex64."""
    tab = _table(freqs_cis)
    return RopeFn.apply(xq, tab), RopeFn.apply(xk, tab)


def repeat_kv(x, n_rep):
    """llama3/LLaMA-jax.ipynb:626-627.  Kept for API parity only: attention() never materialises repeated K/V
    (the kernels index kv_head = q_head // n_rep)."""
    return x if n_rep == 1 else torch.repeat_interleave(x, n_rep, dim=2)


def _cfg(x, n_heads, n_kv_heads, head_dim, mask, norm, eps=1e-6, attn_impl=L.IMPL_AUTO, gemm_impl=L.IMPL_AUTO, residual=False):
    B, T, C = x.shape
    return BlockConfig(B=B, T=T, C=C, Hq=n_heads, Hkv=n_kv_heads, hd=head_dim, dtype=x.dtype,
                       norm_kind=L.NORM_RMS if norm else L.NORM_NONE, norm_eps=eps, rope=True,
                       causal=mask is not None, scale=1.0 / math.sqrt(head_dim), weight_layout=L.W_IN_OUT,
                       attn_impl=attn_impl, gemm_impl=gemm_impl, residual=residual)


_DECODERS = {}


def _decode(params, x, mask, freqs_cis, n_heads, n_kv_heads, cache, position, norm_w=None, eps=1e-6, residual=False):
    """The cache branch of llama3/LLaMA-jax.ipynb:815-819 through spb_block_decode (no autograd: the notebook only samples
    with it).  ``cache`` is the notebook's ``(k, v)`` tuple ([B, Tc, Hkv, hd]; a new, longer tuple is returned, as
    ``jnp.concatenate`` does) or a ``KVCache`` (appended in place; its views are returned)."""
    from .block import DecodeBlock, KVCache
    B, T, C = x.shape
    hd = C // n_heads
    in_place = isinstance(cache, KVCache)
    if in_place:
        kv = cache
    else:
        if cache[0].shape[1] != position:
            raise ValueError(f"cache holds {cache[0].shape[1]} tokens but position = {position} (llama3:815 slices freqs_cis from `position`)")
        kv = KVCache.from_tensors(cache[0], cache[1], extra=T)
    if kv.length != position:
        raise ValueError(f"cache holds {kv.length} tokens but position = {position}")
    if mask is not None and T > 1:
        # llama3:824-825 adds mask[:, :, :T, :T] to scores [B, H, T, Tc + T]: that only broadcasts for T == 1
        raise ValueError("a mask with a non-empty cache is only defined for a single new token (llama3:824-825 does not broadcast otherwise)")
    cfg = BlockConfig(B=B, T=T, C=C, Hq=n_heads, Hkv=n_kv_heads, hd=hd, dtype=x.dtype, norm_kind=L.NORM_RMS if norm_w is not None else L.NORM_NONE,
                      norm_eps=eps, rope=True, causal=False, scale=1.0 / math.sqrt(hd), weight_layout=L.W_IN_OUT, residual=residual)
    key = (tuple(sorted(cfg.__dict__.items())), kv.capacity, str(x.device))
    dec = _DECODERS.get(key)
    if dec is None:
        if len(_DECODERS) > 8:
            _DECODERS.clear()
        dec = _DECODERS[key] = DecodeBlock(cfg, kv.capacity, device=x.device)
    p = {"wq": params["wq"], "wk": params["wk"], "wv": params["wv"], "wo": params["wo"], "rope": _table(freqs_cis[:position + T])}
    if norm_w is not None:
        p["norm_w"] = norm_w
    with torch.no_grad():
        y = dec.step(x.contiguous(), p, kv)
    return y, (kv.k, kv.v)


def attention(params, x, mask, freqs_cis, n_heads, n_kv_heads, cache=None, position=0, comm=None):
    """llama3/LLaMA-jax.ipynb:809-829.  Returns (out [B,T,C], (k, v)) with k, v post-rotary, un-repeated.
    ``cache=None, position=0`` is the training / prefill path (differentiable); with a cache it is one decode step."""
    if cache is not None:
        return _decode(params, x, mask, freqs_cis, n_heads, n_kv_heads, cache, position)
    if position != 0:
        raise ValueError("position != 0 without a cache: llama3:815 would rotate at `position` but attend to nothing before it")
    B, T, C = x.shape
    cfg = _cfg(x, n_heads, n_kv_heads, C // n_heads, mask, norm=False)
    tab = _table(freqs_cis[position:position + T])
    names = ("wq", "wk", "wv", "wo", "rope")
    y, k, v = BlockFn.apply(cfg, names, comm, x, params["wq"], params["wk"], params["wv"], params["wo"], tab)
    return y, (k, v)


def attention_block(block_params, x, mask, freqs_cis, n_heads, n_kv_heads, eps=1e-6, comm=None, residual=False):
    """The fused hot path exactly as transformer_block calls it (llama3/LLaMA-jax.ipynb:881):
    ``attention(params['attention'], rms_norm(x, params['attention_norm']), mask, freqs_cis, n_heads, n_kv_heads)``
    - one C-ABI call forward, one backward.  residual=True returns ``x + attention(...)``, the next line of the caller
    (``h = x + attn_output``, :885), with the add inside the output projection's epilogue (and ``dx += dy`` inside the norm
    VJP) instead of two extra passes over [B,T,C]."""
    B, T, C = x.shape
    cfg = _cfg(x, n_heads, n_kv_heads, C // n_heads, mask, norm=True, eps=eps, residual=residual)
    tab = _table(freqs_cis[:T])
    a = block_params["attention"]
    names = ("wq", "wk", "wv", "wo", "norm_w", "rope")
    y, k, v = BlockFn.apply(cfg, names, comm, x, a["wq"], a["wk"], a["wv"], a["wo"], block_params["attention_norm"], tab)
    return y, (k, v)


def feed_forward(params, x, comm=None):
    """llama3/LLaMA-jax.ipynb:854-855: ``dot(silu(dot(x, w3)) * dot(x, w1), w2)`` with w1, w3 [C, 4C], w2 [4C, C]."""
    from .ffn import FfnConfig, FfnFn
    C_ = x.shape[-1]
    cfg = FfnConfig(N=x.numel() // C_, C=C_, H=params["w1"].shape[1], dtype=x.dtype, norm_kind=L.NORM_NONE)
    y = FfnFn.apply(cfg, ("w1", "w3", "w2"), comm, x.reshape(-1, C_), params["w1"], params["w3"], params["w2"])
    return y.reshape(x.shape)


def ffn_block(block_params, h, eps=1e-6, comm=None, residual=True):
    """The second half of transformer_block (llama3/LLaMA-jax.ipynb:886-890):
    ``out = h + feed_forward(params['ffn'], rms_norm(h, params['ffn_norm']))`` - one C-ABI call forward, one backward,
    the residual add in the w2 GEMM's epilogue (residual=False returns the ffn output alone)."""
    from .ffn import FfnConfig, FfnFn
    C_ = h.shape[-1]
    f = block_params["ffn"]
    cfg = FfnConfig(N=h.numel() // C_, C=C_, H=f["w1"].shape[1], dtype=h.dtype, norm_kind=L.NORM_RMS, norm_eps=eps, residual=residual)
    y = FfnFn.apply(cfg, ("w1", "w3", "w2", "norm_w"), comm, h.reshape(-1, C_), f["w1"], f["w3"], f["w2"], block_params["ffn_norm"])
    return y.reshape(h.shape)


def transformer_block(params, x, mask, freqs_cis, n_heads, n_kv_heads, cache=None, position=0, training=False, dropout_rate=0.0,
                      key=None, comm=None):
    """llama3/LLaMA-jax.ipynb:880-891, same signature.  Two C-ABI calls forward (attention block with ``h = x + attn`` in its
    output projection's epilogue, feed-forward sub-block with ``out = h + ffn`` in the w2 GEMM's epilogue) and two backward.
    The notebook always calls it with ``training=False`` (:927; ``dropout_rate=0.0`` :358): the dropout branch has no kernel."""
    if training and dropout_rate != 0.0:
        raise NotImplementedError("dropout inside transformer_block is dead code in the notebook (training=False, llama3:927) and has no kernel")
    if cache is not None:
        # decode step: h = x + attention(rms_norm(x), cache, position) in one spb_block_decode call, then the feed-forward half
        h, new_cache = _decode(params["attention"], x, mask, freqs_cis, n_heads, n_kv_heads, cache, position, norm_w=params["attention_norm"],
                               residual=True)
        with torch.no_grad():
            return ffn_block(params, h, comm=comm, residual=True), new_cache
    h, new_cache = attention_block(params, x, mask, freqs_cis, n_heads, n_kv_heads, comm=comm, residual=True)
    out = ffn_block(params, h, comm=comm, residual=True)
    return out, new_cache


class _NllFn(torch.autograd.Function):
    """-mean(log_softmax(logits)[targets]) over logits [N, V] (llama3/LLaMA-jax.ipynb:959-967) through spb_nll_fwd / spb_nll_bwd."""

    @staticmethod
    def forward(ctx, logits, targets):
        import ctypes as C
        lib = L.load()
        N, V = logits.shape
        if logits.stride(1) != 1:
            logits = logits.contiguous()
        tg = targets.to(torch.int64).contiguous()
        f32 = dict(dtype=torch.float32, device=logits.device)
        lse, row_nll, loss = torch.empty(N, **f32), torch.empty(N, **f32), torch.empty((), **f32)
        dt = L.F32 if logits.dtype == torch.float32 else L.BF16
        st = C.c_void_p(torch.cuda.current_stream(logits.device).cuda_stream)
        p = lambda t: C.c_void_p(t.data_ptr())
        L.check(lib.spb_nll_fwd(dt, p(logits), logits.stride(0), p(tg), N, V, p(lse), p(row_nll), p(loss), st), "spb_nll_fwd")
        ctx.save_for_backward(logits, tg, lse)
        return loss

    @staticmethod
    def backward(ctx, dloss):
        import ctypes as C
        logits, tg, lse = ctx.saved_tensors
        lib = L.load()
        N, V = logits.shape
        dlogits = torch.empty_like(logits)
        dt = L.F32 if logits.dtype == torch.float32 else L.BF16
        st = C.c_void_p(torch.cuda.current_stream(logits.device).cuda_stream)
        p = lambda t: C.c_void_p(t.data_ptr())
        L.check(lib.spb_nll_bwd(dt, p(logits), logits.stride(0), p(tg), N, V, p(lse), float(dloss), p(dlogits), dlogits.stride(0), st),
                "spb_nll_bwd")
        return dlogits, None


def nll_loss(logits, targets):
    """llama3/LLaMA-jax.ipynb:959-967: ``-mean(take_along_axis(log_softmax(logits), targets[:, None], axis=1))``;
    logits [..., V] are flattened to [N, V] like the notebook's reshape (:959-960)."""
    V = logits.shape[-1]
    return _NllFn.apply(logits.reshape(-1, V), targets.reshape(-1))


class _LossHeadFn(torch.autograd.Function):
    """loss = nll(h @ w, targets) (llama3/LLaMA-jax.ipynb:930 + :959-967) through spb_loss_head_fwd / spb_loss_head_bwd: the [N, V]
    logits are never written (forward: online max / sum-exp / target logit in the GEMM epilogue; backward: dlogits recomputed tile by
    tile straight into the two gradient GEMMs' operand).  bf16; the vocabulary need not be a multiple of 8."""

    @staticmethod
    def forward(ctx, h, w, targets):
        import ctypes as C
        lib = L.load()
        N, Cm = h.shape
        V = w.shape[1]
        if h.dtype != torch.bfloat16 or w.dtype != torch.bfloat16:
            raise L.SpbError("loss_head: the fused head is bf16 (fp32: ops.matmul + nll_loss)")
        h, w = h.contiguous(), w.contiguous()
        tg = targets.to(torch.int64).contiguous()
        f32 = dict(dtype=torch.float32, device=h.device)
        lse, row_nll, loss = torch.empty(N, **f32), torch.empty(N, **f32), torch.empty((), **f32)
        ws = torch.empty(lib.spb_loss_head_ws_bytes(N, Cm, V), dtype=torch.uint8, device=h.device)
        st = C.c_void_p(torch.cuda.current_stream(h.device).cuda_stream)
        p = lambda t: C.c_void_p(t.data_ptr())
        L.check(lib.spb_loss_head_fwd(L.BF16, p(h), h.stride(0), p(w), w.stride(0), p(tg), N, Cm, V, p(lse), p(row_nll), p(loss), p(ws), st),
                "spb_loss_head_fwd")
        ctx.save_for_backward(h, w, tg, lse)
        ctx.ws = ws
        return loss

    @staticmethod
    def backward(ctx, dloss):
        import ctypes as C
        h, w, tg, lse = ctx.saved_tensors
        lib = L.load()
        N, Cm = h.shape
        V = w.shape[1]
        dh = torch.empty_like(h)
        dw = torch.empty((Cm, V), dtype=torch.float32, device=h.device)
        st = C.c_void_p(torch.cuda.current_stream(h.device).cuda_stream)
        p = lambda t: C.c_void_p(t.data_ptr())
        L.check(lib.spb_loss_head_bwd(L.BF16, p(h), h.stride(0), p(w), w.stride(0), p(tg), N, Cm, V, p(lse), float(dloss), p(dh), dh.stride(0),
                                      p(dw), dw.stride(0), p(ctx.ws), st), "spb_loss_head_bwd")
        ctx.ws = None
        return dh, dw.to(w.dtype), None


def loss_head(norm_f, output_w, h, targets):
    """The end of model_forward + compute_loss (llama3/LLaMA-jax.ipynb:929-930, 959-967):
    ``logits = dot(rms_norm(h, norm_f), output)``; ``loss = -mean(log_softmax(logits)[targets])`` - bf16: fused (no logits in memory);
    fp32: the rms_norm / matmul / nll_loss kernels one after the other."""
    Cm = h.shape[-1]
    hn = rms_norm(h, norm_f).reshape(-1, Cm)
    if h.dtype == torch.bfloat16:
        return _LossHeadFn.apply(hn, output_w, targets.reshape(-1))
    from .ops import matmul
    return nll_loss(matmul(hn, output_w), targets)
