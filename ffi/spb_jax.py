"""jax.ffi face of libspb200.so: drop-in replacements for the hot-path cells of llama3/LLaMA-jax.ipynb with the
notebook's own names, argument order and defaults (c15 rms_norm :536, c17 apply_rotary_emb :592, c18 repeat_kv :626,
c24 attention :809), traceable under jax.jit and differentiable under jax.value_and_grad (:993-995) through
jax.custom_vjp.  transformer_block / model_forward / compute_loss / update_step of the notebook stay as they are.

Needs JAX with jax.ffi (>= 0.4.35) and a CUDA backend, and ffi/spb_jax_ffi.cc compiled into libspb_jax_ffi.so (see the
header of that file).  Neither exists in the image this repository is developed in (no jax wheel, no network), so this
module is shipped as source: tests/test_ffi_source.py checks it against include/spb200.h without importing jax.

    import spb_jax
    spb_jax.register()                       # once per process
    from spb_jax import rms_norm, apply_rotary_emb, repeat_kv, attention      # replaces cells c15, c17, c18, c24

Contract carried over from tracing: the mask is a tracer whose values cannot be read, so ``mask is not None``
means causal and ``None`` bidirectional.  ``cache is not None`` or ``position != 0`` raise in ``attention`` (no slow path);
the KV-cache step of the same function (llama3:815-829, caller ``generate`` :499-511) is ``attention_decode``: fixed-capacity
caches appended in place, the number of valid rows a device scalar - one trace under jit serves every token.
"""
import ctypes
import os
from functools import partial

import numpy as np

import jax
import jax.numpy as jnp

_HERE = os.path.dirname(os.path.abspath(__file__))
_CORE = os.environ.get("SPB_LIB_PATH") or os.path.join(_HERE, "..", "solvingpapers_b200", "libspb200.so")
_SHIM = os.environ.get("SPB_JAX_FFI_PATH") or os.path.join(_HERE, "libspb_jax_ffi.so")
_core = None
_registered = False
_comm = 0  # communicator handle (int) for data-parallel runs; 0 = single device


class _BlockDesc(ctypes.Structure):  # spb_block_desc (include/spb200.h)
    _fields_ = [(n, ctypes.c_int32) for n in ("B", "T", "C", "Hq", "Hkv", "hd", "dtype", "norm_kind")] + [("norm_eps", ctypes.c_float)] + \
               [(n, ctypes.c_int32) for n in ("norm_upcast", "rope_kind", "causal")] + [("scale", ctypes.c_float)] + \
               [(n, ctypes.c_int32) for n in ("weight_layout", "qkv_bias", "out_bias", "attn_impl", "gemm_impl", "residual")]


class _WsLayout(ctypes.Structure):  # spb_ws_layout
    _fields_ = [(n, ctypes.c_size_t) for n in ("total", "xhat", "stats", "qkv", "attn_o", "lse", "delta", "dattn_o", "dqkv", "dq_acc",
                                                "dxhat", "partial", "partial_bytes")] + [("ld_qkv", ctypes.c_int64)] + \
               [(n, ctypes.c_size_t) for n in ("attn_scratch", "attn_scratch_bytes", "gemm_scratch", "gemm_scratch_bytes")]


class _DecodeWs(ctypes.Structure):  # spb_decode_ws
    _fields_ = [(n, ctypes.c_size_t) for n in ("total", "xhat", "qkv", "attn_o", "gemm_part", "attn_part")]


def register():
    """Load both libraries and register the FFI targets (idempotent)."""
    global _core, _registered
    if _registered:
        return
    _core = ctypes.CDLL(_CORE, mode=ctypes.RTLD_GLOBAL)  # the shim links against it
    _core.spb_block_ws_layout.restype = ctypes.c_int
    _core.spb_block_ws_layout.argtypes = [ctypes.POINTER(_BlockDesc), ctypes.POINTER(_WsLayout)]
    _core.spb_block_decode_ws_layout.restype = ctypes.c_int
    _core.spb_block_decode_ws_layout.argtypes = [ctypes.POINTER(_BlockDesc), ctypes.c_int32, ctypes.POINTER(_DecodeWs)]
    _core.spb_norm_bwd_partial_bytes.restype = ctypes.c_size_t
    _core.spb_norm_bwd_partial_bytes.argtypes = [ctypes.c_int64, ctypes.c_int32]
    shim = ctypes.CDLL(_SHIM)
    for target, sym in (("spb_block_fwd", "SpbBlockFwd"), ("spb_block_bwd", "SpbBlockBwd"), ("spb_block_decode", "SpbBlockDecode"),
                        ("spb_rms_norm", "SpbRmsNorm"), ("spb_rms_norm_bwd", "SpbRmsNormBwd"), ("spb_rope", "SpbRope")):
        jax.ffi.register_ffi_target(target, jax.ffi.pycapsule(getattr(shim, sym)), platform="CUDA")
    _registered = True


def set_communicator(handle: int):
    """Data parallelism: pass the value of the handle spb_comm_init_rank returned on this process; from then on the
    backward custom call sums the weight gradients over ranks with ncclAllReduce.  0 switches it off."""
    global _comm
    _comm = int(handle)


def _workspace_bytes(B, T, C, n_heads, n_kv_heads, dtype, norm, residual):
    """Host-side, at trace time: spb_block_ws_layout(desc).total."""
    hd = C // n_heads
    d = _BlockDesc(B=B, T=T, C=C, Hq=n_heads, Hkv=n_kv_heads, hd=hd, dtype=1 if dtype == jnp.bfloat16 else 0, norm_kind=1 if norm else 0,
                   norm_eps=1e-6, norm_upcast=0, rope_kind=1, causal=1, scale=1.0 / float(np.sqrt(hd)), weight_layout=0, qkv_bias=0,
                   out_bias=0, attn_impl=0, gemm_impl=0, residual=int(residual))
    lay = _WsLayout()
    rc = _core.spb_block_ws_layout(ctypes.byref(d), ctypes.byref(lay))
    if rc != 0:
        raise ValueError(f"spb_block_ws_layout failed with code {rc}")
    return int(lay.total)


def _decode_workspace_bytes(B, Tq, C, n_heads, n_kv_heads, dtype, norm, residual, Tmax):
    """Host-side, at trace time: spb_block_decode_ws_layout(desc, Tmax).total."""
    hd = C // n_heads
    d = _BlockDesc(B=B, T=Tq, C=C, Hq=n_heads, Hkv=n_kv_heads, hd=hd, dtype=1 if dtype == jnp.bfloat16 else 0, norm_kind=1 if norm else 0,
                   norm_eps=1e-6, norm_upcast=0, rope_kind=1, causal=0, scale=1.0 / float(np.sqrt(hd)), weight_layout=0, qkv_bias=0,
                   out_bias=0, attn_impl=0, gemm_impl=0, residual=int(residual))
    lay = _DecodeWs()
    rc = _core.spb_block_decode_ws_layout(ctypes.byref(d), int(Tmax), ctypes.byref(lay))
    if rc != 0:
        raise ValueError(f"spb_block_decode_ws_layout failed with code {rc}")
    return int(lay.total)


def _rope_table(freqs_cis):
    """complex64 [T, hd/2] (llama3:563-567) -> the library's fp32 [T, hd/2, 4] table (cos, -sin, sin, cos)."""
    c, s = jnp.real(freqs_cis).astype(jnp.float32), jnp.imag(freqs_cis).astype(jnp.float32)
    return jnp.stack([c, -s, s, c], axis=-1)


# ---------------------------------------------------------------------------------------------------- the fused block
@partial(jax.custom_vjp, nondiff_argnums=(7, 8, 9, 10, 11, 12))
def _block(x, wq, wk, wv, wo, norm_w, rope, n_heads, n_kv_heads, eps, causal, norm, residual):
    (y, k, v), _ = _block_fwd(x, wq, wk, wv, wo, norm_w, rope, n_heads, n_kv_heads, eps, causal, norm, residual)
    return y, k, v


def _block_fwd(x, wq, wk, wv, wo, norm_w, rope, n_heads, n_kv_heads, eps, causal, norm, residual):
    B, T, C = x.shape
    hd = C // n_heads
    ws_bytes = _workspace_bytes(B, T, C, n_heads, n_kv_heads, x.dtype, norm, residual)
    y, k, v, ws = jax.ffi.ffi_call(
        "spb_block_fwd",
        (jax.ShapeDtypeStruct(x.shape, x.dtype), jax.ShapeDtypeStruct((B, T, n_kv_heads, hd), x.dtype),
         jax.ShapeDtypeStruct((B, T, n_kv_heads, hd), x.dtype), jax.ShapeDtypeStruct((ws_bytes,), jnp.uint8)),
    )(x, wq, wk, wv, wo, norm_w, rope, n_heads=np.int32(n_heads), n_kv_heads=np.int32(n_kv_heads), eps=np.float32(eps), causal=bool(causal),
      norm=bool(norm), residual=bool(residual))
    return (y, k, v), (x, wq, wk, wv, wo, norm_w, rope, ws)


def _block_bwd(n_heads, n_kv_heads, eps, causal, norm, residual, res, cts):
    x, wq, wk, wv, wo, norm_w, rope, ws = res
    dy, _dk, _dv = cts  # (k, v) are the new_cache: the notebook's loss never differentiates through them (llama3:927 discards it)
    f32 = jnp.float32
    dx, dwq, dwk, dwv, dwo, dnorm_w, _ = jax.ffi.ffi_call(
        "spb_block_bwd",
        (jax.ShapeDtypeStruct(x.shape, x.dtype), jax.ShapeDtypeStruct(wq.shape, f32), jax.ShapeDtypeStruct(wk.shape, f32),
         jax.ShapeDtypeStruct(wv.shape, f32), jax.ShapeDtypeStruct(wo.shape, f32), jax.ShapeDtypeStruct(norm_w.shape, f32),
         jax.ShapeDtypeStruct(ws.shape, jnp.uint8)),
        input_output_aliases={8: 6},  # the workspace is consumed: the backward writes its scratch into it
    )(x, dy, wq, wk, wv, wo, norm_w, rope, ws, n_heads=np.int32(n_heads), n_kv_heads=np.int32(n_kv_heads), eps=np.float32(eps),
      causal=bool(causal), norm=bool(norm), residual=bool(residual), comm=np.int64(_comm))
    cast = lambda g, w: g.astype(w.dtype)
    return dx, cast(dwq, wq), cast(dwk, wk), cast(dwv, wv), cast(dwo, wo), cast(dnorm_w, norm_w), jnp.zeros_like(rope)


_block.defvjp(_block_fwd, _block_bwd)


# ---------------------------------------------------------------------------------------------------- notebook signatures
@partial(jax.custom_vjp, nondiff_argnums=(2,))
def rms_norm(x, weight, eps=1e-6):
    """llama3/LLaMA-jax.ipynb:536-538."""
    return _rms_fwd(x, weight, eps)[0]


def _rms_fwd(x, weight, eps):
    n = int(np.prod(x.shape[:-1]))
    y, stats = jax.ffi.ffi_call("spb_rms_norm", (jax.ShapeDtypeStruct(x.shape, x.dtype), jax.ShapeDtypeStruct((n, 2), jnp.float32)))(
        x, weight, eps=np.float32(eps))
    return y, (x, weight, stats)


def _rms_bwd(eps, res, dy):
    x, weight, stats = res
    n, c = int(np.prod(x.shape[:-1])), int(x.shape[-1])
    scratch = int(_core.spb_norm_bwd_partial_bytes(n, c))
    dx, dw, _ = jax.ffi.ffi_call("spb_rms_norm_bwd", (jax.ShapeDtypeStruct(x.shape, x.dtype), jax.ShapeDtypeStruct(weight.shape, jnp.float32),
                                                      jax.ShapeDtypeStruct((max(scratch, 1),), jnp.uint8)))(x, weight, stats, dy)
    return dx, dw.astype(weight.dtype)


rms_norm.defvjp(_rms_fwd, _rms_bwd)


@jax.custom_vjp
def _rope_one(x, table):
    return jax.ffi.ffi_call("spb_rope", jax.ShapeDtypeStruct(x.shape, x.dtype))(x, table, transpose=False)


def _rope_one_fwd(x, table):
    return _rope_one(x, table), table


def _rope_one_bwd(table, dy):
    return jax.ffi.ffi_call("spb_rope", jax.ShapeDtypeStruct(dy.shape, dy.dtype))(dy, table, transpose=True), jnp.zeros_like(table)


_rope_one.defvjp(_rope_one_fwd, _rope_one_bwd)


def apply_rotary_emb(xq, xk, freqs_cis):
    """llama3/LLaMA-jax.ipynb:592-601; xq [B,T,Hq,hd], xk [B,T,Hkv,hd], freqs_cis [T, hd/2] complex64."""
    table = _rope_table(freqs_cis)
    return _rope_one(xq, table), _rope_one(xk, table)


def repeat_kv(x, n_rep):
    """llama3/LLaMA-jax.ipynb:626-627.  API parity only: attention() never materialises repeated K/V."""
    return x if n_rep == 1 else jnp.repeat(x, n_rep, axis=2)


def attention(params, x, mask, freqs_cis, n_heads, n_kv_heads, cache=None, position=0):
    """llama3/LLaMA-jax.ipynb:809-829: returns (out [B,T,C], (k, v)), k and v post-rotary and un-repeated."""
    if cache is not None or position != 0:
        raise NotImplementedError("attention() is the training path (cache=None, position=0); the KV-cache step is attention_decode()")
    T = x.shape[1]
    dummy_w = jnp.ones((x.shape[-1],), x.dtype)
    y, k, v = _block(x, params["wq"], params["wk"], params["wv"], params["wo"], dummy_w, _rope_table(freqs_cis[position:position + T]),
                     n_heads, n_kv_heads, 1e-6, mask is not None, False, False)
    return y, (k, v)


def attention_block(block_params, x, mask, freqs_cis, n_heads, n_kv_heads, eps=1e-6, residual=False):
    """The call of transformer_block fused into one custom call (llama3/LLaMA-jax.ipynb:881):
    ``attention(params['attention'], rms_norm(x, params['attention_norm']), mask, freqs_cis, n_heads, n_kv_heads)``;
    residual=True also folds the next line, ``h = x + attn_output`` (:885), into the output projection's epilogue."""
    a = block_params["attention"]
    T = x.shape[1]
    y, k, v = _block(x, a["wq"], a["wk"], a["wv"], a["wo"], block_params["attention_norm"], _rope_table(freqs_cis[:T]), n_heads, n_kv_heads,
                     eps, mask is not None, True, residual)
    return y, (k, v)


def attention_decode(block_params, x, k_cache, v_cache, cache_len, freqs_cis, n_heads, n_kv_heads, eps=1e-6, norm=True, residual=False):
    """The KV-cache branch of ``attention`` (llama3/LLaMA-jax.ipynb:815-829) as ``generate`` (:499-511) drives it, for one or a few
    new tokens ``x [B, Tq, C]``: ``attention(params['attention'], rms_norm(x, params['attention_norm']), None, freqs_cis, n_heads,
    n_kv_heads, cache=(k, v), position=cache_len)``.  The notebook grows ``(k, v)`` by ``jnp.concatenate`` every step (a new shape,
    hence a new trace, per token); here the caches have a fixed capacity ``[B, Tmax, n_kv_heads, head_dim]``, the new rows are
    appended in place and ``cache_len`` (int32[1], on the device) is advanced by the call itself, so one jitted step serves every
    token.  ``freqs_cis`` must cover ``Tmax`` positions.  Returns ``(out, (k_cache, v_cache, cache_len))``; not differentiable
    (the notebook only decodes under ``generate``).  ``norm=False`` takes ``x`` as already normalised, ``residual=True`` adds ``x``."""
    a = block_params["attention"] if "attention" in block_params else block_params
    B, Tq, C = x.shape
    Tmax = k_cache.shape[1]
    norm_w = block_params["attention_norm"] if norm else jnp.ones((C,), x.dtype)
    ws_bytes = _decode_workspace_bytes(B, Tq, C, n_heads, n_kv_heads, x.dtype, norm, residual, Tmax)
    y, k_cache, v_cache, cache_len, _ = jax.ffi.ffi_call(
        "spb_block_decode",
        (jax.ShapeDtypeStruct(x.shape, x.dtype), jax.ShapeDtypeStruct(k_cache.shape, k_cache.dtype),
         jax.ShapeDtypeStruct(v_cache.shape, v_cache.dtype), jax.ShapeDtypeStruct((1,), jnp.int32), jax.ShapeDtypeStruct((ws_bytes,), jnp.uint8)),
        input_output_aliases={7: 1, 8: 2, 9: 3},  # the caches and the length are updated in place
    )(x, a["wq"], a["wk"], a["wv"], a["wo"], norm_w, _rope_table(freqs_cis[:Tmax]), k_cache, v_cache, cache_len.astype(jnp.int32).reshape(1),
      n_heads=np.int32(n_heads), n_kv_heads=np.int32(n_kv_heads), eps=np.float32(eps), norm=bool(norm), residual=bool(residual))
    return y, (k_cache, v_cache, cache_len)
