"""Per-stage entry points of the C ABI on torch CUDA tensors (used by the per-kernel parity tests, ncu runs and
bench.py's kernel-level roofline).  Thin marshalling only - every function is one call into libspb200.so."""
import ctypes as C

import torch

from . import _lib as L
from .block import _dt, _ptr, _stream


def _rowmajor(t, name):
    if t.dim() != 2 or t.stride(1) != 1:
        raise L.SpbError(f"{name} must be a 2-D row-major tensor (possibly a column slice)")
    return t


def gemm(A, B, transA=False, transB=False, bias=None, out=None, out_dtype=None, accumulate=False, impl=L.IMPL_AUTO):
    """D[M,N] = opA(A) @ opB(B) (+bias) (+D).  A stored [M,K] (or [K,M] if transA); B stored [K,N] (or [N,K] if transB)."""
    lib = L.load()
    _rowmajor(A, "A"), _rowmajor(B, "B")
    M, K = (A.shape[1], A.shape[0]) if transA else (A.shape[0], A.shape[1])
    N = B.shape[0] if transB else B.shape[1]
    kb = B.shape[1] if transB else B.shape[0]
    if kb != K:
        raise L.SpbError(f"gemm: contraction mismatch {K} vs {kb}")
    out_dtype = out_dtype or (out.dtype if out is not None else A.dtype)
    if out is None:
        if accumulate:
            raise L.SpbError("accumulate needs out")
        out = torch.empty((M, N), dtype=out_dtype, device=A.device)
    _rowmajor(out, "out")
    L.check(lib.spb_gemm(_dt(A.dtype), _dt(out.dtype), impl, M, N, K, _ptr(A), A.stride(0), int(transA), _ptr(B), B.stride(0),
                         int(transB), _ptr(out), out.stride(0), _ptr(bias), int(accumulate), _stream()), "spb_gemm")
    return out


def norm_fwd(x, w, b=None, kind=L.NORM_RMS, eps=1e-6, upcast=False):
    lib = L.load()
    Cn = x.shape[-1]
    N = x.numel() // Cn
    y = torch.empty_like(x)
    stats = torch.empty((N, 2), dtype=torch.float32, device=x.device)
    L.check(lib.spb_norm_fwd(_dt(x.dtype), kind, int(upcast), eps, _ptr(x), _ptr(w), _ptr(b), _ptr(y), _ptr(stats), N, Cn,
                             _stream()), "spb_norm_fwd")
    return y, stats


def norm_bwd(x, w, stats, dy, kind=L.NORM_RMS, upcast=False, want_db=False):
    lib = L.load()
    Cn = x.shape[-1]
    N = x.numel() // Cn
    dx = torch.empty_like(x)
    dw = torch.empty(Cn, dtype=torch.float32, device=x.device)
    db = torch.empty(Cn, dtype=torch.float32, device=x.device) if want_db else None
    partial = torch.empty(lib.spb_norm_bwd_partial_bytes(N, Cn), dtype=torch.uint8, device=x.device)
    L.check(lib.spb_norm_bwd(_dt(x.dtype), kind, int(upcast), _ptr(x), _ptr(w), _ptr(stats), _ptr(dy), _ptr(dx), _ptr(dw),
                             _ptr(db), _ptr(partial), N, Cn, _stream()), "spb_norm_bwd")
    return dx, dw, db


def rope_apply_(x2d, T, heads, hd, table, transpose=False):
    """In place on x2d [N, >= heads*hd] (row stride = x2d.stride(0))."""
    lib = L.load()
    _rowmajor(x2d, "x")
    L.check(lib.spb_rope_apply(_dt(x2d.dtype), _ptr(x2d), x2d.stride(0), x2d.shape[0], T, heads, hd, _ptr(table),
                               int(transpose), _stream()), "spb_rope_apply")
    return x2d


def colsum(x2d):
    lib = L.load()
    _rowmajor(x2d, "x")
    N, cols = x2d.shape
    out = torch.empty(cols, dtype=torch.float32, device=x2d.device)
    partial = torch.empty(lib.spb_norm_bwd_partial_bytes(N, cols), dtype=torch.uint8, device=x2d.device)
    L.check(lib.spb_colsum(_dt(x2d.dtype), _ptr(x2d), x2d.stride(0), N, cols, _ptr(out), _ptr(partial), _stream()), "spb_colsum")
    return out


def flash_fwd(q, k, v, causal, scale, impl=L.IMPL_AUTO):
    """q [B,T,Hq,hd], k/v [B,T,Hkv,hd] (last two dims contiguous, token stride free) -> o [B,T,Hq,hd], lse [B,Hq,T]."""
    lib = L.load()
    B, T, Hq, hd = q.shape
    Hkv = k.shape[2]
    o = torch.empty((B, T, Hq, hd), dtype=q.dtype, device=q.device)
    lse = torch.empty((B, Hq, T), dtype=torch.float32, device=q.device)
    L.check(lib.spb_flash_fwd(_dt(q.dtype), impl, B, T, Hq, Hkv, hd, int(causal), scale, _ptr(q), q.stride(1), _ptr(k),
                              k.stride(1), _ptr(v), v.stride(1), _ptr(o), o.stride(1), _ptr(lse), _stream()), "spb_flash_fwd")
    return o, lse


def flash_bwd(q, k, v, o, do, lse, causal, scale, impl=L.IMPL_AUTO):
    lib = L.load()
    B, T, Hq, hd = q.shape
    Hkv = k.shape[2]
    dq = torch.empty((B, T, Hq, hd), dtype=q.dtype, device=q.device)
    dk = torch.empty((B, T, Hkv, hd), dtype=q.dtype, device=q.device)
    dv = torch.empty((B, T, Hkv, hd), dtype=q.dtype, device=q.device)
    delta = torch.empty((B, Hq, T), dtype=torch.float32, device=q.device)
    dq_acc = torch.empty((B, T, Hq, hd), dtype=torch.float32, device=q.device)
    L.check(lib.spb_flash_bwd(_dt(q.dtype), impl, B, T, Hq, Hkv, hd, int(causal), scale, _ptr(q), q.stride(1), _ptr(k),
                              k.stride(1), _ptr(v), v.stride(1), _ptr(o), o.stride(1), _ptr(do), do.stride(1), _ptr(lse),
                              _ptr(delta), _ptr(dq), dq.stride(1), _ptr(dk), dk.stride(1), _ptr(dv), dv.stride(1),
                              _ptr(dq_acc), _stream()), "spb_flash_bwd")
    return dq, dk, dv


def attn_wide_fwd(q, k, v, causal, scale):
    """Wide-head attention (bf16, hd > 256): q [B,T,Hq,hd], k/v [B,T,Hkv,hd] -> o, lse through spb_attn_wide_fwd."""
    lib = L.load()
    B, T, Hq, hd = q.shape
    Hkv = k.shape[2]
    o = torch.empty((B, T, Hq, hd), dtype=q.dtype, device=q.device)
    lse = torch.empty((B, Hq, T), dtype=torch.float32, device=q.device)
    scratch = torch.empty(lib.spb_attn_wide_scratch_bytes(B, T, Hq, Hkv, hd), dtype=torch.uint8, device=q.device)
    L.check(lib.spb_attn_wide_fwd(B, T, Hq, Hkv, hd, int(causal), scale, _ptr(q), q.stride(1), _ptr(k), k.stride(1), _ptr(v), v.stride(1),
                                  _ptr(o), o.stride(1), _ptr(lse), _ptr(scratch), _stream()), "spb_attn_wide_fwd")
    return o, lse


def attn_wide_bwd(q, k, v, o, do, lse, causal, scale):
    lib = L.load()
    B, T, Hq, hd = q.shape
    Hkv = k.shape[2]
    dq = torch.empty((B, T, Hq, hd), dtype=q.dtype, device=q.device)
    dk = torch.empty((B, T, Hkv, hd), dtype=q.dtype, device=q.device)
    dv = torch.empty((B, T, Hkv, hd), dtype=q.dtype, device=q.device)
    delta = torch.empty((B, Hq, T), dtype=torch.float32, device=q.device)
    scratch = torch.empty(lib.spb_attn_wide_scratch_bytes(B, T, Hq, Hkv, hd), dtype=torch.uint8, device=q.device)
    L.check(lib.spb_attn_wide_bwd(B, T, Hq, Hkv, hd, int(causal), scale, _ptr(q), q.stride(1), _ptr(k), k.stride(1), _ptr(v), v.stride(1),
                                  _ptr(o), o.stride(1), _ptr(do), do.stride(1), _ptr(lse), _ptr(delta), _ptr(dq), dq.stride(1), _ptr(dk),
                                  dk.stride(1), _ptr(dv), dv.stride(1), _ptr(scratch), _stream()), "spb_attn_wide_bwd")
    return dq, dk, dv


def attn_decode(q, k_cache, v_cache, Tk, scale, want_lse=False):
    """llama3:815-829 with a cache: q [B,Tq,Hq,hd] (rotated) against the first Tk rows of k_cache / v_cache [B,Tmax,Hkv,hd]
    (any token / batch strides, last dim contiguous) -> o [B,Tq,Hq,hd] (and lse [B,Hq,Tq])."""
    lib = L.load()
    B, Tq, Hq, hd = q.shape
    Hkv = k_cache.shape[2]
    if q.stride(3) != 1 or q.stride(2) != hd or k_cache.stride(3) != 1 or k_cache.stride(2) != hd or v_cache.stride(3) != 1 or v_cache.stride(2) != hd:
        raise L.SpbError("attn_decode: heads must be contiguous [.., H, hd]")
    if q.stride(0) != Tq * q.stride(1):
        raise L.SpbError("attn_decode: q rows must have one stride")
    o = torch.empty((B, Tq, Hq, hd), dtype=q.dtype, device=q.device)
    lse = torch.empty((B, Hq, Tq), dtype=torch.float32, device=q.device) if want_lse else None
    scratch = torch.empty(lib.spb_attn_decode_scratch_bytes(B, Tq, Hq, Hkv, hd, Tk), dtype=torch.uint8, device=q.device)
    L.check(lib.spb_attn_decode(_dt(q.dtype), B, Tq, Hq, Hkv, hd, Tk, scale, _ptr(q), q.stride(1), _ptr(k_cache), k_cache.stride(1),
                                k_cache.stride(0), _ptr(v_cache), v_cache.stride(1), v_cache.stride(0), _ptr(o), o.stride(1), _ptr(lse),
                                _ptr(scratch), _stream()), "spb_attn_decode")
    return (o, lse) if want_lse else o


class _MatmulFn(torch.autograd.Function):
    """y = x @ w with w stored [in, out] (``jnp.dot(h, params['output'])``, llama3:930): forward and both backward GEMMs through
    spb_gemm (tcgen05 for bf16, FFMA for fp32); the weight gradient is accumulated in fp32."""

    @staticmethod
    def forward(ctx, x, w):
        ctx.save_for_backward(x, w)
        return gemm(x, w)

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        dy = dy.contiguous()
        dx = gemm(dy, w, transB=True)
        dw = gemm(x, dy, transA=True, out_dtype=torch.float32)
        return dx, dw.to(w.dtype)


def matmul(x, w):
    """Differentiable 2-D matmul on the library's GEMMs: x [N, K] @ w [K, M]."""
    return _MatmulFn.apply(x.contiguous(), w.contiguous())
