"""torch.autograd glue: makes the C-ABI block differentiable the way ``jax.value_and_grad`` makes the notebook's
(llama3/LLaMA-jax.ipynb:995).  torch is only the tape; forward and backward both run in libspb200.so."""
import ctypes as C

import torch

from . import _lib as L
from .block import AttentionBlock, BlockConfig, _ptr, _stream

_BLOCKS = {}


def get_block(cfg: BlockConfig, device, comm=None) -> AttentionBlock:
    key = (tuple(sorted(cfg.__dict__.items(), key=lambda kv: kv[0])), str(device), comm)
    blk = _BLOCKS.get(key)
    if blk is None:
        blk = AttentionBlock(cfg, device=device, comm=comm)
        _BLOCKS[key] = blk
    return blk


class BlockFn(torch.autograd.Function):
    """y = block(x; weights).  `names` lists the keys of the tensors that follow, in order."""

    @staticmethod
    def forward(ctx, cfg, names, comm, x, *tensors):
        params = dict(zip(names, tensors))
        blk = get_block(cfg, x.device, comm)
        y = blk.forward(x.contiguous(), params)
        ctx.blk, ctx.names, ctx.params = blk, names, params
        ctx.save_for_backward(x, *tensors)
        # residuals live in the block's workspace: a second forward with the same config before this backward
        # would overwrite them, so remember which forward owns the workspace.
        blk._owner = ctx
        k, v = blk.kv_cache()
        ctx.mark_non_differentiable(k, v)
        return y, k, v

    @staticmethod
    def backward(ctx, dy, _dk, _dv):
        blk = ctx.blk
        if getattr(blk, "_owner", None) is not ctx:
            raise L.SpbError("the block workspace was overwritten by a later forward with the same shape; "
                             "run backward before the next forward of an identically-shaped block")
        g = blk.backward(dy.contiguous())
        out = []
        for n in ctx.names:
            p = ctx.params[n]
            gn = {"wqkv": "dwqkv", "wq": "dwq", "wk": "dwk", "wv": "dwv", "wo": "dwo", "bqkv": "dbqkv", "bo": "dbo",
                  "norm_w": "dnorm_w", "norm_b": "dnorm_b"}.get(n)
            out.append(g[gn].to(p.dtype) if gn in g and gn is not None else None)
        return (None, None, None, g["dx"], *out)


class NormFn(torch.autograd.Function):
    """rms_norm / LayerNorm alone (llama3/LLaMA-jax.ipynb:536-538; flax / torch LayerNorm)."""

    @staticmethod
    def forward(ctx, kind, eps, upcast, x, w, b):
        lib = L.load()
        C_ = x.shape[-1]
        xc = x.contiguous()
        N = xc.numel() // C_
        y = torch.empty_like(xc)
        stats = torch.empty((N, 2), dtype=torch.float32, device=x.device)
        dt = L.F32 if x.dtype == torch.float32 else L.BF16
        L.check(lib.spb_norm_fwd(dt, kind, int(upcast), eps, _ptr(xc), _ptr(w), _ptr(b), _ptr(y), _ptr(stats), N, C_,
                                 _stream()), "spb_norm_fwd")
        ctx.save_for_backward(xc, w, stats)
        ctx.kind, ctx.upcast, ctx.has_b = kind, upcast, b is not None
        return y

    @staticmethod
    def backward(ctx, dy):
        lib = L.load()
        xc, w, stats = ctx.saved_tensors
        C_ = xc.shape[-1]
        N = xc.numel() // C_
        dyc = dy.contiguous()
        dx = torch.empty_like(xc)
        dw = torch.empty(C_, dtype=torch.float32, device=xc.device)
        db = torch.empty(C_, dtype=torch.float32, device=xc.device) if ctx.has_b else None
        partial = torch.empty(lib.spb_norm_bwd_partial_bytes(N, C_), dtype=torch.uint8, device=xc.device)
        dt = L.F32 if xc.dtype == torch.float32 else L.BF16
        L.check(lib.spb_norm_bwd(dt, ctx.kind, int(ctx.upcast), _ptr(xc), _ptr(w), _ptr(stats), _ptr(dyc), _ptr(dx),
                                 _ptr(dw), _ptr(db), _ptr(partial), N, C_, _stream()), "spb_norm_bwd")
        return None, None, None, dx, dw.to(w.dtype), (db.to(w.dtype) if db is not None else None)


class RopeFn(torch.autograd.Function):
    """apply_rotary_emb on one tensor [B,T,H,hd] (llama3/LLaMA-jax.ipynb:592-601)."""

    @staticmethod
    def forward(ctx, x, table):
        lib = L.load()
        B, T, H, hd = x.shape
        y = x.contiguous().clone()
        dt = L.F32 if x.dtype == torch.float32 else L.BF16
        L.check(lib.spb_rope_apply(dt, _ptr(y), H * hd, B * T, T, H, hd, _ptr(table), 0, _stream()), "spb_rope_apply")
        ctx.save_for_backward(table)
        return y

    @staticmethod
    def backward(ctx, dy):
        lib = L.load()
        (table,) = ctx.saved_tensors
        B, T, H, hd = dy.shape
        dx = dy.contiguous().clone()
        dt = L.F32 if dy.dtype == torch.float32 else L.BF16
        L.check(lib.spb_rope_apply(dt, _ptr(dx), H * hd, B * T, T, H, hd, _ptr(table), 1, _stream()), "spb_rope_apply")
        return dx, None
