"""torch.autograd glue: makes the C-ABI block differentiable the way ``jax.value_and_grad`` makes the notebook's
(llama3/LLaMA-jax.ipynb:995).  torch is only the tape; forward and backward both run in libspb200.so."""
import ctypes as C

import torch

from . import _lib as L
from .block import AttentionBlock, BlockConfig, _ptr, _stream

class WorkspacePool:
    """Free-list of workspace owners (AttentionBlock / FeedForward objects) keyed by (config, device, communicator).

    The reference trains through ``model_forward``, which calls ``transformer_block`` once per entry of ``params['blocks']``
    with identical shapes (llama3/LLaMA-jax.ipynb:925-927), so several forwards of one shape are alive at the same time and
    each needs its OWN residual storage until its backward has run.  Every forward leases an owner from the pool (creating
    one when the free-list is empty) and the lease returns it after the backward - or when the autograd node is dropped
    without a backward (inference).  A model with L same-shape layers therefore settles at L workspaces, reused every step."""

    def __init__(self, factory):
        self._factory = factory
        self._free = {}
        self.created = 0

    def lease(self, cfg, device, comm):
        key = (_cfg_key(cfg), str(device), comm)
        free = self._free.setdefault(key, [])
        if free:
            owner = free.pop()
        else:
            owner = self._factory(cfg, device, comm)
            self.created += 1
        return Lease(self, key, owner)

    def _give_back(self, key, owner):
        self._free.setdefault(key, []).append(owner)

    def clear(self):
        self._free.clear()


class Lease:
    """Exclusive use of one workspace owner between a forward and its backward."""

    def __init__(self, pool, key, owner):
        self.pool, self.key, self.owner = pool, key, owner

    def release(self):
        if self.owner is not None:
            self.pool._give_back(self.key, self.owner)
            self.owner = None

    def __del__(self):  # autograd node dropped without a backward
        try:
            self.release()
        except Exception:
            pass


def _cfg_key(cfg):
    return (type(cfg).__name__,) + tuple(sorted(cfg.__dict__.items(), key=lambda kv: kv[0]))


BLOCK_POOL = WorkspacePool(lambda cfg, device, comm: AttentionBlock(cfg, device=device, comm=comm))


def get_block(cfg: BlockConfig, device, comm=None) -> AttentionBlock:
    """A private AttentionBlock (own workspace) for callers that drive forward/backward themselves."""
    return AttentionBlock(cfg, device=device, comm=comm)


class BlockFn(torch.autograd.Function):
    """y = block(x; weights).  `names` lists the keys of the tensors that follow, in order."""

    @staticmethod
    def forward(ctx, cfg, names, comm, x, *tensors):
        params = dict(zip(names, tensors))
        lease = BLOCK_POOL.lease(cfg, x.device, comm)
        blk = lease.owner
        y = blk.forward(x.contiguous(), params)
        ctx.lease, ctx.names, ctx.params = lease, names, params
        ctx.save_for_backward(x, *tensors)
        # (k, v) of llama3:819 are returned as tensors that own their memory: the workspace they were computed in goes
        # back to the pool after the backward and is overwritten by a later forward
        k, v = (t.clone() for t in blk.kv_cache())
        ctx.mark_non_differentiable(k, v)
        return y, k, v

    @staticmethod
    def backward(ctx, dy, _dk, _dv):
        lease = ctx.lease
        if lease is None or lease.owner is None:
            raise L.SpbError("this block's residuals were already released by an earlier backward (retain_graph is not supported)")
        g = lease.owner.backward(dy.contiguous())
        lease.release()
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
