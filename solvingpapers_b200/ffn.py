"""Host mirror of the gated feed-forward sub-block over spb_ffn_fwd / spb_ffn_bwd.

llama3/LLaMA-jax.ipynb:854-855  feed_forward(params, x) = dot(silu(dot(x, w3)) * dot(x, w1), w2)
llama3/LLaMA-jax.ipynb:886,890  ffn_output = feed_forward(params['ffn'], rms_norm(h, params['ffn_norm'])); out = h + ffn_output
"""
import ctypes as C
from dataclasses import dataclass

import torch

from . import _lib as L


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _dt(dtype):
    if dtype == torch.float32:
        return L.F32
    if dtype == torch.bfloat16:
        return L.BF16
    raise L.SpbError(f"unsupported dtype {dtype} (float32 or bfloat16)")


@dataclass(frozen=True)
class FfnConfig:
    N: int                    # tokens (B*T)
    C: int                    # model width
    H: int                    # hidden width (4*C in the notebook, llama3:716-718)
    dtype: torch.dtype = torch.bfloat16
    norm_kind: int = L.NORM_RMS
    norm_eps: float = 1e-6
    norm_upcast: bool = False
    act: int = L.ACT_SWIGLU
    weight_layout: int = L.W_IN_OUT
    residual: bool = False    # out = h + ffn(norm(h)) inside the w2 GEMM's epilogue, dx += dy in the norm VJP
    gemm_impl: int = L.IMPL_AUTO

    def desc(self):
        return L.FfnDesc(self.N, self.C, self.H, _dt(self.dtype), self.norm_kind, self.norm_eps, int(self.norm_upcast), self.act,
                         self.weight_layout, int(self.residual), self.gemm_impl)

    @property
    def gated(self):
        return self.act != L.ACT_GELU_TANH

    def flops_fwd_bwd(self):
        n_mm = 3.0 if self.gated else 2.0  # contractions; each costs forward + two backward GEMMs
        return n_mm * 3.0 * 2.0 * self.N * self.C * self.H


class FeedForward:
    """One feed-forward sub-block bound to fixed shapes; weights are passed per call.
    params: {"w13": [C, 2H] packed (w1 | w3)}  or  {"w1": [C,H], "w3": [C,H]};  "w2": [H, C];  "norm_w" (and "norm_b");
    the plain MLP (ACT_GELU_TANH) takes "w1", "w2" and optional biases "b1" [H], "b2" [C]; W_OUT_IN transposes every weight."""

    def __init__(self, cfg: FfnConfig, device="cuda", comm=None):
        if not torch.cuda.is_available():
            raise L.SpbError("solvingpapers_b200 needs a CUDA device (B200, sm_100a); there is no CPU path")
        self.cfg = cfg
        self.lib = L.load()
        self.device = torch.device(device)
        self.d = cfg.desc()
        self.layout = L.FfnWs()
        L.check(self.lib.spb_ffn_ws_layout(C.byref(self.d), C.byref(self.layout)), "spb_ffn_ws_layout")
        self.ws = torch.empty(self.layout.total, dtype=torch.uint8, device=self.device)
        self.comm = comm
        self._saved = None

    def _params(self, p):
        c = self.cfg
        keep = []

        def chk(t, name):
            if t is None:
                return None
            if t.device.type != "cuda" or t.dtype != c.dtype or t.stride(-1) != 1:
                raise L.SpbError(f"{name}: expected a row-major {c.dtype} CUDA tensor, got {t.dtype} on {t.device}")
            keep.append(t)
            return t

        P = L.FfnParams()
        if not c.gated:
            w1, w3 = chk(p["w1"], "w1"), None
        elif "w13" in p:
            w = chk(p["w13"], "w13")
            w1, w3 = (w[:, :c.H], w[:, c.H:]) if c.weight_layout == L.W_IN_OUT else (w[:c.H], w[c.H:])
        else:
            w1, w3 = chk(p["w1"], "w1"), chk(p["w3"], "w3")
        if w3 is not None and w1.stride(0) != w3.stride(0):
            raise L.SpbError("w1 and w3 must share a row stride")
        w2 = chk(p["w2"], "w2")
        P.w1, P.w3, P.ldw_13 = _ptr(w1), _ptr(w3), w1.stride(0)
        P.b1, P.b2 = _ptr(chk(p.get("b1"), "b1")), _ptr(chk(p.get("b2"), "b2"))
        P.w2, P.ldw_2 = _ptr(w2), w2.stride(0)
        if c.norm_kind != L.NORM_NONE:
            P.norm_w = _ptr(chk(p["norm_w"], "norm_w"))
            P.norm_b = _ptr(chk(p.get("norm_b"), "norm_b"))
        return P, keep

    def forward(self, x, params):
        c = self.cfg
        if tuple(x.shape[-1:]) != (c.C,) or x.numel() != c.N * c.C or x.dtype != c.dtype or not x.is_contiguous():
            raise L.SpbError(f"x: expected contiguous [{c.N} tokens, {c.C}] {c.dtype}, got {tuple(x.shape)} {x.dtype}")
        P, keep = self._params(params)
        y = torch.empty_like(x)
        st = torch.cuda.current_stream(self.device).cuda_stream
        L.check(self.lib.spb_ffn_fwd(C.byref(self.d), C.byref(P), _ptr(x), _ptr(y), _ptr(self.ws), C.c_void_p(st)), "spb_ffn_fwd")
        self._saved = (x, params, P, keep)
        return y

    def alloc_grads(self, params):
        c = self.cfg
        f32 = dict(dtype=torch.float32, device=self.device)
        g = {"dw2": torch.empty(tuple(params["w2"].shape), **f32)}
        if "w13" in params:
            g["dw13"] = torch.empty(tuple(params["w13"].shape), **f32)
        else:
            g["dw1"] = torch.empty(tuple(params["w1"].shape), **f32)
            if c.gated:
                g["dw3"] = torch.empty(tuple(params["w3"].shape), **f32)
        for b in ("b1", "b2"):
            if params.get(b) is not None:
                g["d" + b] = torch.empty(tuple(params[b].shape), **f32)
        if c.norm_kind != L.NORM_NONE:
            g["dnorm_w"] = torch.empty(c.C, **f32)
            if c.norm_kind == L.NORM_LAYER:
                g["dnorm_b"] = torch.empty(c.C, **f32)
        return g

    def backward(self, dy, grads=None):
        c = self.cfg
        if self._saved is None:
            raise L.SpbError("backward() before forward()")
        x, params, P, keep = self._saved
        g = grads if grads is not None else self.alloc_grads(params)
        if "dx" not in g:
            g["dx"] = torch.empty_like(x)
        G = L.FfnGrads()
        G.dx = _ptr(g["dx"])
        G.dnorm_w, G.dnorm_b = _ptr(g.get("dnorm_w")), _ptr(g.get("dnorm_b"))
        if "dw13" in g:
            w = g["dw13"]
            d1, d3 = (w[:, :c.H], w[:, c.H:]) if c.weight_layout == L.W_IN_OUT else (w[:c.H], w[c.H:])
        else:
            d1, d3 = g["dw1"], g.get("dw3")
        G.dw1, G.dw3, G.ldw_13 = _ptr(d1), _ptr(d3), d1.stride(0)
        G.dw2, G.ldw_2 = _ptr(g["dw2"]), g["dw2"].stride(0)
        G.db1, G.db2 = _ptr(g.get("db1")), _ptr(g.get("db2"))
        st = torch.cuda.current_stream(self.device).cuda_stream
        L.check(self.lib.spb_ffn_bwd(C.byref(self.d), C.byref(P), _ptr(x), _ptr(dy.contiguous()), _ptr(self.ws), C.byref(G),
                                     C.c_void_p(self.comm) if self.comm else None, C.c_void_p(st)), "spb_ffn_bwd")
        return g


def _make_ffn(cfg, device, comm):
    return FeedForward(cfg, device=device, comm=comm)


def _pool():
    from .autograd import WorkspacePool
    global _POOL
    if _POOL is None:
        _POOL = WorkspacePool(_make_ffn)
    return _POOL


_POOL = None


def get_ffn(cfg, device, comm=None):
    """A private FeedForward (own workspace) for callers that drive forward/backward themselves."""
    return FeedForward(cfg, device=device, comm=comm)


class FfnFn(torch.autograd.Function):
    """y = ffn(x; weights) as a differentiable torch op.  `names` lists the keys of the tensors that follow.
    Every forward leases its own workspace (autograd.WorkspacePool): same-shape layers of one model do not share residuals."""

    @staticmethod
    def forward(ctx, cfg, names, comm, x, *tensors):
        params = dict(zip(names, tensors))
        lease = _pool().lease(cfg, x.device, comm)
        y = lease.owner.forward(x.contiguous(), params)
        ctx.lease, ctx.names, ctx.params = lease, names, params
        return y

    @staticmethod
    def backward(ctx, dy):
        lease = ctx.lease
        if lease is None or lease.owner is None:
            raise L.SpbError("this feed-forward's residuals were already released by an earlier backward (retain_graph is not supported)")
        g = lease.owner.backward(dy.contiguous())
        lease.release()
        out = []
        for n in ctx.names:
            gn = {"w13": "dw13", "w1": "dw1", "w3": "dw3", "w2": "dw2", "norm_w": "dnorm_w", "norm_b": "dnorm_b", "b1": "db1",
                  "b2": "db2"}.get(n)
            out.append(g[gn].to(ctx.params[n].dtype) if gn in g else None)
        return (None, None, None, g["dx"], *out)
