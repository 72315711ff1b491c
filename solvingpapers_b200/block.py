"""Host side of the attention block: owns the workspace, marshals torch CUDA tensors into the C ABI
(include/spb200.h) and exposes forward / backward.  torch is plumbing only (device memory, streams);
all arithmetic happens in libspb200.so.  There is no fallback path.

Mirrors the static arguments of the reference's block functions (llama3/LLaMA-jax.ipynb:809 attention +
:536 rms_norm; gpt/gpt-jax.ipynb:321-357; gemma/gemma.ipynb:218-259; ViT.ipynb:205-221) - see BlockConfig.
"""
import ctypes as C
import math
from dataclasses import dataclass, field
from typing import Optional

import torch

from . import _lib as L


@dataclass
class BlockConfig:
    B: int
    T: int
    C: int
    Hq: int
    Hkv: int
    hd: int
    dtype: torch.dtype = torch.bfloat16
    norm_kind: int = L.NORM_RMS
    norm_eps: float = 1e-6
    norm_upcast: bool = False
    rope: bool = True
    causal: bool = True
    scale: Optional[float] = None
    weight_layout: int = L.W_IN_OUT
    qkv_bias: bool = False
    out_bias: bool = False
    attn_impl: int = L.IMPL_AUTO
    gemm_impl: int = L.IMPL_AUTO
    residual: bool = False  # y = x + block(x) and dx += dy inside the kernels (the reference's `h = x + attention(...)`, llama3:885)

    def desc(self) -> L.BlockDesc:
        scale = self.scale if self.scale is not None else 1.0 / math.sqrt(self.hd)
        return L.BlockDesc(self.B, self.T, self.C, self.Hq, self.Hkv, self.hd, _dt(self.dtype), self.norm_kind,
                           self.norm_eps, int(self.norm_upcast), L.ROPE_TABLE if self.rope else L.ROPE_NONE,
                           int(self.causal), scale, self.weight_layout, int(self.qkv_bias), int(self.out_bias),
                           self.attn_impl, self.gemm_impl, int(self.residual))

    @property
    def n_q(self):
        return self.Hq * self.hd

    @property
    def n_kv(self):
        return self.Hkv * self.hd

    @property
    def n_tot(self):
        return self.n_q + 2 * self.n_kv

    def flops_fwd_bwd(self) -> float:
        """Algorithmic FLOPs of one forward+backward: 3*F_gemm + 3.5*F_attn."""
        N = self.B * self.T
        f_g = 2.0 * N * self.C * self.n_tot + 2.0 * N * self.n_q * self.C
        m = (self.T + 1) / (2.0 * self.T) if self.causal else 1.0
        f_a = 4.0 * self.B * self.Hq * self.T * self.T * self.hd * m
        return 3.0 * f_g + 3.5 * f_a


def _dt(dtype):
    if dtype == torch.float32:
        return L.F32
    if dtype == torch.bfloat16:
        return L.BF16
    raise L.SpbError(f"unsupported dtype {dtype} (fp32 and bf16 only)")


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(None)


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def llama_rope_table(cos, sin):
    """freqs_cis of llama3/LLaMA-jax.ipynb:563-567 as the library's 2x2 table: (c, -s, s, c) per (t, pair)."""
    return torch.stack([cos, -sin, sin, cos], dim=-1).float().contiguous()


def gemma_rope_table(cos_t, sin_t, pairs):
    """The gemma notebook's matrix (gemma/gemma.ipynb:182-200): block [[c, c], [-s, s]], same angle for every pair."""
    c = cos_t.float().reshape(-1, 1).expand(-1, pairs)
    s = sin_t.float().reshape(-1, 1).expand(-1, pairs)
    return torch.stack([c, c, -s, s], dim=-1).contiguous()


class AttentionBlock:
    """One pre-norm attention block bound to fixed shapes.  Weights are passed per call (they change every step)."""

    def __init__(self, cfg: BlockConfig, device="cuda", comm=None):
        if not torch.cuda.is_available():
            raise L.SpbError("solvingpapers_b200 needs a CUDA device (B200, sm_100a); there is no CPU path")
        self.cfg = cfg
        self.lib = L.load()
        self.device = torch.device(device)
        self.d = cfg.desc()
        self.layout = L.WsLayout()
        L.check(self.lib.spb_block_ws_layout(C.byref(self.d), C.byref(self.layout)), "spb_block_ws_layout")
        self.ws = torch.empty(self.layout.total, dtype=torch.uint8, device=self.device)
        assert self.ws.data_ptr() % 256 == 0
        self.comm = comm  # opaque handle from dist.Communicator or None
        self._saved = None

    # ---- views into the workspace --------------------------------------------------------------------
    def _view(self, off, shape, dtype):
        n = 1
        for s in shape:
            n *= s
        return self.ws[off: off + n * torch.empty((), dtype=dtype).element_size()].view(dtype).view(*shape)

    def qkv_view(self):
        c = self.cfg
        return self._view(self.layout.qkv, (c.B, c.T, c.n_tot), c.dtype)

    def kv_cache(self):
        """(k, v) of llama3/LLaMA-jax.ipynb:819: post-rotary, un-repeated, [B,T,Hkv,hd] (strided views)."""
        c = self.cfg
        qkv = self.qkv_view()
        k = qkv[..., c.n_q: c.n_q + c.n_kv].unflatten(-1, (c.Hkv, c.hd))
        v = qkv[..., c.n_q + c.n_kv:].unflatten(-1, (c.Hkv, c.hd))
        return k, v

    def residual_views(self):
        """2-D views of what forward/backward left in the workspace (spb_ws_layout): the operands of the weight-gradient
        GEMMs.  Used by bench.py's multi-rank check to recompute the LOCAL gradients after an all-reduced backward."""
        c = self.cfg
        N = c.B * c.T
        lay = self.layout
        return {"xhat": self._view(lay.xhat, (N, c.C), c.dtype), "stats": self._view(lay.stats, (N, 2), torch.float32),
                "qkv": self._view(lay.qkv, (N, c.n_tot), c.dtype), "attn_o": self._view(lay.attn_o, (N, c.n_q), c.dtype),
                "dqkv": self._view(lay.dqkv, (N, c.n_tot), c.dtype), "dxhat": self._view(lay.dxhat, (N, c.C), c.dtype)}

    def lse_view(self):
        c = self.cfg
        return self._view(self.layout.lse, (c.B, c.Hq, c.T), torch.float32)

    # ---- params ----------------------------------------------------------------------------------------
    def _params(self, p, check_rope=True):
        c = self.cfg
        P = L.BlockParams()
        keep = []

        def chk(t, name):
            if t is None:
                return None
            if t.device.type != "cuda" or t.dtype != c.dtype:
                raise L.SpbError(f"{name}: expected a {c.dtype} CUDA tensor, got {t.dtype} on {t.device}")
            keep.append(t)
            return t

        if "wqkv" in p:
            w = chk(p["wqkv"], "wqkv")
            if w.stride(-1) != 1:
                raise L.SpbError("wqkv must be row-major")
            if c.weight_layout == L.W_IN_OUT:
                wq, wk, wv = w[:, :c.n_q], w[:, c.n_q:c.n_q + c.n_kv], w[:, c.n_q + c.n_kv:]
            else:
                wq, wk, wv = w[:c.n_q], w[c.n_q:c.n_q + c.n_kv], w[c.n_q + c.n_kv:]
        else:
            wq, wk, wv = chk(p["wq"], "wq"), chk(p["wk"], "wk"), chk(p["wv"], "wv")
        for t in (wq, wk, wv):
            if t.stride(-1) != 1:
                raise L.SpbError("weights must be row-major (last stride 1)")
        P.wq, P.wk, P.wv = _ptr(wq), _ptr(wk), _ptr(wv)
        P.ldw_q, P.ldw_k, P.ldw_v = wq.stride(0), wk.stride(0), wv.stride(0)
        wo = chk(p["wo"], "wo")
        P.wo, P.ldw_o = _ptr(wo), wo.stride(0)
        if c.qkv_bias:
            if "bqkv" in p:
                b = chk(p["bqkv"], "bqkv")
                bq, bk, bv = b[:c.n_q], b[c.n_q:c.n_q + c.n_kv], b[c.n_q + c.n_kv:]
            else:
                bq, bk, bv = chk(p["bq"], "bq"), chk(p["bk"], "bk"), chk(p["bv"], "bv")
            P.bq, P.bk, P.bv = _ptr(bq), _ptr(bk), _ptr(bv)
        if c.out_bias:
            P.bo = _ptr(chk(p["bo"], "bo"))
        if c.norm_kind != L.NORM_NONE:
            P.norm_w = _ptr(chk(p["norm_w"], "norm_w"))
            if c.norm_kind == L.NORM_LAYER and p.get("norm_b") is not None:
                P.norm_b = _ptr(chk(p["norm_b"], "norm_b"))
        if c.rope:
            r = p["rope"]
            if check_rope and (r.dtype != torch.float32 or tuple(r.shape) != (c.T, c.hd // 2, 4) or not r.is_contiguous()):
                raise L.SpbError(f"rope table must be contiguous fp32 [T, hd/2, 4], got {tuple(r.shape)} {r.dtype}")
            keep.append(r)
            P.rope = _ptr(r)
        return P, keep

    # ---- forward / backward -----------------------------------------------------------------------------
    def forward(self, x, params):
        c = self.cfg
        if tuple(x.shape) != (c.B, c.T, c.C) or x.dtype != c.dtype or not x.is_contiguous():
            raise L.SpbError(f"x must be contiguous {c.dtype} [{c.B},{c.T},{c.C}], got {tuple(x.shape)} {x.dtype}")
        P, keep = self._params(params)
        y = torch.empty_like(x)
        L.check(self.lib.spb_block_fwd(C.byref(self.d), C.byref(P), _ptr(x), _ptr(y), _ptr(self.ws), _stream()),
                "spb_block_fwd")
        self._saved = (x, params, P, keep)
        return y

    def alloc_grads(self, params):
        """fp32 gradient buffers with the same packing as the weights."""
        c = self.cfg
        dev = self.device
        g = {"dx": None}
        if "wqkv" in params:
            g["dwqkv"] = torch.empty(params["wqkv"].shape, dtype=torch.float32, device=dev)
        else:
            for k in ("wq", "wk", "wv"):
                g["d" + k] = torch.empty(params[k].shape, dtype=torch.float32, device=dev)
        g["dwo"] = torch.empty(params["wo"].shape, dtype=torch.float32, device=dev)
        if c.qkv_bias:
            g["dbqkv"] = torch.empty(c.n_tot, dtype=torch.float32, device=dev)
        if c.out_bias:
            g["dbo"] = torch.empty(c.C, dtype=torch.float32, device=dev)
        if c.norm_kind != L.NORM_NONE:
            g["dnorm_w"] = torch.empty(c.C, dtype=torch.float32, device=dev)
            if c.norm_kind == L.NORM_LAYER:
                g["dnorm_b"] = torch.empty(c.C, dtype=torch.float32, device=dev)
        return g

    def backward(self, dy, grads=None, need_dx=True):
        """VJP for the last forward().  Returns a dict of gradients (weight grads fp32, dx in the activation dtype)."""
        if self._saved is None:
            raise L.SpbError("backward() before forward()")
        c = self.cfg
        x, params, P, keep = self._saved
        if tuple(dy.shape) != (c.B, c.T, c.C) or dy.dtype != c.dtype or not dy.is_contiguous():
            raise L.SpbError("dy must match x")
        g = grads if grads is not None else self.alloc_grads(params)
        if need_dx and g.get("dx") is None:
            g["dx"] = torch.empty_like(x)
        G = L.BlockGrads()
        G.dx = _ptr(g.get("dx"))
        if "dwqkv" in g:
            w = g["dwqkv"]
            if c.weight_layout == L.W_IN_OUT:
                dwq, dwk, dwv = w[:, :c.n_q], w[:, c.n_q:c.n_q + c.n_kv], w[:, c.n_q + c.n_kv:]
            else:
                dwq, dwk, dwv = w[:c.n_q], w[c.n_q:c.n_q + c.n_kv], w[c.n_q + c.n_kv:]
        else:
            dwq, dwk, dwv = g["dwq"], g["dwk"], g["dwv"]
        G.dwq, G.dwk, G.dwv = _ptr(dwq), _ptr(dwk), _ptr(dwv)
        G.ldw_q, G.ldw_k, G.ldw_v = dwq.stride(0), dwk.stride(0), dwv.stride(0)
        G.dwo, G.ldw_o = _ptr(g["dwo"]), g["dwo"].stride(0)
        if c.qkv_bias:
            b = g["dbqkv"]
            G.dbq, G.dbk, G.dbv = _ptr(b[:c.n_q]), _ptr(b[c.n_q:c.n_q + c.n_kv]), _ptr(b[c.n_q + c.n_kv:])
        if c.out_bias:
            G.dbo = _ptr(g["dbo"])
        if c.norm_kind != L.NORM_NONE:
            G.dnorm_w = _ptr(g["dnorm_w"])
            if c.norm_kind == L.NORM_LAYER:
                G.dnorm_b = _ptr(g["dnorm_b"])
        comm = C.c_void_p(self.comm) if self.comm else C.c_void_p(None)
        L.check(self.lib.spb_block_bwd(C.byref(self.d), C.byref(P), _ptr(x), _ptr(dy), _ptr(self.ws), C.byref(G), comm,
                                       _stream()), "spb_block_bwd")
        return g


    def capture_step(self, x, dy, params, grads=None, warmup=3):
        """Forward + backward over fixed buffers recorded once into a CUDA graph; ``replay()`` then re-launches the whole step
        with one driver call.  For launch-bound shapes (the notebook's own dims: ~25 kernels of a few microseconds each) the
        replay removes the per-launch host cost; at the headline shape the step is device-bound and it changes nothing.  The
        buffers are the ones passed here: refill ``x`` / ``dy`` / the weights in place between replays."""
        if self.comm:
            raise L.SpbError("capture_step: the gradient all-reduce is not captured; use forward()/backward() with a communicator")
        g = grads if grads is not None else self.alloc_grads(params)
        if g.get("dx") is None:
            g["dx"] = torch.empty_like(x)
        side = torch.cuda.Stream(device=self.device)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(max(warmup, 1)):          # first-use set-up (function attributes, tensor maps) happens outside capture
                self.forward(x, params)
                self.backward(dy, grads=g)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            y = self.forward(x, params)
            self.backward(dy, grads=g)
        return CapturedStep(graph, y, g)


class CapturedStep:
    """A recorded forward+backward (AttentionBlock.capture_step): ``y`` and ``grads`` are rewritten by every ``replay()``."""

    def __init__(self, graph, y, grads):
        self.graph, self.y, self.grads = graph, y, grads

    def replay(self):
        self.graph.replay()
        return self.y, self.grads


class KVCache:
    """Preallocated key/value cache [B, Tmax, Hkv, hd] that spb_block_decode appends to in place - the O(1)-per-token form of
    the notebook's ``cache`` tuple, whose ``jnp.concatenate`` (llama3/LLaMA-jax.ipynb:817-818) copies every old row on every
    step.  ``length`` rows are valid; ``k`` / ``v`` are views of them."""

    def __init__(self, B, Tmax, Hkv, hd, dtype=torch.bfloat16, device="cuda"):
        self.kbuf = torch.zeros((B, Tmax, Hkv, hd), dtype=dtype, device=device)
        self.vbuf = torch.zeros((B, Tmax, Hkv, hd), dtype=dtype, device=device)
        self.length = 0
        self._len_dev = None     # int32 on the device, read and advanced by recorded steps (DecodeBlock.capture)
        self._len_dev_is = -1    # the value the device copy is known to hold

    def length_dev(self):
        """The cache length as one int32 in device memory, brought up to date with ``length`` (a copy only when eager steps
        or the caller changed ``length`` since the last recorded step)."""
        if self._len_dev is None:
            self._len_dev = torch.zeros(1, dtype=torch.int32, device=self.kbuf.device)
            self._len_dev_is = 0
        if self._len_dev_is != self.length:
            self._len_dev.fill_(self.length)
            self._len_dev_is = self.length
        return self._len_dev

    @property
    def capacity(self):
        return self.kbuf.shape[1]

    @property
    def k(self):
        return self.kbuf[:, :self.length]

    @property
    def v(self):
        return self.vbuf[:, :self.length]

    @classmethod
    def from_tensors(cls, k, v, extra):
        """A cache holding the rows of the notebook-style ``(k, v)`` tuple ([B, Tc, Hkv, hd]) with room for ``extra`` more."""
        B, Tc, Hkv, hd = k.shape
        c = cls(B, Tc + extra, Hkv, hd, dtype=k.dtype, device=k.device)
        c.kbuf[:, :Tc].copy_(k)
        c.vbuf[:, :Tc].copy_(v)
        c.length = Tc
        return c


class DecodeBlock:
    """One decode step of the block bound to fixed shapes (B sequences, Tq new tokens each, cache capacity Tmax):
    spb_block_decode.  The config's T is Tq.  Weights are passed per call like AttentionBlock's."""

    def __init__(self, cfg: BlockConfig, Tmax: int, device="cuda"):
        if not torch.cuda.is_available():
            raise L.SpbError("solvingpapers_b200 needs a CUDA device (B200, sm_100a); there is no CPU path")
        self.cfg, self.Tmax = cfg, int(Tmax)
        self.lib = L.load()
        self.device = torch.device(device)
        self.d = cfg.desc()
        self.layout = L.DecodeWs()
        L.check(self.lib.spb_block_decode_ws_layout(C.byref(self.d), self.Tmax, C.byref(self.layout)), "spb_block_decode_ws_layout")
        self.ws = torch.empty(self.layout.total, dtype=torch.uint8, device=self.device)
        self._marshal = AttentionBlock._params  # same parameter packing rules

    def step(self, x, params, cache: KVCache):
        c = self.cfg
        if tuple(x.shape) != (c.B, c.T, c.C) or x.dtype != c.dtype or not x.is_contiguous():
            raise L.SpbError(f"x must be contiguous {c.dtype} [{c.B},{c.T},{c.C}], got {tuple(x.shape)} {x.dtype}")
        if cache.capacity != self.Tmax or cache.kbuf.dtype != c.dtype or tuple(cache.kbuf.shape) != (c.B, self.Tmax, c.Hkv, c.hd):
            raise L.SpbError("cache does not match this DecodeBlock (shape / capacity / dtype)")
        if c.rope:
            need = cache.length + c.T
            r = params["rope"]
            if r.dtype != torch.float32 or r.dim() != 3 or r.shape[0] < need or tuple(r.shape[1:]) != (c.hd // 2, 4) or not r.is_contiguous():
                raise L.SpbError(f"rope table must be contiguous fp32 [>= {need}, hd/2, 4]")
        P, keep = self._marshal(self, params, check_rope=False)
        y = torch.empty_like(x)
        L.check(self.lib.spb_block_decode(C.byref(self.d), C.byref(P), _ptr(x), _ptr(cache.kbuf), _ptr(cache.vbuf), cache.length, self.Tmax,
                                          _ptr(y), _ptr(self.ws), _stream()), "spb_block_decode")
        cache.length += c.T
        return y

    def capture(self, x, params, cache: KVCache):
        """Record one decode step over fixed buffers (``x``, the weights, ``cache``) into a CUDA graph through
        spb_block_decode_dev: the cache length is read from device memory and advanced by the step itself, so the same graph
        serves every token.  ``replay()`` -> y (a fixed buffer, rewritten by every replay); refill ``x`` in place between
        replays.  The eager step is bound by the cost of its seven launches; the replay is one driver call."""
        c = self.cfg
        if tuple(x.shape) != (c.B, c.T, c.C) or x.dtype != c.dtype or not x.is_contiguous():
            raise L.SpbError(f"x must be contiguous {c.dtype} [{c.B},{c.T},{c.C}], got {tuple(x.shape)} {x.dtype}")
        if cache.capacity != self.Tmax or cache.kbuf.dtype != c.dtype or tuple(cache.kbuf.shape) != (c.B, self.Tmax, c.Hkv, c.hd):
            raise L.SpbError("cache does not match this DecodeBlock (shape / capacity / dtype)")
        if c.rope:
            r = params["rope"]
            if r.dtype != torch.float32 or r.dim() != 3 or r.shape[0] < self.Tmax or tuple(r.shape[1:]) != (c.hd // 2, 4) or not r.is_contiguous():
                raise L.SpbError(f"rope table must be contiguous fp32 [>= Tmax = {self.Tmax}, hd/2, 4] for a recorded step")
        P, keep = self._marshal(self, params, check_rope=False)
        y = torch.empty_like(x)
        start = cache.length
        len_dev = cache.length_dev()

        def run():
            L.check(self.lib.spb_block_decode_dev(C.byref(self.d), C.byref(P), _ptr(x), _ptr(cache.kbuf), _ptr(cache.vbuf), _ptr(len_dev),
                                                  self.Tmax, _ptr(y), _ptr(self.ws), _stream()), "spb_block_decode_dev")

        side = torch.cuda.Stream(device=self.device)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            run()                      # first-use set-up (function attributes) outside the capture; it appends a row ...
            len_dev.fill_(start)       # ... which the first replay overwrites: the length goes back to where it was
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            run()
        return CapturedDecode(self, graph, y, cache, (x, params, keep, len_dev))

    def weight_bytes(self):
        c = self.cfg
        return (c.C * c.n_tot + c.n_q * c.C) * (2 if c.dtype == torch.bfloat16 else 4)


class CapturedDecode:
    """A recorded decode step (DecodeBlock.capture).  The host mirrors the cache length so that ``cache.length`` / ``cache.k``
    stay usable and a replay past the capacity is refused."""

    def __init__(self, block, graph, y, cache, keep):
        self.block, self.graph, self.y, self.cache, self._keep = block, graph, y, cache, keep

    def replay(self):
        c, cache = self.block.cfg, self.cache
        if cache.length + c.T > cache.capacity:
            raise L.SpbError(f"decode replay: cache_len {cache.length} + Tq {c.T} exceeds the cache capacity {cache.capacity}")
        cache.length_dev()            # re-synchronise if eager steps moved the length since the last replay
        self.graph.replay()
        cache.length += c.T
        cache._len_dev_is = cache.length
        return self.y
