"""Drop-in for gemma/gemma.ipynb ``RMSNorm`` (:139-150), the closed form of ``RotaryEmbeddings`` (:169-208) and
``MQA`` (:218-259).  Weights keep torch's [out,in] layout and the module's attribute names.  The literal module is
n_q = no_of_heads // no_kv_heads full-width (D x D) query projections sharing ONE key and ONE value projection;
for the library that is Hq = n_q, Hkv = 1, hd = D with the q|k|v weights packed as one [(n_q+2)*D, D] array.
The dense [T,D,D] "rotary" matrix is never built: it is the 2x2 table (c, c, -s, s) with one angle per position."""
import math

import torch

from . import _lib as L
from .autograd import BlockFn, NormFn
from .block import BlockConfig, gemma_rope_table


def rotary_coeffs(T, D, device="cuda"):
    """cos/sin of a_t = t * 10000^(-2(t-1)/D) (gemma/gemma.ipynb:184-188), fp32 like the notebook."""
    pos = torch.arange(T).unsqueeze(1)
    ang = pos * (10000 ** (-2 * (pos - 1) / D))
    return torch.cos(ang).squeeze(1).to(device), torch.sin(ang).squeeze(1).to(device)


class RMSNorm:
    def __init__(self, dim: int, eps: float = 1e-6):
        self.dim, self.eps = dim, eps

    def __call__(self, weight, x):
        return NormFn.apply(L.NORM_RMS, float(self.eps), True, x, weight, None)


class MQA:
    """forward(x) of gemma/gemma.ipynb:251-259 in eval mode (the dropouts are identity)."""

    def __init__(self, embeddings_dims=768, block_size=128, no_of_kv_heads=2, no_of_heads=4):
        self.D = embeddings_dims
        self.no_of_q_heads = no_of_heads // no_of_kv_heads if no_of_kv_heads > 0 else 1  # gemma:228
        self.training = False

    @staticmethod
    def pack(multi_query, key, value):
        """[(n_q+2)*D, D] packed weight from the module's Linear weights (one-off, at load time)."""
        return torch.cat(list(multi_query) + [key, value], dim=0).contiguous()

    def _cfg(self, x, norm):
        B, T, D = x.shape
        return BlockConfig(B=B, T=T, C=D, Hq=self.no_of_q_heads, Hkv=1, hd=D, dtype=x.dtype,
                           norm_kind=L.NORM_RMS if norm else L.NORM_NONE, norm_eps=1e-6, norm_upcast=True, rope=True,
                           causal=True, scale=1.0 / math.sqrt(D), weight_layout=L.W_OUT_IN)

    def forward(self, params, x, norm_weight=None):
        """params = {'wqkv': packed [(n_q+2)D, D], 'linear_layer': [D, n_q*D]}; with norm_weight the call is
        ``mqa(norm1(x))`` of DecoderLayer (gemma/gemma.ipynb:335) fused."""
        if self.training:
            raise NotImplementedError("train-mode dropout is not reproduced; call in eval mode")
        B, T, D = x.shape
        c, s = rotary_coeffs(T, D, x.device)
        tab = gemma_rope_table(c, s, D // 2)
        if norm_weight is None:
            y, _, _ = BlockFn.apply(self._cfg(x, False), ("wqkv", "wo", "rope"), None, x, params["wqkv"],
                                    params["linear_layer"], tab)
        else:
            y, _, _ = BlockFn.apply(self._cfg(x, True), ("wqkv", "wo", "norm_w", "rope"), None, x, params["wqkv"],
                                    params["linear_layer"], norm_weight, tab)
        return y

    __call__ = forward


class GeGLU:
    """gemma/gemma.ipynb:269-286, same call surface: ``GeGLU(...)(x)`` = linear_layer3(gelu(linear_layer1(x)) * linear_layer2(x)).
    `params` holds the torch ``nn.Linear`` weights as they are stored, [out, in]: linear_layer1, linear_layer2 [4D, D],
    linear_layer3 [D, 4D] (one spb_ffn_fwd / spb_ffn_bwd pair, erf-GELU gate kernel)."""

    def __init__(self, params, comm=None):
        self.params, self.comm = params, comm

    def __call__(self, x):
        from .ffn import FfnConfig, FfnFn
        p = self.params
        D = x.shape[-1]
        cfg = FfnConfig(N=x.numel() // D, C=D, H=p["linear_layer1"].shape[0], dtype=x.dtype, norm_kind=L.NORM_NONE, act=L.ACT_GEGLU,
                        weight_layout=L.W_OUT_IN)
        # gate = linear_layer1 (w3), value = linear_layer2 (w1), down projection = linear_layer3 (w2)
        y = FfnFn.apply(cfg, ("w3", "w1", "w2"), self.comm, x.reshape(-1, D), p["linear_layer1"], p["linear_layer2"], p["linear_layer3"])
        return y.reshape(x.shape)

    forward = __call__
