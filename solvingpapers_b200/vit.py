"""Drop-in for the attention half of vision transformer/ViT.ipynb ``TransformerEncoder`` (:202-227):
``nn.LayerNorm(E)`` (torch eps 1e-5) followed by ``nn.MultiheadAttention(E, H, batch_first=True)(x, x, x)[0]``.
Parameter names are torch's (in_proj_weight [3E,E], in_proj_bias, out_proj.weight, out_proj.bias)."""
from . import _lib as L
from .autograd import BlockFn, NormFn
from .block import BlockConfig


def layer_norm(weight, bias, x, eps=1e-5):
    return NormFn.apply(L.NORM_LAYER, float(eps), False, x, weight, bias)


class MultiheadAttention:
    def __init__(self, embed_dim, num_heads, batch_first=True):
        if not batch_first:
            raise NotImplementedError("the notebook uses batch_first=True (ViT.ipynb:208)")
        assert embed_dim % num_heads == 0
        self.E, self.H, self.hd = embed_dim, num_heads, embed_dim // num_heads

    def _cfg(self, x, norm):
        B, T, E = x.shape
        return BlockConfig(B=B, T=T, C=E, Hq=self.H, Hkv=self.H, hd=self.hd, dtype=x.dtype,
                           norm_kind=L.NORM_LAYER if norm else L.NORM_NONE, norm_eps=1e-5, rope=False, causal=False,
                           scale=self.hd ** -0.5, weight_layout=L.W_OUT_IN, qkv_bias=True, out_bias=True)

    def forward(self, params, query, key=None, value=None):
        """Self-attention only (the notebook passes x three times, ViT.ipynb:221).  Returns (out, None) like torch
        with need_weights=False; the averaged attention map the notebook computes and discards is never built."""
        if (key is not None and key is not query) or (value is not None and value is not query):
            raise NotImplementedError("only self-attention (q is k is v) is on the hot path")
        y, _, _ = BlockFn.apply(self._cfg(query, False), ("wqkv", "bqkv", "wo", "bo"), None, query,
                                params["in_proj_weight"], params["in_proj_bias"], params["out_proj.weight"],
                                params["out_proj.bias"])
        return y, None

    __call__ = forward

    def pre_ln(self, params, x):
        """``multihead_attention(layer_norm1(x))`` (ViT.ipynb:219-221) fused; params also has 'ln.weight','ln.bias'."""
        y, _, _ = BlockFn.apply(self._cfg(x, True), ("wqkv", "bqkv", "wo", "bo", "norm_w", "norm_b"), None, x,
                                params["in_proj_weight"], params["in_proj_bias"], params["out_proj.weight"],
                                params["out_proj.bias"], params["ln.weight"], params["ln.bias"])
        return y
