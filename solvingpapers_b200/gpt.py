"""Drop-in for gpt/gpt-jax.ipynb:321-357 ``CausalSelfAttention`` and the pre-LN call of ``DecoderBlock`` (:414,420).
Same constructor fields and call signature as the flax module; parameters use flax's tree
({'qkv': {'kernel' [C,3C]}, 'proj': {'kernel' [C,C], 'bias' [C]}}; LayerNorm {'scale','bias'}), as torch CUDA tensors."""
import math

from . import _lib as L
from .autograd import BlockFn, NormFn
from .block import BlockConfig


def layer_norm(params, x, eps=1e-6):
    """flax nn.LayerNorm() (eps 1e-6) as used at gpt/gpt-jax.ipynb:414."""
    return NormFn.apply(L.NORM_LAYER, float(eps), False, x, params["scale"], params["bias"])


class CausalSelfAttention:
    def __init__(self, emb_dim: int, num_heads: int, dropout_rate: float = 0.0):
        assert emb_dim % num_heads == 0  # gpt/gpt-jax.ipynb:327
        self.emb_dim, self.num_heads, self.dropout_rate = emb_dim, num_heads, dropout_rate
        self.head_dim = emb_dim // num_heads

    def _cfg(self, x, norm):
        B, T, C = x.shape
        return BlockConfig(B=B, T=T, C=C, Hq=self.num_heads, Hkv=self.num_heads, hd=self.head_dim, dtype=x.dtype,
                           norm_kind=L.NORM_LAYER if norm else L.NORM_NONE, norm_eps=1e-6, rope=False, causal=True,
                           scale=1.0 / math.sqrt(self.head_dim), weight_layout=L.W_IN_OUT, out_bias=True)

    def __call__(self, params, x, deterministic=True):
        if not deterministic and self.dropout_rate > 0.0:
            raise NotImplementedError("dropout (deterministic=False) is not reproduced: flax's RNG stream cannot be matched")
        y, _, _ = BlockFn.apply(self._cfg(x, False), ("wqkv", "wo", "bo"), None, x, params["qkv"]["kernel"],
                                params["proj"]["kernel"], params["proj"]["bias"])
        return y

    def pre_ln(self, params, x, deterministic=True):
        """``self.attn(self.ln1(x))`` of DecoderBlock (gpt/gpt-jax.ipynb:420) as one fused block call.
        params = {'ln1': {'scale','bias'}, 'attn': {...}}."""
        if not deterministic and self.dropout_rate > 0.0:
            raise NotImplementedError("dropout (deterministic=False) is not reproduced")
        a = params["attn"]
        y, _, _ = BlockFn.apply(self._cfg(x, True), ("wqkv", "wo", "bo", "norm_w", "norm_b"), None, x, a["qkv"]["kernel"],
                                a["proj"]["kernel"], a["proj"]["bias"], params["ln1"]["scale"], params["ln1"]["bias"])
        return y


class MLP:
    """gpt/gpt-jax.ipynb:376-389 with deterministic=True: ``fc2(nn.gelu(fc1(x)))``.  `params` is the flax parameter tree of the
    module: {'fc1': {'kernel' [C,4C], 'bias' [4C]}, 'fc2': {'kernel' [4C,C], 'bias' [C]}} (one spb_ffn_fwd /
    spb_ffn_bwd pair: bias in the GEMM epilogues, tanh-GELU kernel, bias gradients as column sums)."""

    def __init__(self, emb_dim=None, dropout_rate=0.0, comm=None):
        self.emb_dim, self.dropout_rate, self.comm = emb_dim, dropout_rate, comm

    def apply(self, variables, x, deterministic=True):
        if not deterministic and self.dropout_rate != 0.0:
            raise NotImplementedError("dropout has no kernel (its RNG stream cannot match flax's); call with deterministic=True")
        from .ffn import FfnConfig, FfnFn
        p = variables.get("params", variables)
        C_ = x.shape[-1]
        cfg = FfnConfig(N=x.numel() // C_, C=C_, H=p["fc1"]["kernel"].shape[1], dtype=x.dtype, norm_kind=L.NORM_NONE, act=L.ACT_GELU_TANH)
        y = FfnFn.apply(cfg, ("w1", "b1", "w2", "b2"), self.comm, x.reshape(-1, C_), p["fc1"]["kernel"], p["fc1"]["bias"], p["fc2"]["kernel"],
                        p["fc2"]["bias"])
        return y.reshape(x.shape)

    __call__ = apply
