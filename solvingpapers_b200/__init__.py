"""solvingpapers_b200 - B200-native (sm_100a) attention block behind the notebooks' own call surface.

Product code only: hand-written CUDA in csrc/ behind the C ABI of include/spb200.h, plus thin host-side mirrors of
the reference's functions (llama, gpt, gemma, vit).  No CPU fallback, no oracle import."""
from . import _lib  # noqa: F401
from .block import AttentionBlock, BlockConfig, DecodeBlock, KVCache, llama_rope_table, gemma_rope_table  # noqa: F401
from .ffn import FeedForward, FfnConfig  # noqa: F401

__all__ = ["AttentionBlock", "BlockConfig", "DecodeBlock", "KVCache", "FeedForward", "FfnConfig", "llama_rope_table", "gemma_rope_table"]
