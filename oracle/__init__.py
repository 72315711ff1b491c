"""CPU oracle for the attention-block hot path.  TEST INFRASTRUCTURE ONLY.

Everything under ``oracle/`` is a checker: a plain torch-CPU restatement of the
arithmetic the reference notebooks perform for the pre-norm attention block
(``rms_norm``/``LayerNorm`` -> QKV projection -> rotary -> softmax attention ->
output projection) with autograd for the backward.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl
reference`` legs may import it.  The product package ``solvingpapers_b200``
never imports it and has no CPU fallback.

Pinning status (see DESIGN.md "Oracle"):
  * llama  : PINNED  against the notebook's own cell source (c15-c18, c24)
             executed under numpy (jnp.* -> numpy.*), forward values in fp64 and
             gradients by central finite differences of that literal source
             (tests/golden/make_golden.py, tests/golden/llama_*.npz).
  * gemma  : PINNED  against the notebook's own torch modules executed on CPU
             (cells c0,c1,c6,c7,c8; tests/golden/gemma_*.npz).
  * vit    : PINNED  against torch.nn.LayerNorm + torch.nn.MultiheadAttention as the
             notebook calls them (tests/golden/vit_*.npz).
  * gpt    : PINNED  against the notebook's own module cells (gpt/gpt-jax.ipynb c9
             CausalSelfAttention, c10 MLP, c11 DecoderBlock) exec'd with jnp -> numpy and
             nn -> a numpy stand-in for the flax.linen names they use; forward in fp32 and
             fp64, gradients by finite differences of that source (tests/golden/gpt_*.npz).
             flax itself is not installable: its Dense / LayerNorm / gelu semantics are what
             the stand-in restates (third-party, version unpinned by the reference).
"""
from .metrics import max_rel_err  # noqa: F401
from . import llama_oracle, gpt_oracle, gemma_oracle, vit_oracle  # noqa: F401
