"""CPU: host-side logic - config/FLOP accounting, rotary tables, workspace layout, sharding, loud failure without CUDA."""
import ctypes
import math

import numpy as np
import pytest
import torch

import solvingpapers_b200 as sp
from solvingpapers_b200 import _lib as L
from solvingpapers_b200 import dist
from oracle import llama_oracle, gemma_oracle


def test_flop_accounting_matches_baseline_md():
    c3 = sp.BlockConfig(4, 2048, 4096, 32, 8, 128)
    assert abs(c3.flops_fwd_bwd() / 1e9 - 2542.9) < 0.1
    assert abs(c3.flops_fwd_bwd() / (4 * 2048) / 1e6 - 310.407) < 0.01
    c1 = sp.BlockConfig(16, 128, 256, 4, 2, 64)
    assert abs(c1.flops_fwd_bwd() / 1e9 - 2.889) < 0.001
    c5 = sp.BlockConfig(256, 197, 768, 12, 12, 64, causal=False)
    assert abs(c5.flops_fwd_bwd() / 1e9 - 820.7) < 0.1


def test_rope_tables():
    cos, sin = llama_oracle.precompute_freqs_cis(8, 5)
    t = sp.llama_rope_table(cos, sin)
    assert t.shape == (5, 4, 4) and t.dtype == torch.float32
    assert torch.equal(t[..., 0], cos) and torch.equal(t[..., 1], -sin) and torch.equal(t[..., 2], sin) and torch.equal(t[..., 3], cos)
    c, s = gemma_oracle.rotary_coeffs(6, 8)
    g = sp.gemma_rope_table(c, s, 4)
    assert g.shape == (6, 4, 4)
    assert torch.equal(g[:, 0], g[:, 3])                 # same angle for every pair
    assert torch.equal(g[..., 0], g[..., 1]) and torch.equal(g[..., 2], -g[..., 3])
    assert float(g[0, 0, 2]) == 0.0 and float(g[0, 0, 3]) == 0.0   # t=0: every odd output is 0


def test_llama_dropin_freqs_match_oracle_table():
    from solvingpapers_b200 import llama
    fc = llama.precompute_freqs_cis(64, 128, device="cpu")
    cos, sin = llama_oracle.precompute_freqs_cis(64, 128)
    assert fc.dtype == torch.complex64 and fc.shape == (128, 32)
    assert float((fc.real - cos).abs().max()) < 1e-6 and float((fc.imag - sin).abs().max()) < 1e-6


def test_workspace_layout_is_aligned_and_disjoint():
    lib = L.load()
    d = sp.BlockConfig(2, 197, 768, 12, 12, 64, dtype=torch.float32).desc()
    lay = L.WsLayout()
    assert lib.spb_block_ws_layout(ctypes.byref(d), ctypes.byref(lay)) == 0
    offs = [lay.xhat, lay.stats, lay.qkv, lay.attn_o, lay.lse, lay.delta, lay.dattn_o, lay.dqkv, lay.dq_acc, lay.dxhat, lay.partial]
    assert all(o % 256 == 0 for o in offs) and offs == sorted(offs) and len(set(offs)) == len(offs)
    assert lay.total >= lay.partial + lay.partial_bytes and lay.ld_qkv == 3 * 768
    N = 2 * 197
    assert lay.stats - lay.xhat >= N * 768 * 4


def test_no_cuda_means_loud_failure_not_fallback():
    if torch.cuda.is_available():
        pytest.skip("this box has a GPU")
    with pytest.raises(L.SpbError, match="no CPU path"):
        sp.AttentionBlock(sp.BlockConfig(1, 8, 8, 1, 1, 8))


def test_product_package_does_not_import_the_oracle():
    import os, re
    root = os.path.dirname(sp.__file__)
    for dp, _, files in os.walk(root):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dp, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f


def test_shard_bounds():
    assert dist.shard_bounds(32, 8, 0) == (0, 4) and dist.shard_bounds(32, 8, 7) == (28, 32)
    with pytest.raises(ValueError):
        dist.shard_bounds(30, 8, 0)
    with pytest.raises(ValueError):
        dist.shard_bounds(32, 8, 8)
    x = torch.arange(12).reshape(6, 2)
    assert torch.equal(torch.cat([dist.shard_batch(x, 3, r) for r in range(3)]), x)


def test_dropin_error_behaviour_is_checked_before_any_kernel():
    from solvingpapers_b200 import llama, gpt, vit
    x = torch.zeros(1, 4, 8)
    kv = torch.zeros(1, 5, 1, 4)
    with pytest.raises(ValueError):  # cache length and position disagree (llama3:815 slices the table from `position`)
        llama.attention({}, x, None, None, 2, 1, cache=(kv, kv), position=3)
    with pytest.raises(ValueError):  # position without a cache
        llama.attention({}, x, None, None, 2, 1, position=3)
    with pytest.raises(NotImplementedError):
        gpt.CausalSelfAttention(8, 2, 0.1)({}, x, deterministic=False)
    with pytest.raises(NotImplementedError):
        vit.MultiheadAttention(8, 2)({}, x, torch.zeros(1, 4, 8), x)
    with pytest.raises(AssertionError):
        gpt.CausalSelfAttention(10, 3, 0.0)


def test_workspace_pool_gives_every_live_forward_its_own_workspace():
    """llama3/LLaMA-jax.ipynb:925-927 calls transformer_block once per layer with identical shapes: leases of one
    config must be distinct while alive and be reused once released (no GPU needed: the factory is a stand-in)."""
    from solvingpapers_b200.autograd import WorkspacePool

    class Owner:
        pass

    pool = WorkspacePool(lambda cfg, device, comm: Owner())
    cfg = sp.BlockConfig(2, 64, 128, 4, 2, 32)
    a, b = pool.lease(cfg, "cuda:0", None), pool.lease(cfg, "cuda:0", None)
    assert a.owner is not b.owner and pool.created == 2
    first, second = a.owner, b.owner
    a.release()
    a.release()                                   # idempotent
    c = pool.lease(cfg, "cuda:0", None)
    assert c.owner is first and pool.created == 2  # reused, not re-created
    other = pool.lease(sp.BlockConfig(2, 64, 128, 4, 4, 32), "cuda:0", None)
    assert other.owner is not first and other.owner is not second and pool.created == 3
    del b                                          # dropped without a backward: goes back to the pool
    d = pool.lease(cfg, "cuda:0", None)
    assert d.owner is second and pool.created == 3
