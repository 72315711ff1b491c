"""GPU parity of the KV-cache decode path (llama3/LLaMA-jax.ipynb:815-819 + caller :499-511):
spb_attn_decode, spb_block_decode and the drop-in ``attention(..., cache=, position=)`` against the oracle's cache branch
and fixtures made from the notebook's own attention cell."""
import math
import os

import numpy as np
import pytest
import torch

from oracle import llama_oracle, max_rel_err
from oracle.metrics import TOL_BF16, TOL_F32

pytestmark = pytest.mark.gpu
DEV = "cuda"
T_ = torch.from_numpy


def rnd(shape, seed, dtype=torch.float32, scale=1.0):
    g = np.random.default_rng(seed)
    return T_(g.standard_normal(shape).astype(np.float32) * scale).to(dtype)


def tol(dtype):
    return TOL_BF16 if dtype == torch.bfloat16 else TOL_F32


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("B,Tq,Hq,Hkv,hd,Tk", [
    (2, 1, 4, 2, 64, 22),       # the notebook's head geometry, short cache
    (4, 1, 32, 8, 128, 2048),   # Llama-8B head geometry at the C3 length: 16 key tiles, split over CTAs
    (1, 1, 8, 8, 128, 1000),    # MHA, ragged length
    (2, 3, 4, 1, 64, 131),      # several new tokens, MQA: 12 query rows per kv head -> two row groups
    (1, 1, 1, 1, 256, 300),     # one head of 256
    (3, 2, 2, 2, 64, 1),        # a single key
    (2, 5, 8, 2, 128, 777),     # 20 query rows per kv head: more than the 16 rows of one MMA tile -> two row groups (bf16 path)
    (1, 2, 4, 1, 256, 1500),    # wide heads, several key tiles, MQA
])
def test_attn_decode_against_oracle_math(dtype, B, Tq, Hq, Hkv, hd, Tk):
    from solvingpapers_b200 import ops
    Tmax = Tk + 5  # the cache is larger than what is valid: rows past Tk must not be read
    q = rnd((B, Tq, Hq, hd), 1, dtype)
    kc = rnd((B, Tmax, Hkv, hd), 2, dtype)
    vc = rnd((B, Tmax, Hkv, hd), 3, dtype)
    kc[:, Tk:] = float("nan")
    vc[:, Tk:] = float("nan")
    scale = 1.0 / math.sqrt(hd)
    n_rep = Hq // Hkv
    kk = llama_oracle.repeat_kv(kc[:, :Tk].double(), n_rep).permute(0, 2, 1, 3)      # llama3:820-822
    vv = llama_oracle.repeat_kv(vc[:, :Tk].double(), n_rep).permute(0, 2, 1, 3)
    s = torch.matmul(q.double().permute(0, 2, 1, 3), kk.transpose(-1, -2)) / math.sqrt(hd)
    ref = torch.matmul(torch.softmax(s, dim=-1), vv).permute(0, 2, 1, 3)
    lse_ref = torch.logsumexp(s, dim=-1)
    o, lse = ops.attn_decode(q.to(DEV), kc.to(DEV), vc.to(DEV), Tk, scale, want_lse=True)
    torch.cuda.synchronize()
    assert torch.isfinite(o.float()).all()
    assert max_rel_err(o, ref) < tol(dtype) and max_rel_err(lse, lse_ref) < tol(dtype)


@pytest.mark.parametrize("tag", ["one_token", "chunk_nomask"])
def test_block_decode_fp32_against_golden_from_notebook_source(golden_dir, tag):
    """spb_block_decode (norm -> projections -> rotary at `position` -> in-place append -> attention over the cache -> output
    projection) against the notebook's own attention cell run with cache=(k, v), position=Tc."""
    import solvingpapers_b200 as sp
    from solvingpapers_b200 import _lib as L
    g = dict(np.load(os.path.join(golden_dir, f"llama_decode_{tag}.npz")))
    B, Tc, Tq, C, H, Hkv, use_mask = (int(v) for v in g["meta"])
    hd = C // H
    cache = sp.KVCache.from_tensors(T_(g["cache_k"]).to(DEV), T_(g["cache_v"]).to(DEV), extra=Tq + 3)
    cfg = sp.BlockConfig(B, Tq, C, H, Hkv, hd, dtype=torch.float32, norm_kind=L.NORM_RMS, rope=True, causal=False)
    dec = sp.DecodeBlock(cfg, cache.capacity)
    cos, sin = llama_oracle.precompute_freqs_cis(hd, Tc + Tq)
    p = {"wq": T_(g["wq"]).to(DEV), "wk": T_(g["wk"]).to(DEV), "wv": T_(g["wv"]).to(DEV), "wo": T_(g["wo"]).to(DEV),
         "norm_w": T_(g["norm_w"]).to(DEV), "rope": sp.llama_rope_table(cos, sin).to(DEV)}
    y = dec.step(T_(g["x1"]).to(DEV), p, cache)
    torch.cuda.synchronize()
    assert cache.length == Tc + Tq
    assert max_rel_err(y, T_(g["y64"])) < TOL_F32
    assert max_rel_err(cache.k, T_(g["k64"])) < TOL_F32 and max_rel_err(cache.v, T_(g["v64"])) < TOL_F32


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_llama_attention_with_cache_reads_like_the_notebook(dtype):
    """The drop-in's cache branch with the notebook's calling convention: prefill with cache=None, then token-by-token
    decode with cache=new_cache, position=len.  Property the domain offers: decoding token t with the cache of tokens < t gives
    row t of the causal prefill over the whole sequence - checked against the oracle's full causal run."""
    from solvingpapers_b200 import llama
    rng = np.random.default_rng(4)
    B, T, C, H, Hkv, T0 = 2, 40, 256, 4, 2, 33
    hd = C // H
    rd = lambda t: t.to(dtype).float()
    bp = llama_oracle.init_block_params(rng, C, H, Hkv)
    a = {k: rd(v) for k, v in bp["attention"].items()}
    x = rd(T_(rng.standard_normal((B, T, C)).astype(np.float32)))
    fc = llama_oracle.precompute_freqs_cis(hd, T)
    ref, (kref, vref) = llama_oracle.attention(a, x, llama_oracle.make_mask(T), fc, H, Hkv)   # full causal run on the host
    p = {k: v.to(dtype).to(DEV) for k, v in a.items()}
    xd = x.to(dtype).to(DEV)
    freqs_cis = llama.precompute_freqs_cis(hd, T)
    mask = torch.zeros(1, 1, T, T, device=DEV)
    out0, cache = llama.attention(p, xd[:, :T0], mask, freqs_cis, H, Hkv)                      # prefill (cache=None)
    outs = [out0]
    for t in range(T0, T):                                                                     # generate-style loop
        o, cache = llama.attention(p, xd[:, t:t + 1], mask, freqs_cis, H, Hkv, cache=cache, position=t)
        assert cache[0].shape == (B, t + 1, Hkv, hd)
        outs.append(o)
    got = torch.cat(outs, dim=1)
    torch.cuda.synchronize()
    assert max_rel_err(got, ref) < tol(dtype)
    assert max_rel_err(cache[0], kref) < tol(dtype) and max_rel_err(cache[1], vref) < tol(dtype)
    # the same with an in-place KVCache: no copy of the old rows per step
    from solvingpapers_b200 import KVCache
    kv = KVCache.from_tensors(*llama.attention(p, xd[:, :T0], mask, freqs_cis, H, Hkv)[1], extra=T - T0)
    for t in range(T0, T):
        o, (k_view, _) = llama.attention(p, xd[:, t:t + 1], mask, freqs_cis, H, Hkv, cache=kv, position=t)
        assert k_view.data_ptr() == kv.kbuf.data_ptr() and k_view.shape[1] == t + 1
        assert max_rel_err(o, ref[:, t:t + 1]) < tol(dtype) * (ref.abs().max() / ref[:, t:t + 1].abs().max()).item()
    # what the notebook's own code cannot compute raises instead of guessing
    with pytest.raises(ValueError):
        llama.attention(p, xd[:, :2], mask, freqs_cis, H, Hkv, cache=(cache[0][:, :5], cache[1][:, :5]), position=5)   # mask with 2 new tokens
    with pytest.raises(ValueError):
        llama.attention(p, xd[:, :1], None, freqs_cis, H, Hkv, cache=(cache[0][:, :5], cache[1][:, :5]), position=7)   # position != cache length
    with pytest.raises(ValueError):
        llama.attention(p, xd[:, :1], None, freqs_cis, H, Hkv, position=3)


def test_block_decode_c3_shape_bf16_against_oracle():
    """One decode step at the C3 model shape (d4096, 32 q / 8 kv heads x 128) with a 1023-token cache, B = 4: bf16 against
    the oracle's cache branch in fp32 on the rounded inputs."""
    import solvingpapers_b200 as sp
    from solvingpapers_b200 import _lib as L
    B, Tc, C, H, Hkv = 4, 1023, 4096, 32, 8
    hd = C // H
    rng = np.random.default_rng(6)
    rd = lambda t: t.to(torch.bfloat16).float()
    bp = llama_oracle.init_block_params(rng, C, H, Hkv)
    a = {k: rd(v) for k, v in bp["attention"].items()}
    nw = rd(bp["attention_norm"])
    x = rd(T_(rng.standard_normal((B, 1, C)).astype(np.float32)))
    ck = rd(T_(rng.standard_normal((B, Tc, Hkv, hd)).astype(np.float32)))
    cv = rd(T_(rng.standard_normal((B, Tc, Hkv, hd)).astype(np.float32)))
    fc = llama_oracle.precompute_freqs_cis(hd, Tc + 1)
    ref, (kref, vref) = llama_oracle.attention(a, llama_oracle.rms_norm(x, nw), None, fc, H, Hkv, cache=(ck, cv), position=Tc)
    ref = x + ref                                                                               # llama3:885
    dev = lambda t: t.to(torch.bfloat16).to(DEV)
    cache = sp.KVCache.from_tensors(dev(ck), dev(cv), extra=1)
    cfg = sp.BlockConfig(B, 1, C, H, Hkv, hd, dtype=torch.bfloat16, norm_kind=L.NORM_RMS, rope=True, causal=False, residual=True)
    dec = sp.DecodeBlock(cfg, cache.capacity)
    p = {"wqkv": dev(torch.cat([a["wq"], a["wk"], a["wv"]], dim=1)), "wo": dev(a["wo"]), "norm_w": dev(nw),
         "rope": sp.llama_rope_table(*fc).to(DEV)}
    y = dec.step(dev(x), p, cache)
    torch.cuda.synchronize()
    assert max_rel_err(y, ref) < TOL_BF16
    assert max_rel_err(cache.k[:, -1:], kref[:, -1:]) < TOL_BF16 and max_rel_err(cache.v[:, -1:], vref[:, -1:]) < TOL_BF16
    assert torch.equal(cache.k[:, :-1].cpu(), dev(ck).cpu())   # old rows untouched


@pytest.mark.parametrize("dtype,B,C,H,Hkv,Tc,steps", [
    (torch.bfloat16, 4, 1024, 8, 2, 300, 9),     # crosses a 64-key tile boundary (300 -> 309 .. 320 would; 9 steps stay inside one more tile)
    (torch.bfloat16, 2, 512, 4, 4, 60, 8),       # 60 -> 68: the tile count of the valid range changes under the recorded grid
    (torch.float32, 16, 256, 4, 2, 22, 6),       # the notebook's dims (C1-decode), fp32 kernels
])
def test_recorded_decode_step_replays_like_eager_steps(dtype, B, C, H, Hkv, Tc, steps):
    """DecodeBlock.capture (spb_block_decode_dev in a CUDA graph: length in device memory, advanced by the step) against the
    same number of eager spb_block_decode steps on a twin cache: outputs per token, the appended rows and the final length."""
    import solvingpapers_b200 as sp
    from solvingpapers_b200 import _lib as L
    hd = C // H
    rng = np.random.default_rng(11)
    bp = llama_oracle.init_block_params(rng, C, H, Hkv)
    a = bp["attention"]
    dev = lambda t: t.to(dtype).to(DEV)
    Tmax = Tc + steps + 3
    fc = llama_oracle.precompute_freqs_cis(hd, Tmax)
    p = {"wqkv": dev(torch.cat([a["wq"], a["wk"], a["wv"]], dim=1)), "wo": dev(a["wo"]), "norm_w": dev(bp["attention_norm"]),
         "rope": sp.llama_rope_table(*fc).to(DEV)}
    ck = dev(T_(rng.standard_normal((B, Tc, Hkv, hd)).astype(np.float32)))
    cv = dev(T_(rng.standard_normal((B, Tc, Hkv, hd)).astype(np.float32)))
    xs = [dev(T_(rng.standard_normal((B, 1, C)).astype(np.float32))) for _ in range(steps)]
    cfg = sp.BlockConfig(B, 1, C, H, Hkv, hd, dtype=dtype, norm_kind=L.NORM_RMS, rope=True, causal=False, residual=True)
    # eager
    cache_e = sp.KVCache.from_tensors(ck, cv, extra=Tmax - Tc)
    dec_e = sp.DecodeBlock(cfg, Tmax)
    ys_e = [dec_e.step(x, p, cache_e).clone() for x in xs]
    # recorded
    cache_g = sp.KVCache.from_tensors(ck, cv, extra=Tmax - Tc)
    dec_g = sp.DecodeBlock(cfg, Tmax)
    xbuf = torch.empty_like(xs[0])
    cap = dec_g.capture(xbuf, p, cache_g)
    assert cache_g.length == Tc                      # recording appends nothing that stays
    ys_g = []
    for x in xs:
        xbuf.copy_(x)
        ys_g.append(cap.replay().clone())
    torch.cuda.synchronize()
    assert cache_g.length == cache_e.length == Tc + steps
    assert int(cache_g.length_dev().item()) == Tc + steps
    tol_same = 1e-5 if dtype == torch.float32 else 1e-2   # same kernels; the key splits (and so the fp32 merge order) differ
    for i, (ye, yg) in enumerate(zip(ys_e, ys_g)):
        assert max_rel_err(yg, ye) < tol_same, i
    assert max_rel_err(cache_g.k, cache_e.k) < tol_same and max_rel_err(cache_g.v, cache_e.v) < tol_same
    assert torch.equal(cache_g.k[:, :Tc].cpu(), ck.cpu())
    # an eager step in between moves the length on the host only: the next replay re-synchronises the device copy
    x_extra = xs[0]
    y_e = dec_e.step(x_extra, p, cache_e)
    y_mixed = dec_g.step(x_extra, p, cache_g)
    assert max_rel_err(y_mixed, y_e) < tol_same
    xbuf.copy_(xs[1])
    y_g2 = cap.replay().clone()
    y_e2 = dec_e.step(xs[1], p, cache_e)
    torch.cuda.synchronize()
    assert cache_g.length == cache_e.length == Tc + steps + 2
    assert max_rel_err(y_g2, y_e2) < tol_same
    with pytest.raises(L.SpbError, match="exceeds the cache capacity"):
        for _ in range(3):
            cap.replay()


def test_decode_rejects_what_it_cannot_do():
    import solvingpapers_b200 as sp
    from solvingpapers_b200 import _lib as L
    with pytest.raises(L.SpbError):   # causal mask over several new tokens: no counterpart in the notebook
        sp.DecodeBlock(sp.BlockConfig(1, 2, 128, 2, 2, 64, dtype=torch.float32, causal=True), 16).step(
            torch.zeros(1, 2, 128, device=DEV), {"wq": torch.zeros(128, 128, device=DEV), "wk": torch.zeros(128, 128, device=DEV),
                                                 "wv": torch.zeros(128, 128, device=DEV), "wo": torch.zeros(128, 128, device=DEV),
                                                 "norm_w": torch.ones(128, device=DEV), "rope": torch.zeros(2, 32, 4, device=DEV)},
            sp.KVCache(1, 16, 2, 64, dtype=torch.float32))
    with pytest.raises(L.SpbError):   # more rows than the small-M path takes
        sp.DecodeBlock(sp.BlockConfig(65, 1, 128, 2, 2, 64, dtype=torch.float32, causal=False), 16)
