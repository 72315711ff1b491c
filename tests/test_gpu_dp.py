"""GPU: the NCCL path of the C library.  On a 1-GPU box this runs with a communicator of size 1 (SURVEY section 4:
"world=1 on CI, world in {2,4,8} on the GPU box" - the multi-rank run is bench.py under torchrun); it proves libnccl is
found by dlopen, the communicator is created from the C ABI, the all-reduce is issued from inside spb_block_bwd on the
side stream and joined, and gradients are unchanged by a size-1 SUM."""
import numpy as np
import pytest
import torch

from oracle import llama_oracle, max_rel_err

pytestmark = pytest.mark.gpu


def test_block_bwd_with_size1_communicator_matches_no_communicator():
    import solvingpapers_b200 as sp
    from solvingpapers_b200 import dist as spd
    comm = spd.Communicator(0, 1)
    try:
        B, T, C, H, Hkv = 2, 64, 128, 4, 2
        hd = C // H
        gen = torch.Generator().manual_seed(0)
        dev = lambda t: t.cuda()
        cos, sin = llama_oracle.precompute_freqs_cis(hd, T)
        p = {"wqkv": dev(torch.randn(C, (H + 2 * Hkv) * hd, generator=gen) / 11), "wo": dev(torch.randn(H * hd, C, generator=gen) / 11),
             "norm_w": dev(torch.randn(C, generator=gen)), "rope": sp.llama_rope_table(cos, sin).cuda()}
        x, dy = dev(torch.randn(B, T, C, generator=gen)), dev(torch.randn(B, T, C, generator=gen))
        cfg = sp.BlockConfig(B, T, C, H, Hkv, hd, dtype=torch.float32)
        g0 = sp.AttentionBlock(cfg)
        g0.forward(x, p)
        ref = {k: v.clone() for k, v in g0.backward(dy).items()}
        b1 = sp.AttentionBlock(cfg, comm=comm.handle)
        b1.forward(x, p)
        got = b1.backward(dy)
        torch.cuda.synchronize()
        for k in ref:
            assert max_rel_err(got[k], ref[k]) == 0.0, k
        t = torch.arange(10, dtype=torch.float32, device="cuda")
        comm.allreduce_([t])
        torch.cuda.synchronize()
        assert torch.equal(t.cpu(), torch.arange(10, dtype=torch.float32))
    finally:
        comm.close()
