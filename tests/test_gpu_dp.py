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


def test_registered_gradient_buffers_size1_communicator():
    """spb_comm_alloc / spb_comm_free: memory from NCCL's allocator registered with the communicator, usable as an ordinary
    tensor (the weight-gradient GEMMs write it, the all-reduce works on it in place)."""
    from solvingpapers_b200 import dist as spd, ops
    comm = spd.Communicator(0, 1)
    try:
        g = comm.alloc((256, 384))
        assert g.shape == (256, 384) and g.dtype == torch.float32 and g.is_cuda and g.data_ptr() % 4096 == 0
        assert comm.buffer_kind(g) in ("registered", "symmetric window", "unregistered")
        a = torch.randn(512, 256, device="cuda").bfloat16()
        b = torch.randn(512, 384, device="cuda").bfloat16()
        ops.gemm(a, b, transA=True, out=g)                      # a weight-gradient GEMM straight into the registered buffer
        ref = a.float().t() @ b.float()
        comm.allreduce_([g])                                     # size-1 SUM: unchanged
        torch.cuda.synchronize()
        assert max_rel_err(g, ref) < 1e-5
        assert comm.buffer_kind(torch.empty(4, device="cuda")) == "unregistered"
    finally:
        comm.close()


def _two_rank_worker(rank, world, port, out):
    import os
    import torch.distributed as tdist
    import solvingpapers_b200 as sp
    from solvingpapers_b200 import dist as spd
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    tdist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    comm = spd.Communicator(rank, world)
    B, T, C, H, Hkv = 4, 128, 256, 4, 2
    hd = C // H
    gen = torch.Generator().manual_seed(0)  # same seed on both ranks: replicated weights, same global batch
    cos, sin = llama_oracle.precompute_freqs_cis(hd, T)
    dev = lambda t: t.bfloat16().cuda()
    p = {"wqkv": dev(torch.randn(C, (H + 2 * Hkv) * hd, generator=gen) / 16), "wo": dev(torch.randn(H * hd, C, generator=gen) / 16),
         "norm_w": dev(torch.randn(C, generator=gen)), "rope": sp.llama_rope_table(cos, sin).cuda()}
    x, dy = dev(torch.randn(B, T, C, generator=gen)), dev(torch.randn(B, T, C, generator=gen))
    per = B // world
    blk = sp.AttentionBlock(sp.BlockConfig(per, T, C, H, Hkv, hd, dtype=torch.bfloat16), comm=comm.handle)
    blk.forward(spd.shard_batch(x, world, rank).contiguous(), p)
    g = blk.backward(spd.shard_batch(dy, world, rank).contiguous())
    torch.cuda.synchronize()
    if rank == 0:
        full = sp.AttentionBlock(sp.BlockConfig(B, T, C, H, Hkv, hd, dtype=torch.bfloat16))
        full.forward(x, p)
        ref = full.backward(dy)
        errs = {k: max_rel_err(g[k], ref[k]) for k in ("dwqkv", "dwo", "dnorm_w")}
        errs["dx"] = max_rel_err(g["dx"], ref["dx"][:per])
    # the feed-forward sub-block's gradient all-reduce: sharded tokens, summed dw13 / dw2 / dnorm
    Hh = 4 * C
    pf = {"w13": dev(torch.randn(C, 2 * Hh, generator=gen) / 16), "w2": dev(torch.randn(Hh, C, generator=gen) / 32),
          "norm_w": dev(torch.randn(C, generator=gen))}
    xs, dys = spd.shard_batch(x, world, rank).reshape(-1, C).contiguous(), spd.shard_batch(dy, world, rank).reshape(-1, C).contiguous()
    ff = sp.FeedForward(sp.FfnConfig(N=per * T, C=C, H=Hh, residual=True), comm=comm.handle)
    ff.forward(xs, pf)
    gf = ff.backward(dys)
    torch.cuda.synchronize()
    if rank == 0:
        full = sp.FeedForward(sp.FfnConfig(N=B * T, C=C, H=Hh, residual=True))
        full.forward(x.reshape(-1, C).contiguous(), pf)
        ref = full.backward(dy.reshape(-1, C).contiguous())
        errs.update({"ffn_" + k: max_rel_err(gf[k], ref[k]) for k in ("dw13", "dw2", "dnorm_w")})
        errs["ffn_dx"] = max_rel_err(gf["dx"], ref["dx"][:per * T])
        torch.save(errs, out)
    tdist.barrier()
    comm.close()
    tdist.destroy_process_group()


def test_two_gpu_allreduced_gradients_equal_full_batch_gradients(tmp_path):
    """"8-GPU allreduced dW == 1-GPU full-batch dW" - here with the GPUs this box has (needs >= 2)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs at least 2 GPUs (run under gpurun --gpus 2)")
    import socket
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    out = str(tmp_path / "errs.pt")
    mp.spawn(_two_rank_worker, args=(2, port, out), nprocs=2, join=True)
    errs = torch.load(out)
    assert max(errs.values()) < 2e-2, errs
