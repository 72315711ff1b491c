"""CPU, world_size 2, gloo: the data-parallel contract of SURVEY 8(e) - batch sharded contiguously, weights replicated,
weight gradients SUMMED over ranks equal the single-process full-batch gradients; activations of a shard do not depend
on the other shard.  The per-rank arithmetic here is the oracle (no GPU in this container); the same sharding helpers
(solvingpapers_b200.dist) drive the NCCL path on the GPU box (tests/test_gpu_dp.py)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as tdist
import torch.multiprocessing as mp

from oracle import llama_oracle, max_rel_err


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out):
    from solvingpapers_b200 import dist as spd
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    tdist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(0)  # same seed on every rank: replicated weights, same global batch
    B, T, C, H, Hkv = 4, 16, 32, 4, 2
    params = llama_oracle.init_block_params(rng, C, H, Hkv, dtype=torch.float64)
    x = torch.from_numpy(rng.standard_normal((B, T, C)))
    dy = torch.from_numpy(rng.standard_normal((B, T, C)))
    xs, dys = spd.shard_batch(x, world, rank), spd.shard_batch(dy, world, rank)
    r = llama_oracle.block_fwd_bwd(params, xs, dys, H, Hkv, dtype=torch.float64)
    grads = [r[k].clone() for k in ("dwq", "dwk", "dwv", "dwo", "dnorm")]
    for g in grads:
        tdist.all_reduce(g, op=tdist.ReduceOp.SUM)  # SUM, not mean (dy is defined globally)
    ys = [torch.zeros_like(r["y"]) for _ in range(world)]
    tdist.all_gather(ys, r["y"])
    if rank == 0:
        full = llama_oracle.block_fwd_bwd(params, x, dy, H, Hkv, dtype=torch.float64)
        errs = [max_rel_err(g, full[k]) for g, k in zip(grads, ("dwq", "dwk", "dwv", "dwo", "dnorm"))]
        errs.append(max_rel_err(torch.cat(ys), full["y"]))
        torch.save(errs, out)
    tdist.destroy_process_group()


def test_sharded_gradients_sum_to_full_batch_gradients(tmp_path):
    out = str(tmp_path / "errs.pt")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    errs = torch.load(out)
    assert max(errs) < 1e-12, errs
