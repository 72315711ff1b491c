#!/usr/bin/env python
"""Host ceiling of bench.py's e2e leg, alone: N ranks copy 2 x 67 MB pinned host -> device and 2 x 67 MB device -> pinned
host per iteration at the same time, no kernels (the bytes one C3 e2e step moves per rank).

    python tools/diag_pcie.py                                   # one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/diag_pcie.py
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch
    import torch.distributed as dist
    import bench
    rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    x = torch.empty(4, 2048, 4096, dtype=torch.bfloat16, device=dev)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    aff = bench.bind_near_gpu(torch, local)
    r = bench.measure_host_ceiling(torch, dist if world > 1 else None, x, world, barrier, iters=16)
    r["host_affinity_rank0"] = aff
    if rank == 0:
        print(json.dumps(r))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
