"""GPU timing (not a test; launch under torchrun with N ranks): the library's NCCL all-reduce of the C3 weight-gradient buckets
alone on the GPUs - 100 MB (dWqkv), 64 MB (dWo), both - so that the exposed/overlapped share in bench.py can be judged."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from solvingpapers_b200 import dist as spd

rank, local, world = spd.env_rank_world()
torch.cuda.set_device(local)
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
comm = spd.Communicator(rank, world)
bufs = {"dwqkv 100MB": [torch.ones(4096 * 6144, device="cuda")], "dwo 64MB": [torch.ones(4096 * 4096, device="cuda")],
        "both 164MB": [torch.ones(4096 * 6144, device="cuda"), torch.ones(4096 * 4096, device="cuda")]}
for name, ts in bufs.items():
    for _ in range(5):
        comm.allreduce_(ts)
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        comm.allreduce_(ts)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    nbytes = sum(t.numel() * 4 for t in ts)
    if rank == 0:
        print(f"world {world} allreduce {name}: {ms:.3f} ms  algbw {nbytes/ms/1e6:.0f} GB/s  busbw {nbytes/ms/1e6*2*(world-1)/world:.0f} GB/s")
dist.barrier()
comm.close()
dist.destroy_process_group()
