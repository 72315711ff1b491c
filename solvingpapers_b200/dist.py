"""Batch-sharded data parallelism for the block.  One process per GPU.  The batch is split contiguously
along B; weights and the rotary table are replicated; the only exchange step is the all-reduce (SUM, fp32) of the
weight gradients, issued from inside libspb200.so (spb_block_bwd with a communicator) via NCCL over NVLink/NVSwitch.
torch.distributed is used only to hand the 128-byte NCCL unique id from rank 0 to the other ranks."""
import ctypes as C
import os

import torch

from . import _lib as L


def shard_bounds(global_batch: int, world_size: int, rank: int):
    """Rank r owns x[lo:hi] of the global batch (contiguous, equal shards)."""
    if world_size < 1 or not (0 <= rank < world_size):
        raise ValueError(f"rank {rank} of {world_size}")
    if global_batch % world_size != 0:
        raise ValueError(f"global batch {global_batch} is not divisible by world size {world_size}")
    per = global_batch // world_size
    return rank * per, (rank + 1) * per


def shard_batch(t: torch.Tensor, world_size: int, rank: int):
    lo, hi = shard_bounds(t.shape[0], world_size, rank)
    return t[lo:hi]


class Communicator:
    """NCCL communicator owned by the C library (dlopen'ed libnccl; inside a torch process that is torch's own copy)."""

    def __init__(self, rank: int, world_size: int, group=None):
        import torch.distributed as dist
        self.lib = L.load()
        self.rank, self.world_size = rank, world_size
        idbuf = (C.c_ubyte * 128)()
        if rank == 0:
            L.check(self.lib.spb_comm_unique_id(C.cast(idbuf, C.c_void_p)), "spb_comm_unique_id")
        t = torch.tensor(list(bytes(idbuf)), dtype=torch.uint8)
        if world_size > 1:
            dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl" else torch.device("cpu")
            t = t.to(dev)
            dist.broadcast(t, src=0, group=group)
            t = t.cpu()
        raw = bytes(t.tolist())
        buf = (C.c_ubyte * 128).from_buffer_copy(raw)
        h = C.c_void_p()
        L.check(self.lib.spb_comm_init_rank(C.cast(buf, C.c_void_p), world_size, rank, C.byref(h)), "spb_comm_init_rank")
        self.handle = h.value

    def allreduce_(self, tensors):
        """In-place SUM over ranks of fp32 CUDA tensors (one NCCL group)."""
        n = len(tensors)
        for t in tensors:
            if t.dtype != torch.float32 or not t.is_contiguous() or t.device.type != "cuda":
                raise L.SpbError("allreduce_ takes contiguous fp32 CUDA tensors")
        ptrs = (C.c_void_p * n)(*[t.data_ptr() for t in tensors])
        counts = (C.c_size_t * n)(*[t.numel() for t in tensors])
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        L.check(self.lib.spb_comm_allreduce_f32(C.c_void_p(self.handle), ptrs, counts, n, st), "spb_comm_allreduce_f32")

    def alloc(self, shape, dtype=torch.float32):
        """A tensor in memory that the communicator's all-reduce works on in place (spb_comm_alloc: NCCL allocator + window /
        buffer registration).  Collective: every rank allocates the same sizes in the same order.  Lives until close()."""
        n = 1
        for d in shape:
            n *= int(d)
        esz = torch.empty((), dtype=dtype).element_size()
        ptr = C.c_void_p()
        L.check(self.lib.spb_comm_alloc(C.c_void_p(self.handle), n * esz, C.byref(ptr)), "spb_comm_alloc")
        typestr = {torch.float32: "<f4", torch.bfloat16: "<u2", torch.uint8: "|u1"}[dtype]

        class _Raw:
            __cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr.value, False), "version": 2}

        t = torch.as_tensor(_Raw(), device=torch.device("cuda", torch.cuda.current_device()))
        if dtype == torch.bfloat16:
            t = t.view(torch.bfloat16)
        t = t.view(*shape)
        self._bufs = getattr(self, "_bufs", [])
        self._bufs.append(t)
        return t

    def buffer_kind(self, t):
        return {0: "unregistered", 1: "registered", 2: "symmetric window"}[self.lib.spb_comm_buffer_kind(C.c_void_p(self.handle), C.c_void_p(t.data_ptr()))]

    def close(self):
        if getattr(self, "handle", None):
            self.lib.spb_comm_destroy(C.c_void_p(self.handle))
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def env_rank_world():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
