"""Multi-GPU host logic: the VA space is partitioned across the GPUs of one box, one process per GPU.

No data-path collective exists on this path: every GPU owns a contiguous range of va_blocks, its fault ring and its
state table.  torch.distributed is only the control plane: a barrier, the max over ranks of a timing, and the
exchange of the 64-byte CUDA IPC handles that let a GPU address a peer's HBM arena for point-to-point page moves.
"""
import numpy as np


def block_range(n_blocks, world, rank):
    """Contiguous range of va_blocks owned by `rank`: block b belongs to GPU b * world // n_blocks."""
    first = (rank * n_blocks + world - 1) // world
    outer = ((rank + 1) * n_blocks + world - 1) // world
    return first, outer


def owner_of_block(block, n_blocks, world):
    return (np.asarray(block, np.int64) * world) // n_blocks


def owner_of_address(addr, va_base, n_blocks, world):
    return owner_of_block((np.asarray(addr, np.uint64) - np.uint64(va_base)) >> np.uint64(21), n_blocks, world)


def split_trace_by_owner(packets, va_base, n_blocks, world):
    """Splits a global fault trace into per-GPU sub-traces by the owner of the faulting va_block, keeping ring
    order inside each (the trace generator tags entries with a GPU by address; cmp_gpu, :713-719)."""
    w = packets.astype(np.uint64)
    addr = ((w[:, 3] << np.uint64(20)) | (w[:, 2] >> np.uint64(12))) << np.uint64(12)
    own = owner_of_address(addr, va_base, n_blocks, world)
    return [packets[own == r] for r in range(world)]


def exchange_handles(dist, handle_bytes):
    """all_gather of one 64-byte IPC handle per rank (CPU tensors: works on gloo and nccl-with-cpu alike)."""
    import torch
    mine = torch.tensor(list(handle_bytes), dtype=torch.uint8)
    if dist.get_backend() == "nccl":
        mine = mine.cuda()
    out = [torch.zeros_like(mine) for _ in range(dist.get_world_size())]
    dist.all_gather(out, mine)
    return [bytes(t.cpu().tolist()) for t in out]


def max_over_ranks(dist, values):
    import torch
    t = torch.tensor(values, dtype=torch.float64)
    if dist.get_backend() == "nccl":
        t = t.cuda()
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [float(x) for x in t.tolist()]
