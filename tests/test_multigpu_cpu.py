"""N>1 host logic on CPU: world_size-2 gloo process group (no GPU).  Covers the VA sharding, the per-owner trace
split (each rank preprocesses only its own sub-trace; the union equals the single-GPU result per va_block), the IPC
handle exchange and the max-over-ranks reduction that bench.py uses."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import faultlib as F
import uvmtier_testlib as T
from uvmtier import sharding as SH, tracegen as TG


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, n_blocks, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # 1) IPC handle exchange: every rank ends up with every handle, in rank order
        handles = SH.exchange_handles(dist, bytes([rank]) * 64)
        assert handles == [bytes([r]) * 64 for r in range(world)]
        # 2) max over ranks
        assert SH.max_over_ranks(dist, [1.0 + rank, 5.0 - rank]) == [float(world), 5.0]
        # 3) every rank preprocesses its own sub-trace with the CPU checker; results go back to the parent
        pk, tbl, lay = TG.fault_trace(4000, 3, "uniform", 1, total_va_bytes=n_blocks << 21)
        mine = SH.split_trace_by_owner(pk, lay["va_base"], n_blocks, world)[rank]
        gp = TG.gpu_params(256)
        r = F.run_fetch_preprocess(T.load_oracle(), "oracle", np.ascontiguousarray(mine), gp, tbl)
        first, outer = SH.block_range(n_blocks, world, rank)
        blk = ((r["entries"]["fault_address"] - np.uint64(lay["va_base"])) >> np.uint64(21)).astype(np.int64)
        assert ((blk >= first) & (blk < outer)).all(), "a rank saw a fault of a va_block it does not own"
        ret[rank] = (len(mine), int(r["batch"]["num_cached_faults"].sum()))
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_world_size_2_gloo_sharding_and_control_plane():
    world, n_blocks = 2, 64
    with mp.Manager() as m:
        ret = m.dict()
        mp.spawn(_worker, args=(world, _free_port(), n_blocks, ret), nprocs=world, join=True)
        sizes = [ret[r] for r in range(world)]
    assert sum(s[0] for s in sizes) == 4000 and all(s[0] == s[1] for s in sizes)


def test_block_ranges_partition_the_va_space():
    for n_blocks, world in ((32768, 8), (32768, 3), (7, 4), (64, 2)):
        seen = np.zeros(n_blocks, int)
        for r in range(world):
            a, b = SH.block_range(n_blocks, world, r)
            seen[a:b] += 1
            if b > a:
                assert (SH.owner_of_block(np.arange(a, b), n_blocks, world) == r).all()
        assert (seen == 1).all()
