"""Synthetic replayable-fault traces, block state and page lists for the harness.

All randomness comes from the reference's own test generator (uvm_test_rng, UVM/uvm_test_rng.c:37-62,
exposed by the library as uvmtier_test_rng_fill): one u32 stream per seed, shaped with numpy.
Packets are raw 32-byte NVC369_BUF_ENTRY records (UVM/clc369.h:33-66).
"""
import ctypes

import numpy as np

from . import lib, P

# HW access type encodings, hwref/volta/gv100/dev_fault.h:246-251
HW_VIRT_READ, HW_VIRT_WRITE, HW_VIRT_ATOMIC_STRONG, HW_VIRT_PREFETCH, HW_VIRT_ATOMIC_WEAK = 0, 1, 2, 3, 4
HW_FAULT_TYPE_PTE = 2            # hwref/blackwell/gb100/dev_fault.h:272
HW_FAULT_TYPE_PDE_SIZE = 1       # fatal class (UVM_FAULT_TYPE_PDE_SIZE)
ENG_ID_GRAPHICS = 384            # hwref/blackwell/gb100/dev_fault.h:29
APERTURE_VID = 10                # uvm_aperture_t UVM_APERTURE_VID (UVM/uvm_hal_types.h:33-51)
VA_BASE = 0x7F0000000000
# GPC client ids whose uTLB exists with 10 TPCs per GPC: T1_0..T1_19 (dev_fault.h:287-333)
T1_CLIENTS = np.array(list(range(0x00, 0x08)) + list(range(0x21, 0x29)) + list(range(0x35, 0x39)), np.uint32)

INSTANCE_PTR_DT = np.dtype([("instance_ptr", "<u8"), ("aperture", "u1"), ("ve_id", "u1"), ("va_space", "<u2"),
                            ("gpu_id", "<u2"), ("sort_rank", "<u2")])
assert INSTANCE_PTR_DT.itemsize == 16
GPU_PARAMS_DT = np.dtype([("max_batch_size", "<u4"), ("utlb_per_gpc_count", "<u4"), ("gpc_count", "<u4"),
                          ("va_bits", "<u4"), ("coalesce", "<u4"), ("reserved", "<u4", 3)])
assert GPU_PARAMS_DT.itemsize == 32


def rng_stream(seed, n):
    out = np.empty(n, np.uint32)
    lib().uvmtier_test_rng_fill(ctypes.c_uint32(seed & 0xFFFFFFFF), ctypes.c_size_t(n), P(out))
    return out


class Stream:
    """Sequential consumer of the KISS u32 stream."""

    def __init__(self, seed, n):
        self.buf = rng_stream(seed, n)
        self.pos = 0

    def take(self, n):
        assert self.pos + n <= len(self.buf), "rng stream exhausted"
        v = self.buf[self.pos:self.pos + n]
        self.pos += n
        return v

    def below(self, n, bound):
        return (self.take(n).astype(np.uint64) % np.uint64(bound)).astype(np.int64)

    def unit(self, n):
        return self.take(n).astype(np.float64) / 4294967296.0


def gpu_params(max_batch_size=256, tpc_per_gpc=10, gpc_count=8, va_bits=57, coalesce=1):
    """uvm_parent_gpu_t fields the fault path reads (Blackwell: utlb_per_gpc = 2*TPC/GPC + 1,
    UVM/uvm_blackwell_fault_buffer.h:85-90)."""
    g = np.zeros(1, GPU_PARAMS_DT)
    g["max_batch_size"] = max_batch_size
    g["utlb_per_gpc_count"] = 2 * tpc_per_gpc + 1
    g["gpc_count"] = gpc_count
    g["va_bits"] = va_bits
    g["coalesce"] = coalesce
    return g


def instance_ptr_table(n_spaces, gpu_id=1):
    """One instance pointer (channel) per VA space, all on `gpu_id`."""
    t = np.zeros(n_spaces, INSTANCE_PTR_DT)
    t["instance_ptr"] = (0x1000 + np.arange(n_spaces, dtype=np.uint64) * 7) << np.uint64(12)
    t["aperture"] = APERTURE_VID
    t["ve_id"] = 0xFF
    t["va_space"] = np.arange(n_spaces)
    t["gpu_id"] = gpu_id
    t["sort_rank"] = np.arange(n_spaces)     # (va_space, gpu) pairs are already in order
    return t


def encode_packets(inst_ptr, address, access_hw, gpc, client, timestamp, veid=None, fault_type=None, valid=None,
                   inst_aperture_hw=None):
    """Vectorised NVC369_BUF_ENTRY encoder (UVM/clc369.h:33-66). address/inst_ptr in bytes (4K aligned)."""
    n = len(address)
    w = np.zeros((n, 8), np.uint32)
    ip = np.asarray(inst_ptr, np.uint64) >> np.uint64(12)
    ap = np.zeros(n, np.uint32) if inst_aperture_hw is None else np.asarray(inst_aperture_hw, np.uint32)
    w[:, 0] = ((ip & np.uint64(0xFFFFF)) << np.uint64(12)).astype(np.uint32) | (ap << 8)
    w[:, 1] = (ip >> np.uint64(20)).astype(np.uint32)
    pg = (np.asarray(address, np.uint64) >> np.uint64(12)) & np.uint64((1 << 52) - 1)
    w[:, 2] = ((pg & np.uint64(0xFFFFF)) << np.uint64(12)).astype(np.uint32)
    w[:, 3] = (pg >> np.uint64(20)).astype(np.uint32)
    ts = np.asarray(timestamp, np.uint64)
    w[:, 4] = (ts & np.uint64(0xFFFFFFFF)).astype(np.uint32)
    w[:, 5] = (ts >> np.uint64(32)).astype(np.uint32)
    ve = np.zeros(n, np.uint32) if veid is None else np.asarray(veid, np.uint32)
    w[:, 6] = ENG_ID_GRAPHICS + ve
    ft = np.full(n, HW_FAULT_TYPE_PTE, np.uint32) if fault_type is None else np.asarray(fault_type, np.uint32)
    vd = np.ones(n, np.uint32) if valid is None else np.asarray(valid, np.uint32)
    w[:, 7] = (ft | (1 << 7) | (np.asarray(client, np.uint32) << 8) | (np.asarray(access_hw, np.uint32) << 16) |
               (0 << 20) | (np.asarray(gpc, np.uint32) << 24) | (1 << 30) | (vd << 31))
    return np.ascontiguousarray(w)


def fault_trace(n, seed=1, distribution="uniform", n_spaces=4, total_va_bytes=64 << 30, gpc_count=8,
                gpu_id=1):
    """Returns (packets[n,8] u32, table, layout) for one GPU's ring.

    Access mix 60/25/5/8/2 % READ/WRITE/ATOMIC_STRONG/PREFETCH/ATOMIC_WEAK.  The VA of every VA space
    starts at VA_BASE and spans total_va_bytes / n_spaces (2 MB aligned).
    distribution: uniform | stream | hotset | adversarial
    """
    s = Stream(seed, 12 * n + 64)
    blocks_per_space = max(1, (total_va_bytes >> 21) // n_spaces)
    pages_per_space = blocks_per_space * 512
    table = instance_ptr_table(n_spaces, gpu_id)
    space = s.below(n, n_spaces)
    gpc = s.below(n, gpc_count)
    client = T1_CLIENTS[s.below(n, len(T1_CLIENTS))]
    u = s.unit(n)
    access = np.select([u < 0.60, u < 0.85, u < 0.90, u < 0.98], [HW_VIRT_READ, HW_VIRT_WRITE, HW_VIRT_ATOMIC_STRONG,
                                                                 HW_VIRT_PREFETCH], HW_VIRT_ATOMIC_WEAK)
    if distribution == "uniform":
        page = s.below(n, pages_per_space)
    elif distribution == "stream":
        # sequential pages, each touched d in [1,8] times (different uTLBs come from gpc/client draws)
        d = 1 + s.below(n, 8)
        reps = np.repeat(np.arange(n), d)[:n]
        page = reps % pages_per_space
        space = (reps // pages_per_space) % n_spaces
    elif distribution == "hotset":
        hot_blocks = max(1, blocks_per_space // 100)
        hot = s.unit(n) < 0.9
        blk = np.where(hot, s.below(n, hot_blocks) * (blocks_per_space // hot_blocks), s.below(n, blocks_per_space))
        page = blk * 512 + s.below(n, 512)
    elif distribution == "adversarial":
        # few pages, rising/falling access types, alternating uTLBs: stresses the fetch-merge rules
        page = s.below(n, 6) + 512 * s.below(n, 2)
        gpc = s.below(n, 2)
        client = T1_CLIENTS[s.below(n, 3)]
        space = s.below(n, min(2, n_spaces))
        access = s.below(n, 5)
    else:
        raise ValueError(distribution)
    addr = np.uint64(VA_BASE) + (page.astype(np.uint64) << np.uint64(12))
    inst = table["instance_ptr"][space]
    ts = np.uint64(1000) + np.arange(n, dtype=np.uint64) * np.uint64(3)
    pk = encode_packets(inst, addr, access, gpc, client, ts)
    layout = dict(va_base=VA_BASE, blocks_per_space=blocks_per_space, n_spaces=n_spaces, gpu_id=gpu_id)
    return pk, table, layout
