"""Helpers for the non-replayable fault tests: synthetic shadow buffers and runners for the CPU checkers."""
import ctypes

import numpy as np

import faultlib as F
from uvmtier import tracegen as TG
from uvmtier.blockstate import COPY_RUN_DT, SERVICE_GROUP_DT
from uvmtier_testlib import P


def shadow_buffer(n, seed, lay, tbl, fatal_frac=0.03, stale_frac=0.03, outside_frac=0.03, blocks=None):
    """n non-replayable packets: HUB (copy-engine) clients, no PREFETCH accesses (the reference asserts there are none,
    uvm_gpu_non_replayable_faults.c:536-539), REPLAYABLE clear; a few fatal fault types, a few stale channels (their
    instance pointer is in no table row), a few addresses outside every va_range."""
    rng = np.random.default_rng(seed)
    bps, ns = lay["blocks_per_space"], lay["n_spaces"]
    space = rng.integers(0, ns, n)
    blk = rng.integers(0, bps if blocks is None else min(blocks, bps), n)
    page = rng.integers(0, 512, n)
    # clustered pages so that faults of one va_block interact (prefetch, already-resident pages)
    page = np.where(rng.random(n) < 0.5, (page // 64) * 64 + rng.integers(0, 8, n), page)
    addr = (np.uint64(lay["va_base"]) + (blk.astype(np.uint64) << np.uint64(21)) + (page.astype(np.uint64) << np.uint64(12))).astype(np.uint64)
    outside = rng.random(n) < outside_frac
    addr[outside] += np.uint64(bps << 21)
    acc = rng.choice([TG.HW_VIRT_READ, TG.HW_VIRT_WRITE, TG.HW_VIRT_ATOMIC_STRONG, TG.HW_VIRT_ATOMIC_WEAK], n, p=[0.5, 0.3, 0.1, 0.1])
    inst = tbl["instance_ptr"][space].copy()
    stale = rng.random(n) < stale_frac
    inst[stale] ^= np.uint64(0x5000)
    pk = TG.encode_packets(inst, addr, acc, np.zeros(n, np.uint32), np.full(n, 0x20, np.uint32), np.arange(n, dtype=np.uint64) + 1)
    pk[:, 7] &= ~np.uint32(1 << 7)                     # REPLAYABLE = 0
    pk[:, 7] |= np.uint32(1 << 20)                     # CLIENT_TYPE = HUB
    pk[:, 6] = 65 + (np.arange(n) % 20)                # NV_PFAULT_MMU_ENG_ID_CE0..CE19 (hwref/blackwell/gb100/dev_fault.h)
    fatal = rng.random(n) < fatal_frac
    pk[fatal, 7] = (pk[fatal, 7] & ~np.uint32(0x1F)) | TG.HW_FAULT_TYPE_PDE_SIZE
    return np.ascontiguousarray(pk)


def run_checker(lib, prefix, pk, gp, tbl, layout, state, pf=(1, 51, 1), run_cap=1 << 18):
    """oracle_ / ref_service_non_replayable_faults on a copy of the state."""
    n = len(pk)
    entries = np.zeros(n, F.FAULT_ENTRY_DT)
    flags = np.zeros(n, "<u2")
    svc = np.zeros(n, SERVICE_GROUP_DT)
    runs = np.zeros(run_cap, COPY_RUN_DT)
    n_runs = ctypes.c_uint32(0)
    st = state.copy()
    ptrs = st.ptrs()
    fn = getattr(lib, prefix + "_service_non_replayable_faults")
    fn.restype = ctypes.c_uint32
    examined = fn(ctypes.c_uint32(n), P(pk), P(gp), P(tbl), ctypes.c_size_t(len(tbl)), P(layout), ctypes.byref(ptrs), ctypes.c_uint(pf[0]),
                  ctypes.c_uint(pf[1]), ctypes.c_uint(pf[2]), P(entries), P(flags), P(svc), P(runs), ctypes.c_uint32(run_cap),
                  ctypes.byref(n_runs))
    assert examined != 0xFFFFFFFF
    return dict(entries=entries, flags=flags, service=svc, runs=runs, n_runs=n_runs.value, state=st, examined=examined)


RECORD_FIELDS = ("slot", "region_first", "region_outer", "page_fault_count", "prefetch_count", "migrated", "first_touch",
                 "hint_residency", "run_count", "batch", "group", "new_residency")


def assert_equal(want, got, what=""):
    n = len(want["flags"])
    assert np.array_equal(want["flags"], got["flags"][:n]), what + " entry flags"
    for f in F.DECODE_FIELDS + ("flags", "num_instances", "access_type_mask"):
        assert np.array_equal(want["entries"][f], got["entries"][f][:n]), "%s entries.%s" % (what, f)
    want["state"].assert_equal(got["state"], what)
    a, b = want["service"], got["service"][:n]
    for f in RECORD_FIELDS:
        assert np.array_equal(a[f], b[f]), "%s record.%s" % (what, f)
    assert want["n_runs"] == got["n_runs"], "%s n_runs %d vs %d" % (what, want["n_runs"], got["n_runs"])
    has = np.nonzero(a["run_count"] > 0)[0]
    if len(has):
        within = np.concatenate([np.arange(c) for c in a["run_count"][has]])
        wo = np.repeat(a["run_offset"][has].astype(np.int64), a["run_count"][has]) + within
        go = np.repeat(b["run_offset"][has].astype(np.int64), b["run_count"][has]) + within
        assert np.array_equal(want["runs"][wo], got["runs"][go]), what + " copy runs"


def run_simt(sim, pk, gp, tbl, layout, state, pf=(1, 51, 1), run_cap=1 << 18, heavy_min=96):
    """The product's kernels (k_nonreplayable_prepare + the servicing pass) under the CPU SIMT emulator."""
    n = len(pk)
    B = int(gp["max_batch_size"][0])
    total = (n + B - 1) // B * B
    entries = np.zeros(n, F.FAULT_ENTRY_DT)
    flags = np.zeros(total, "<u2")
    svc = np.zeros(total, SERVICE_GROUP_DT)
    runs = np.zeros(run_cap, COPY_RUN_DT)
    counters = np.zeros(2, "<u4")
    st = state.copy()
    ptrs = st.ptrs()
    sim.simt_service_non_replayable_faults(P(pk), ctypes.c_uint32(n), P(gp), P(tbl), ctypes.c_uint32(len(tbl)), P(layout), ctypes.byref(ptrs),
                                           ctypes.c_uint32(pf[0]), ctypes.c_uint32(pf[1]), ctypes.c_uint32(pf[2]), P(entries), P(flags),
                                           P(svc), P(runs), ctypes.c_uint32(run_cap), P(counters), ctypes.c_uint32(heavy_min))
    return dict(entries=entries, flags=flags, service=svc, runs=runs, n_runs=int(counters[0]), state=st)
