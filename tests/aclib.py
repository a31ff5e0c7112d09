"""Helpers for the access-counter tests: NVC365 packet encoder, dtypes, runners (oracle / SIMT emulator)."""
import ctypes

import numpy as np

from uvmtier_testlib import P

from uvmtier.blockstate import COPY_RUN_DT, SERVICE_GROUP_DT
from uvmtier.layouts import BATCH_DT, GROUP_DT

AC_ENTRY_DT = np.dtype([("address", "<u8"), ("instance_ptr", "<u8"), ("instance_aperture", "<u4"), ("mmu_engine_id", "<u4"),
                        ("ve_id", "<u4"), ("counter_value", "<u4"), ("sub_granularity", "<u4"), ("bank", "<u4"), ("tag", "<u4"),
                        ("valid", "<u4")])
ORDERED_NOTIF_DT = np.dtype([("address", "<u8"), ("index", "<u2"), ("ctx", "<u2"), ("sub_granularity", "<u4")])
assert (AC_ENTRY_DT.itemsize, ORDERED_NOTIF_DT.itemsize) == (48, 16)


def encode_notifications(inst_ptr, address, sub_granularity, ve_id=None, counter=None, tag=None, bank=None, valid=None, va_bits=57):
    """NVC365_NOTIFY_BUF_ENTRY encoder (UVM/clc365.h:33-80): TYPE = GPU, ADDR_TYPE = GVA, instance block in video memory."""
    n = len(address)
    w = np.zeros((n, 8), np.uint32)
    ip = np.asarray(inst_ptr, np.uint64) >> np.uint64(12)
    bank = np.zeros(n, np.uint32) if bank is None else np.asarray(bank, np.uint32)
    w[:, 0] = 1 | (bank << 2) | ((ip & np.uint64(0xFFFFF)) << np.uint64(12)).astype(np.uint32)
    w[:, 1] = (ip >> np.uint64(20)).astype(np.uint32)
    a = np.asarray(address, np.uint64) & np.uint64((1 << va_bits) - 1)
    w[:, 2] = (a & np.uint64(0xFFFFFFFF)).astype(np.uint32)
    w[:, 3] = (a >> np.uint64(32)).astype(np.uint32)
    w[:, 4] = np.asarray(sub_granularity, np.uint32)
    w[:, 5] = (np.arange(n) if tag is None else np.asarray(tag)).astype(np.uint32) & 0xFFFFF
    w[:, 6] = (np.full(n, 256) if counter is None else np.asarray(counter)).astype(np.uint32) & 0xFFFF
    ve = np.zeros(n, np.uint32) if ve_id is None else np.asarray(ve_id, np.uint32)
    v = np.ones(n, np.uint32) if valid is None else np.asarray(valid, np.uint32)
    w[:, 7] = ((384 + ve) << 20) | (v << 31)
    return w


def _outputs(n, B):
    nb = (n + B - 1) // B
    return (np.zeros(nb * B, ORDERED_NOTIF_DT), np.zeros(nb * B, GROUP_DT), np.zeros(nb, BATCH_DT), np.zeros(nb * B, AC_ENTRY_DT))


def run_preprocess(lib, fn_name, ring, get, put, B, tbl, va_bits=57, clear_valid=True):
    """oracle_access_counters_fetch_preprocess / simt_access_counters_fetch_preprocess on a copy of the ring."""
    ring = ring.copy()
    n = put - get if put >= get else len(ring) - get + put
    ordered, groups, batch, entries = _outputs(n, B)
    getattr(lib, fn_name)(P(ring), ctypes.c_uint32(len(ring)), ctypes.c_uint32(get), ctypes.c_uint32(put), ctypes.c_uint32(B),
                          ctypes.c_uint32(va_bits), P(tbl), ctypes.c_uint32(len(tbl)), ctypes.c_int(1 if clear_valid else 0),
                          P(ordered), P(groups), P(batch), P(entries))
    return dict(ordered=ordered, groups=groups, batch=batch, entries=entries, ring=ring, n=n)


def assert_preprocess_equal(a, b, B):
    assert np.array_equal(a["batch"], b["batch"]) and np.array_equal(a["ring"], b["ring"])
    assert np.array_equal(a["entries"][:a["n"]], b["entries"][:b["n"]])
    for bi in range(len(a["batch"])):                      # slots past the counts are unspecified
        m, g = int(a["batch"]["num_cached_faults"][bi]), int(a["batch"]["num_block_groups"][bi])
        assert np.array_equal(a["ordered"][bi * B:bi * B + m], b["ordered"][bi * B:bi * B + m]), bi
        assert np.array_equal(a["groups"][bi * B:bi * B + g], b["groups"][bi * B:bi * B + g]), bi


def run_service_oracle(orc, k, B, tbl, layout, state, pf=(1, 51, 1), granularity_2m=True, cpu_chunk_2m=True, run_cap=1 << 20):
    nb = len(k["batch"])
    svc, runs = np.zeros(nb * B, SERVICE_GROUP_DT), np.zeros(run_cap, COPY_RUN_DT)
    st = state.copy()
    ptrs = st.ptrs()
    n_runs = ctypes.c_uint32(0)
    fn = orc.oracle_access_counters_service
    fn.restype = ctypes.c_uint32
    nsvc = fn(ctypes.c_uint32(nb), ctypes.c_uint32(B), P(k["ordered"]), P(k["groups"]), P(k["batch"]), P(tbl), P(layout),
              ctypes.byref(ptrs), ctypes.c_uint(pf[0]), ctypes.c_uint(pf[1]), ctypes.c_uint(pf[2]), ctypes.c_int(granularity_2m),
              ctypes.c_int(cpu_chunk_2m), P(svc), P(runs), ctypes.c_uint32(run_cap), ctypes.byref(n_runs))
    return dict(records=svc[:nsvc], runs=runs, n_runs=n_runs.value, state=st, flags=np.zeros(nb * B, "<u2"), batch=k["batch"].copy())


def run_service_ref(ref, k, B, tbl, layout, state, pf=(1, 51, 1), granularity_2m=True, cpu_chunk_2m=True, run_cap=1 << 20):
    """oracle/_ref/libref_access_counters.so: the reference's own expand_notification_block and service_va_block_locked
    (verbatim); same contract as run_service_oracle."""
    nb = len(k["batch"])
    svc, runs = np.zeros(nb * B, SERVICE_GROUP_DT), np.zeros(run_cap, COPY_RUN_DT)
    st = state.copy()
    ptrs = st.ptrs()
    n_runs = ctypes.c_uint32(0)
    fn = ref.ref_access_counters_service
    fn.restype = ctypes.c_uint32
    nsvc = fn(ctypes.c_uint32(nb), ctypes.c_uint32(B), P(k["ordered"]), P(k["groups"]), P(k["batch"]), P(tbl), P(layout),
              ctypes.byref(ptrs), ctypes.c_uint(pf[0]), ctypes.c_uint(pf[1]), ctypes.c_uint(pf[2]), ctypes.c_int(granularity_2m),
              ctypes.c_int(cpu_chunk_2m), P(svc), P(runs), ctypes.c_uint32(run_cap), ctypes.byref(n_runs))
    return dict(records=svc[:nsvc], runs=runs, n_runs=n_runs.value, state=st)


def run_service_simt(sim, k, B, tbl, layout, state, pf=(1, 51, 1), granularity_2m=True, cpu_chunk_2m=True, run_cap=1 << 20,
                     heavy_min=96):
    nb = len(k["batch"])
    svc, runs, counters = np.zeros(nb * B, SERVICE_GROUP_DT), np.zeros(run_cap, COPY_RUN_DT), np.zeros(2, "<u4")
    st = state.copy()
    ptrs = st.ptrs()
    batch = k["batch"].copy()
    sim.simt_access_counters_service(ctypes.c_uint32(nb), ctypes.c_uint32(B), P(k["ordered"]), P(k["groups"]), P(batch), P(tbl),
                                     P(layout), ctypes.byref(ptrs), ctypes.c_uint32(pf[0]), ctypes.c_uint32(pf[1]),
                                     ctypes.c_uint32(pf[2]), ctypes.c_int(granularity_2m), ctypes.c_int(cpu_chunk_2m), P(svc), P(runs),
                                     ctypes.c_uint32(run_cap), P(counters), ctypes.c_uint32(heavy_min))
    return dict(service=svc, runs=runs, n_runs=int(counters[0]), state=st, flags=np.zeros(nb * B, "<u2"), batch=batch)
