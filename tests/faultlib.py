"""Helpers for the fault-preprocessing tests: struct layouts and runners for the CPU checkers."""
import ctypes

import numpy as np

from uvmtier_testlib import P
from uvmtier import tracegen as TG

FE_FATAL, FE_REPLAYABLE, FE_VIRTUAL, FE_PROTECTED, FE_FILTERED = 0x01, 0x02, 0x04, 0x08, 0x10
NV_WARN_MORE_PROCESSING_REQUIRED = 0x10005

FAULT_ENTRY_DT = np.dtype([
    ("fault_address", "<u8"), ("timestamp", "<u8"), ("instance_ptr_address", "<u8"),
    ("client_id", "<u2"), ("mmu_engine_id", "<u2"), ("utlb_id", "<u2"), ("num_instances", "<u2"),
    ("instance_ptr_aperture", "u1"), ("fault_type", "u1"), ("fault_access_type", "u1"), ("client_type", "u1"),
    ("mmu_engine_type", "u1"), ("gpc_id", "u1"), ("ve_id", "u1"), ("flags", "u1"),
    ("access_type_mask", "u1"), ("pad", "u1", 7)])
assert FAULT_ENTRY_DT.itemsize == 48
BATCH_DT = np.dtype([("num_cached_faults", "<u4"), ("num_coalesced_faults", "<u4"), ("is_single_instance_ptr", "<u4"),
                     ("max_utlb_id", "<u4"), ("status", "<u4"), ("num_duplicate_faults", "<u4"),
                     ("num_invalid_prefetch_faults", "<u4"), ("num_block_groups", "<u4")])
assert BATCH_DT.itemsize == 32

DECODE_FIELDS = ("fault_address", "timestamp", "instance_ptr_address", "client_id", "mmu_engine_id", "utlb_id",
                 "instance_ptr_aperture", "fault_type", "fault_access_type", "client_type", "mmu_engine_type",
                 "gpc_id", "ve_id")


def n_batches(n, B):
    return (n + B - 1) // B


def run_parse(lib, prefix, packets, gp):
    n = len(packets)
    out = np.zeros(n, FAULT_ENTRY_DT)
    getattr(lib, prefix + "_faults_parse")(P(packets), ctypes.c_uint32(n), P(gp), P(out))
    return out


def run_fetch_preprocess(lib, prefix, packets, gp, table):
    """CPU checker run. Returns dict(entries, ordered, va_space, gpu, batch, ring_after)."""
    n = len(packets)
    B = int(gp["max_batch_size"][0])
    ring = packets.copy()
    entries = np.zeros(n, FAULT_ENTRY_DT)
    ordered = np.full(n, 0xFFFFFFFF, "<u4")
    vs = np.full(n, 0xFFFFFFFF, "<u4")
    gpu = np.full(n, 0xFFFFFFFF, "<u4")
    batch = np.zeros(n_batches(n, B) + 1, BATCH_DT)
    fn = getattr(lib, prefix + "_faults_fetch_preprocess")
    fn.restype = ctypes.c_int
    # utlbs[].num_pending_faults per batch (optional output of both checkers)
    U = utlb_count(gp)
    utlb = np.zeros((len(batch), U), "<u2")
    setter = getattr(lib, prefix + "_faults_set_utlb_output", None)
    if setter is not None:
        setter(P(utlb))
    nb = fn(P(ring), ctypes.c_uint32(n), P(gp), P(table), ctypes.c_size_t(len(table)), P(entries), P(ordered), P(vs),
            P(gpu), P(batch))
    if setter is not None:
        setter(None)
    return dict(entries=entries, ordered=ordered, va_space=vs, gpu=gpu, batch=batch[:nb], ring_after=ring,
                utlb_pending=utlb[:nb] if setter is not None else None)


def utlb_count(gp):
    return int(gp["gpc_count"][0]) * int(gp["utlb_per_gpc_count"][0])


def assert_fetch_equal(a, b, what=""):
    """Compares two fetch+preprocess results on everything the reference defines."""
    assert len(a["batch"]) == len(b["batch"]), what
    for f in ("num_cached_faults", "num_coalesced_faults", "is_single_instance_ptr", "max_utlb_id", "status"):
        assert np.array_equal(a["batch"][f], b["batch"][f]), "%s batch.%s" % (what, f)
    ea, eb = a["entries"], b["entries"]
    for f in DECODE_FIELDS + ("flags",):
        assert np.array_equal(ea[f], eb[f]), "%s entries.%s" % (what, f)
    live = (ea["flags"] & FE_FILTERED) == 0      # num_instances / mask are only defined on representatives
    assert np.array_equal(ea["num_instances"][live], eb["num_instances"][live]), what + " num_instances"
    assert np.array_equal(ea["access_type_mask"][live], eb["access_type_mask"][live]), what + " access_type_mask"
    # ordered view: only for batches that translated
    B = None
    base = 0
    for k, bt in enumerate(a["batch"]):
        nco = int(bt["num_coalesced_faults"])
        if bt["status"] == 0:
            for key in ("ordered", "va_space", "gpu"):
                assert np.array_equal(a[key][base:base + nco], b[key][base:base + nco]), "%s %s batch %d" % (what, key, k)
        base += int(bt["num_cached_faults"])
    assert np.array_equal(a["ring_after"], b["ring_after"]), what + " ring valid bits"
    if a.get("utlb_pending") is not None and b.get("utlb_pending") is not None:
        assert np.array_equal(a["utlb_pending"], b["utlb_pending"]), what + " utlbs[].num_pending_faults"


def directed_quirk_trace():
    """READ@uTLB1, WRITE@uTLB1, READ@uTLB2 on one page: the WRITE takes over as
    representative with num_instances 2, and the now-filtered first entry absorbs the third."""
    tbl = TG.instance_ptr_table(1)
    addr = np.full(3, TG.VA_BASE + 0x5000, np.uint64)
    pk = TG.encode_packets(tbl["instance_ptr"][[0, 0, 0]], addr, [TG.HW_VIRT_READ, TG.HW_VIRT_WRITE, TG.HW_VIRT_READ],
                           [0, 0, 0], [0x00, 0x00, 0x01], [10, 11, 12])
    return pk, tbl


# ----------------------------------------------------------------------------- product-side layouts
ORDERED_DT = np.dtype([("fault_address", "<u8"), ("fault_index", "<u2"), ("ctx", "<u2"), ("num_instances", "<u2"),
                       ("access_type_mask", "u1"), ("fault_access_type", "u1")])
assert ORDERED_DT.itemsize == 16
GROUP_DT = np.dtype([("block_start", "<u8"), ("ctx", "<u2"), ("first_fault_index", "<u2"), ("block_faults", "<u2"),
                     ("flags", "<u2")])
assert GROUP_DT.itemsize == 16
DISP_FILTERED, DISP_FATAL, DISP_VALID = 0x10000, 0x20000, 0x40000


def product_to_checker_view(n, B, table, disp, ordered, groups, batch, entries, ring_after):
    """Re-expresses the product's outputs in the dict layout run_fetch_preprocess() returns."""
    nb = n_batches(n, B)
    o = np.full(n, 0xFFFFFFFF, "<u4"); vs = o.copy(); gpu = o.copy()
    for b in range(nb):
        base = b * B
        if batch["status"][b] != 0:
            continue
        m = int(batch["num_coalesced_faults"][b])
        rec = ordered[base:base + m]
        o[base:base + m] = rec["fault_index"]
        vs[base:base + m] = table["va_space"][rec["ctx"]]
        gpu[base:base + m] = table["gpu_id"][rec["ctx"]]
    return dict(entries=entries, ordered=o, va_space=vs, gpu=gpu, batch=batch[:nb], ring_after=ring_after, disp=disp,
                ordered_rec=ordered, groups=groups)


def run_fetch_preprocess_simt(sim, packets, gp, table, clear_valid=1, get=0, force_cta_kernel=0):
    n = len(packets)
    B = int(gp["max_batch_size"][0])
    ring = packets.copy()
    disp = np.zeros(n, "<u4")
    ordered = np.full(n, 0xFF, "u1").repeat(16).view(ORDERED_DT)[:n].copy()
    groups = np.full(n * 16, 0xFF, "u1").view(GROUP_DT).copy()
    batch = np.zeros(n_batches(n, B), BATCH_DT)
    entries = np.zeros(n, FAULT_ENTRY_DT)
    sim.simt_fault_fetch_preprocess.restype = ctypes.c_int
    # ring of n+1 entries so that put = n is a legal index (the ring never wraps here)
    ring1 = np.zeros((n + 1, 8), np.uint32); ring1[:n] = ring
    utlb = np.zeros((len(batch), utlb_count(gp)), "<u2")
    r = sim.simt_fault_fetch_preprocess(P(ring1), ctypes.c_uint32(n + 1), ctypes.c_uint32(get), ctypes.c_uint32(n), P(gp),
                                        P(table), ctypes.c_uint32(len(table)), ctypes.c_int(clear_valid), P(disp),
                                        P(ordered), P(groups), P(batch), P(entries), P(utlb), ctypes.c_int(force_cta_kernel))
    assert r == len(batch), r
    res = product_to_checker_view(n, B, table, disp, ordered, groups, batch, entries, ring1[:n].copy())
    res["utlb_pending"] = utlb
    return res


def check_groups(res, table, B):
    """Structural check of the va_block runs against the sorted view (reference rule: same va_space, gpu and
    2 MB block; fatal entries stand alone)."""
    for b, bt in enumerate(res["batch"]):
        if bt["status"] != 0:
            assert bt["num_block_groups"] == 0
            continue
        base = b * B
        m = int(bt["num_coalesced_faults"])
        rec = res["ordered_rec"][base:base + m]
        ng = int(bt["num_block_groups"])
        g = res["groups"][base:base + ng]
        assert int(g["block_faults"].sum()) == m
        pos = 0
        for k in range(ng):
            assert g["first_fault_index"][k] == pos
            run = rec[pos:pos + int(g["block_faults"][k])]
            assert len(run) >= 1
            assert (run["fault_address"] >> np.uint64(21) == g["block_start"][k] >> np.uint64(21)).all()
            vs = table["va_space"][run["ctx"]]; gp = table["gpu_id"][run["ctx"]]
            assert (vs == vs[0]).all() and (gp == gp[0]).all()
            fatal = (run["fault_access_type"] & 0x80) != 0
            assert not fatal.any() or len(run) == 1
            assert g["flags"][k] == (1 if fatal.any() else 0)
            if k + 1 < ng and not fatal.any():
                nxt = rec[pos + len(run)]
                same = (nxt["fault_address"] >> np.uint64(21) == g["block_start"][k] >> np.uint64(21)) and \
                    table["va_space"][nxt["ctx"]] == vs[0] and table["gpu_id"][nxt["ctx"]] == gp[0]
                assert (not same) or (nxt["fault_access_type"] & 0x80), "run split although the next entry belongs to it"
            pos += len(run)
