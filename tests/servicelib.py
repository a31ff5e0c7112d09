"""Helpers for the servicing-pass tests (stage C/D/E): layouts, block-state tables, runners, comparison."""
import ctypes

import numpy as np

from uvmtier_testlib import P, bool_from_masks, masks_from_bool

from uvmtier.blockstate import (BlockStatePtrs, COPY_RUN_DT, HostBlockState, PREFETCH_INFO_DT, SERVICE_GROUP_DT,  # noqa: F401
                                STATE_FIELDS, VA_LAYOUT_DT, make_layout)


def random_state(n_slots, seed, gpu_id=1, populated=0.7, ro_frac=0.1, none_frac=0.02, pref_frac=0.3):
    """Random but consistent residency: every page lives in at most one place; GPU PTE bits are nested and only
    on GPU-resident pages; some blocks are read-only / have no permissions; some prefer the GPU."""
    rng = np.random.default_rng(seed)
    st = HostBlockState(n_slots)
    where = rng.choice(4, (n_slots, 512), p=[1 - populated, populated * 0.5, populated * 0.2, populated * 0.3])
    untouched = rng.random(n_slots) < 0.25                  # never-populated blocks: first touch
    where[untouched] = 0
    runs = rng.random(n_slots) < 0.3                        # run-structured GPU residency
    for i in np.nonzero(runs & ~untouched)[0]:
        a = rng.integers(0, 512)
        b = rng.integers(a, 512) + 1
        where[i, a:b] = 3
    st.cpu_resident0 = masks_from_bool(where == 1)
    st.cpu_resident1 = masks_from_bool(where == 2)
    gres = where == 3
    st.gpu_resident = masks_from_bool(gres)
    r = rng.random((n_slots, 512))
    rd = gres & (r < 0.8)
    wr = rd & (r < 0.5)
    at = wr & (r < 0.3)
    # a few remote mappings: the GPU maps pages that live in CPU memory (make_resident unmaps them when they move)
    remote = ((where == 1) | (where == 2)) & (rng.random((n_slots, 512)) < 0.02)
    rd |= remote
    wr |= remote & (r < 0.5)
    st.gpu_pte0, st.gpu_pte1, st.gpu_pte2 = masks_from_bool(rd), masks_from_bool(wr), masks_from_bool(at)
    u = rng.random(n_slots)
    st.logical_prot = np.where(u < none_frac, 0, np.where(u < none_frac + ro_frac, 1, 3)).astype("u1")
    ro = st.logical_prot != 3                                # no write/atomic mappings under a read-only VMA
    st.gpu_pte1[ro] = 0
    st.gpu_pte2[ro] = 0
    st.gpu_pte0[st.logical_prot == 0] = 0
    st.preferred_location = np.where(rng.random(n_slots) < pref_frac, gpu_id, 257).astype("<u2")
    st.prefetch_info["last_migration_proc_id"] = np.where(rng.random(n_slots) < 0.5, gpu_id, 257)
    st.prefetch_info["fault_migrations_to_last_proc"] = rng.integers(0, 3, n_slots)
    return st


def _alloc_outputs(n_batches, B, run_cap):
    total = n_batches * B
    return (np.zeros(total, "<u2"), np.zeros(total, SERVICE_GROUP_DT), np.zeros(max(run_cap, 1), COPY_RUN_DT), np.zeros(2, "<u4"))


def run_service_oracle(orc, k1, gp, table, layout, state, pf=(1, 51, 1), run_cap=1 << 20):
    """oracle_service_fault_batches on the product-layout outputs of the preprocess stage (dict k1)."""
    B = int(gp["max_batch_size"][0])
    nb = len(k1["batch"])
    flags, _, runs, _ = _alloc_outputs(nb, B, run_cap)
    svc = np.zeros(nb * B, SERVICE_GROUP_DT)
    batch = k1["batch"].copy()
    st = state.copy()
    ptrs = st.ptrs()
    n_runs = ctypes.c_uint32(0)
    fn = orc.oracle_service_fault_batches
    fn.restype = ctypes.c_uint32
    nsvc = fn(ctypes.c_uint32(nb), ctypes.c_uint32(B), P(k1["ordered_rec"]), P(k1["groups"]), P(batch), P(table), P(layout),
              ctypes.byref(ptrs), ctypes.c_uint(pf[0]), ctypes.c_uint(pf[1]), ctypes.c_uint(pf[2]), P(flags), P(svc), P(runs),
              ctypes.c_uint32(run_cap), ctypes.byref(n_runs))
    return dict(flags=flags, records=svc[:nsvc], runs=runs, n_runs=n_runs.value, batch=batch, state=st)


def run_service_ref(ref, k1, gp, table, layout, state, pf=(1, 51, 1), run_cap=1 << 20):
    """oracle/_ref/libref_service.so: the reference's own service_fault_batch_block_locked (verbatim) drives the
    entry loop; the back half is the CPU restatement.  Same outputs as run_service_oracle, except that entry
    flags carry only what the reference keeps per entry (fatal + reason + cancel mode, invalid prefetch)."""
    B = int(gp["max_batch_size"][0])
    nb = len(k1["batch"])
    flags, _, runs, _ = _alloc_outputs(nb, B, run_cap)
    svc = np.zeros(nb * B, SERVICE_GROUP_DT)
    batch = k1["batch"].copy()
    st = state.copy()
    ptrs = st.ptrs()
    n_runs = ctypes.c_uint32(0)
    fn = ref.ref_service_fault_batches
    fn.restype = ctypes.c_uint32
    nsvc = fn(ctypes.c_uint32(nb), ctypes.c_uint32(B), P(k1["ordered_rec"]), P(k1["groups"]), P(batch), P(table), P(layout),
              ctypes.byref(ptrs), ctypes.c_uint(pf[0]), ctypes.c_uint(pf[1]), ctypes.c_uint(pf[2]), P(flags), P(svc), P(runs),
              ctypes.c_uint32(run_cap), ctypes.byref(n_runs))
    return dict(flags=flags, records=svc[:nsvc], runs=runs, n_runs=n_runs.value, batch=batch, state=st)


def use_reference_back_half(svc_lib, ref_prefetch, ref_resident, pf=(1, 51, 1)):
    """Routes the back half of the servicing checker `svc_lib` (liboracle.so or libref_service.so) to the reference's
    own code: uvm_perf_prefetch.c compiled unmodified (libref_prefetch.so) and uvm_va_block_make_resident compiled
    verbatim (libref_resident.so).  Pass None, None to go back to the restatements.  Returns True when hooked."""
    fn = svc_lib.oracle_service_set_back_half
    fn.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
    fn.restype = None
    if ref_prefetch is None or ref_resident is None:
        fn(None, None)
        return False
    ref_prefetch.ref_prefetch_init(ctypes.c_uint(pf[0]), ctypes.c_uint(pf[1]), ctypes.c_uint(pf[2]))
    fn(ctypes.cast(ref_prefetch.ref_prefetch_get_hint_va_block, ctypes.c_void_p), ctypes.cast(ref_resident.ref_make_resident, ctypes.c_void_p))
    return True


REF_FLAG_MASK = 0x0004 | 0x0008 | 0x0700 | 0x1000     # FATAL, INVALID_PREFETCH, fatal reason, cancel mode


def assert_ref_equals_oracle(ref_res, orc_res, what=""):
    """Reference loop vs restated loop: everything except the flags the reference does not keep per entry and
    page_fault_count (a local variable of the reference's function)."""
    assert np.array_equal(ref_res["flags"] & REF_FLAG_MASK, orc_res["flags"] & REF_FLAG_MASK), what + " entry flags"
    for f in ("num_duplicate_faults", "num_invalid_prefetch_faults"):
        assert np.array_equal(ref_res["batch"][f], orc_res["batch"][f]), "%s batch.%s" % (what, f)
    ref_res["state"].assert_equal(orc_res["state"], what)
    a, b = ref_res["records"], orc_res["records"]
    assert len(a) == len(b), "%s number of records %d vs %d" % (what, len(a), len(b))
    for f in RECORD_FIELDS:
        if f != "page_fault_count":
            assert np.array_equal(a[f], b[f]), "%s record.%s" % (what, f)
    assert ref_res["n_runs"] == orc_res["n_runs"]
    assert np.array_equal(ref_res["runs"][:ref_res["n_runs"]], orc_res["runs"][:orc_res["n_runs"]]), what + " copy runs"


def run_service_simt(sim, k1, gp, table, layout, state, pf=(1, 51, 1), run_cap=1 << 20, lanes=1, heavy_min=96):
    """The product's servicing kernels under the CPU SIMT emulator.  Slots holding heavy_min entries or more go to the
    warp-per-block kernel (0 = all of them, 0xffffffff = none), the others to the lane-group kernel (`lanes` is unused)."""
    B = int(gp["max_batch_size"][0])
    nb = len(k1["batch"])
    flags, svc, runs, counters = _alloc_outputs(nb, B, run_cap)
    batch = k1["batch"].copy()
    st = state.copy()
    ptrs = st.ptrs()
    sim.simt_service_fault_batches(ctypes.c_uint32(nb), ctypes.c_uint32(B), P(k1["ordered_rec"]), P(k1["groups"]), P(batch),
                                   P(table), P(layout), ctypes.byref(ptrs), ctypes.c_uint32(pf[0]), ctypes.c_uint32(pf[1]),
                                   ctypes.c_uint32(pf[2]), P(flags), P(svc), P(runs), ctypes.c_uint32(run_cap), P(counters),
                                   ctypes.c_int(lanes), ctypes.c_uint32(heavy_min))
    return dict(flags=flags, service=svc, runs=runs, n_runs=int(counters[0]), batch=batch, state=st)


RECORD_FIELDS = ("slot", "region_first", "region_outer", "page_fault_count", "prefetch_count", "migrated", "first_touch",
                 "hint_residency", "run_count", "batch", "group", "new_residency")


def assert_service_equal(want, got, B, what=""):
    """want: oracle result (compact records in (batch, group) order); got: product result (record of run g of
    batch b at b*B + g).  Copy runs are compared per record (their placement in the run array is free)."""
    assert np.array_equal(want["flags"], got["flags"]), what + " entry flags"
    for f in ("num_duplicate_faults", "num_invalid_prefetch_faults"):
        assert np.array_equal(want["batch"][f], got["batch"][f]), "%s batch.%s" % (what, f)
    want["state"].assert_equal(got["state"], what)
    rec = want["records"]
    pos = rec["batch"].astype(np.int64) * B + rec["group"].astype(np.int64)
    grec = got["service"][pos]
    for f in RECORD_FIELDS:
        assert np.array_equal(rec[f], grec[f]), "%s record.%s" % (what, f)
    assert want["n_runs"] == got["n_runs"], "%s n_runs %d vs %d" % (what, want["n_runs"], got["n_runs"])
    has = np.nonzero(rec["run_count"] > 0)[0]
    if len(has):
        within = np.concatenate([np.arange(c) for c in rec["run_count"][has]])
        wo = np.repeat(rec["run_offset"][has].astype(np.int64), rec["run_count"][has]) + within
        go = np.repeat(grec["run_offset"][has].astype(np.int64), rec["run_count"][has]) + within
        assert np.array_equal(want["runs"][wo], got["runs"][go]), what + " copy runs"


def check_service_invariants(res, state_before):
    """Size-independent properties of a serviced trace."""
    st = res["state"]
    g, c0, c1 = (bool_from_masks(m) for m in (st.gpu_resident, st.cpu_resident0, st.cpu_resident1))
    assert not (g & c0).any() and not (g & c1).any() and not (c0 & c1).any(), "a page is resident in two places"
    p0, p1, p2 = (bool_from_masks(m) for m in (st.gpu_pte0, st.gpu_pte1, st.gpu_pte2))
    assert not (p2 & ~p1).any() and not (p1 & ~p0).any(), "GPU PTE bits are not nested"
    gb = bool_from_masks(state_before.gpu_resident)
    assert not (gb & ~g).any(), "servicing never takes pages away from the faulting GPU"
