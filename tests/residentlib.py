"""Helpers for the make_resident tests: request builders, runners (oracle / SIMT emulator), comparison."""
import ctypes

import numpy as np

from uvmtier_testlib import P, bool_from_masks, masks_from_bool

from uvmtier.blockstate import (CAUSE_API_MIGRATE, CAUSE_EVICTION, CAUSE_REPLAYABLE_FAULT, COPY_RUN_DT,  # noqa: F401
                                MAKE_RESIDENT_REQ_DT, MAKE_RESIDENT_RESULT_DT, MR_EVICT_GPU_PAGES, MR_HAS_PAGE_MASK,
                                MR_HAS_PREFETCH_MASK, MR_READ_DUPLICATE, TIER_CXL, TIER_DRAM, TIER_HBM, HostBlockState)

RESULT_FIELDS = ("slot", "migrated", "migrated_peer", "first_touch", "run_count", "unmapped", "pages_migrated",
                 "pages_changed_residency")


def with_peers(st, seed, peers=((3, True), (5, False)), share=0.25):
    """Adds other GPUs to a random state: `peers` = (gpu id, direct copy access) in ascending id; a share of the pages
    that live on the CPU or nowhere move to a peer's memory (every page still lives in one place)."""
    rng = np.random.default_rng(seed + 900)
    free = ~bool_from_masks(st.gpu_resident)
    if st.read_duplicated is not None:
        free &= ~bool_from_masks(st.read_duplicated)
    pick = free & (rng.random(free.shape) < share)
    owner = rng.integers(0, len(peers), free.shape)
    for k, (gid, direct) in enumerate(peers):
        mine = pick & (owner == k)
        st.cpu_resident0 &= ~masks_from_bool(mine)
        st.cpu_resident1 &= ~masks_from_bool(mine)
        st.add_peer(gid, direct, masks_from_bool(mine))
    return st


def with_read_duplication(st, seed, dup=0.15, quirk=0.01):
    """Turns read duplication on for a random state: a share of the CPU-resident pages also have a GPU copy (the GPU
    maps such pages read-only at most), and a few CPU-only pages carry the mark a NUMA node switch leaves
    (uvm_va_block.c:5288-5310).  Every read_duplicated page has a CPU copy."""
    rng = np.random.default_rng(seed + 500)
    st.enable_read_duplication()
    cpu = bool_from_masks(st.cpu_resident0) | bool_from_masks(st.cpu_resident1)
    both = cpu & (rng.random(cpu.shape) < dup)
    st.gpu_resident |= masks_from_bool(both)
    st.gpu_pte1 &= ~masks_from_bool(both)
    st.gpu_pte2 &= ~masks_from_bool(both)
    st.gpu_evicted &= ~masks_from_bool(both)
    st.read_duplicated = masks_from_bool(both | (cpu & (rng.random(cpu.shape) < quirk)))
    return st


def random_requests(n_slots, seed, frac=0.6, evict_frac=0.25, bad=0, rdup_frac=0.0, evict_cause_frac=0.0):
    """At most one request per slot (a va_block is locked during make_resident), random destination, region,
    optional page / prefetch masks; a share of them are evictions of whatever the GPU holds.  rdup_frac of the others
    are read-duplicating copies; evict_cause_frac of the moves to the CPU carry cause EVICTION without the
    "GPU pages only" flag (the eviction marks then cover everything that migrated, :5023-5037)."""
    rng = np.random.default_rng(seed)
    slots = np.nonzero(rng.random(n_slots) < frac)[0]
    rng.shuffle(slots)
    n = len(slots)
    rq = np.zeros(n + bad, MAKE_RESIDENT_REQ_DT)
    rq["slot"][:n] = slots
    a = rng.integers(0, 512, n)
    b = rng.integers(0, 513, n)
    whole = rng.random(n) < 0.3
    rq["region_first"][:n] = np.where(whole, 0, np.minimum(a, b))
    rq["region_outer"][:n] = np.where(whole, 512, np.maximum(a, b))
    evict = rng.random(n) < evict_frac
    rq["dest_tier"][:n] = np.where(evict, rng.choice([TIER_DRAM, TIER_CXL], n), rng.choice([TIER_HBM, TIER_DRAM, TIER_CXL], n))
    rq["cause"][:n] = np.where(evict, CAUSE_EVICTION, rng.choice([CAUSE_API_MIGRATE, CAUSE_REPLAYABLE_FAULT], n))
    has_pm = ~evict & (rng.random(n) < 0.5)
    has_pf = ~evict & (rq["cause"][:n] == CAUSE_REPLAYABLE_FAULT) & (rng.random(n) < 0.6)
    rdup = ~evict & (rng.random(n) < rdup_frac)
    plain_evict = ~evict & ~rdup & ~has_pf & (rq["dest_tier"][:n] != TIER_HBM) & (rng.random(n) < evict_cause_frac)
    rq["cause"][:n] = np.where(plain_evict, CAUSE_EVICTION, rq["cause"][:n])
    rq["flags"][:n] = np.where(evict, MR_EVICT_GPU_PAGES,
                               has_pm * MR_HAS_PAGE_MASK + has_pf * MR_HAS_PREFETCH_MASK + rdup * MR_READ_DUPLICATE)
    rq["mask_row"][:n] = np.arange(n)
    page_masks = masks_from_bool(rng.random((n, 512)) < rng.random((n, 1)))
    prefetch_masks = masks_from_bool(rng.random((n, 512)) < 0.5)
    if bad:                                           # rejected requests: slot out of range, inverted region, bad tier
        rq["slot"][n:] = [n_slots + 3, 0, 0][:bad]
        rq["region_first"][n:] = [0, 9, 0][:bad]
        rq["region_outer"][n:] = [4, 3, 4][:bad]
        rq["dest_tier"][n:] = [0, 0, 5][:bad]
    return rq, page_masks, prefetch_masks


def run_oracle(orc, rq, page_masks, prefetch_masks, state, run_cap=1 << 18, gpu_id=1):
    st = state.copy()
    ptrs = st.ptrs()
    orc.oracle_make_resident_set_gpu_id(ctypes.c_uint32(gpu_id))
    res = np.zeros(len(rq), MAKE_RESIDENT_RESULT_DT)
    runs = np.zeros(run_cap, COPY_RUN_DT)
    fn = orc.oracle_make_resident
    fn.restype = ctypes.c_uint32
    n_runs = fn(ctypes.c_uint32(len(rq)), P(rq), P(page_masks), P(prefetch_masks), ctypes.c_uint32(st.n), ctypes.byref(ptrs),
                P(res), P(runs), ctypes.c_uint32(run_cap))
    return dict(results=res, runs=runs, n_runs=int(n_runs), state=st)


def run_ref(ref, rq, page_masks, prefetch_masks, state, run_cap=1 << 18, gpu_id=1):
    """oracle/_ref/libref_resident.so: the reference's own uvm_va_block_make_resident (verbatim) on the flat table."""
    st = state.copy()
    ptrs = st.ptrs()
    res = np.zeros(len(rq), MAKE_RESIDENT_RESULT_DT)
    runs = np.zeros(run_cap, COPY_RUN_DT)
    fn = ref.ref_make_resident
    fn.restype = ctypes.c_uint32
    n_runs = fn(ctypes.c_uint32(len(rq)), P(rq), P(page_masks), P(prefetch_masks), ctypes.c_uint32(st.n), ctypes.c_uint32(gpu_id),
                ctypes.byref(ptrs), P(res), P(runs), ctypes.c_uint32(run_cap))
    return dict(results=res, runs=runs, n_runs=int(n_runs), state=st)


def run_simt(sim, rq, page_masks, prefetch_masks, state, run_cap=1 << 18, gpu_id=1):
    st = state.copy()
    ptrs = st.ptrs()
    res = np.zeros(len(rq), MAKE_RESIDENT_RESULT_DT)
    runs = np.zeros(run_cap, COPY_RUN_DT)
    counters = np.zeros(2, "<u4")
    sim.simt_make_resident(ctypes.c_uint32(len(rq)), P(rq), P(page_masks), P(prefetch_masks), ctypes.c_uint32(st.n),
                           ctypes.byref(ptrs), P(res), P(runs), ctypes.c_uint32(run_cap), P(counters), ctypes.c_uint32(gpu_id))
    return dict(results=res, runs=runs, n_runs=int(counters[0]), state=st)


def assert_equal(want, got, what=""):
    want["state"].assert_equal(got["state"], what)
    a, b = want["results"], got["results"]
    for f in RESULT_FIELDS:
        assert np.array_equal(a[f], b[f]), "%s result.%s" % (what, f)
    assert want["n_runs"] == got["n_runs"], "%s n_runs %d vs %d" % (what, want["n_runs"], got["n_runs"])
    has = np.nonzero(a["run_count"] > 0)[0]
    if len(has):                                      # runs per request (their place in the array is free)
        within = np.concatenate([np.arange(c) for c in a["run_count"][has]])
        ia = np.repeat(a["run_offset"][has].astype(np.int64), a["run_count"][has]) + within
        ib = np.repeat(b["run_offset"][has].astype(np.int64), a["run_count"][has]) + within
        assert np.array_equal(want["runs"][ia], got["runs"][ib]), what + " copy runs"


def check_invariants(before, res):
    """Size-independent properties of make_resident."""
    st = res["state"]
    g, c0, c1 = (bool_from_masks(m) for m in (st.gpu_resident, st.cpu_resident0, st.cpu_resident1))
    rd = bool_from_masks(st.read_duplicated) if st.read_duplicated is not None else np.zeros_like(g)
    assert not (g & (c0 | c1) & ~rd).any(), "a page is resident in two places without being read-duplicated"
    assert not (c0 & c1).any(), "a page is resident on two CPU nodes"
    assert not (rd & ~(c0 | c1)).any(), "a read-duplicated page has no CPU copy"
    p1 = bool_from_masks(st.gpu_pte1)
    assert not (p1 & g & (c0 | c1)).any(), "the GPU can write a page that has a second copy"
    gb, c0b, c1b = (bool_from_masks(m) for m in (before.gpu_resident, before.cpu_resident0, before.cpu_resident1))
    # nothing is lost: a page resident somewhere before is resident somewhere after
    assert not ((gb | c0b | c1b) & ~(g | c0 | c1)).any(), "a resident page vanished"
    p0 = bool_from_masks(st.gpu_pte0)
    assert not (p0 & gb & ~g).any(), "the GPU still maps a page that left it"
    ev = bool_from_masks(st.gpu_evicted)
    assert not (ev & g).any(), "an evicted page is GPU resident"
