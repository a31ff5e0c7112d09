"""CPU tests of uvm_va_block_make_resident on the state table (section 1c of include/uvmtier.h): the product's
kernel under the SIMT emulator vs the CPU restatement, directed cases for the reference's rules (source order,
first touch, eviction marks, unmapping, page_mask / region clipping), and a promotion / eviction round trip."""
import numpy as np
import pytest

import residentlib as R
import servicelib as S
import uvmtier_testlib as T


def _state(n_slots, seed, rdup=False):
    st = S.random_state(n_slots, seed)
    rng = np.random.default_rng(seed + 100)
    # some pages carry an eviction mark: always pages that live on the CPU now
    cpu = T.bool_from_masks(st.cpu_resident0) | T.bool_from_masks(st.cpu_resident1)
    st.gpu_evicted = T.masks_from_bool(cpu & (rng.random(cpu.shape) < 0.3))
    return R.with_read_duplication(st, seed) if rdup else st


@pytest.mark.parametrize("seed,rdup", [(1, False), (2, False), (3, False), (4, True), (5, True), (6, True)])
def test_simt_make_resident_matches_oracle(seed, rdup):
    n_slots = 96
    st0 = _state(n_slots, seed, rdup)
    rq, pm, pf = R.random_requests(n_slots, seed, bad=3, rdup_frac=0.4 if rdup else 0.0, evict_cause_frac=0.3)
    want = R.run_oracle(T.load_oracle(), rq, pm, pf, st0)
    got = R.run_simt(T.load_simt(), rq, pm, pf, st0)
    R.assert_equal(want, got, "seed %d" % seed)
    R.check_invariants(st0, got)
    assert (want["results"]["slot"][-3:] == 0xFFFFFFFF).all()          # the three malformed requests are rejected
    assert int(want["results"]["migrated"].sum()) > 1000 and int(want["results"]["first_touch"].sum()) > 100


def _one(st, **kw):
    rq = np.zeros(1, R.MAKE_RESIDENT_REQ_DT)
    for k, v in kw.items():
        rq[k] = v
    return rq


def _bits(mask_row):
    return set(np.nonzero(T.bool_from_masks(mask_row.reshape(1, 8))[0])[0].tolist())


def test_directed_rules():
    orc = T.load_oracle()
    st = R.HostBlockState(2)
    setb = lambda m, s, pages: [m.__setitem__((s, p // 64), m[s, p // 64] | np.uint64(1 << (p % 64))) for p in pages]
    setb(st.cpu_resident0, 0, [1, 2, 3]); setb(st.cpu_resident1, 0, [4, 5, 9]); setb(st.gpu_resident, 0, [6, 7, 8])
    for m in (st.gpu_pte0, st.gpu_pte1):
        setb(m, 0, [6, 7])
    # (a) to the GPU, region [0, 12): node 0 then node 1 as sources, pages 0, 10, 11 are first touch
    r = R.run_oracle(orc, _one(st, slot=0, region_first=0, region_outer=12, dest_tier=R.TIER_HBM, cause=R.CAUSE_API_MIGRATE), None, None, st)
    res = r["results"][0]
    assert list(res["migrated"]) == [3, 3, 0] and res["first_touch"] == 3 and res["unmapped"] == 0
    runs = r["runs"][:res["run_count"]]
    assert [(int(x["first_page"]), int(x["bits"]) & 1023, (int(x["bits"]) >> 10) & 7, int(x["bits"]) >> 14) for x in runs] == \
        [(1, 3, 0, 0), (4, 2, 1, 0), (9, 1, 1, 0), (0, 1, 7, 0), (10, 2, 7, 0)]
    assert _bits(r["state"].gpu_resident[0]) == set(range(12)) and not r["state"].cpu_resident0[0].any()
    # (b) eviction to the CXL node: only what the GPU holds moves, mappings go, the pages are marked evicted
    r = R.run_oracle(orc, _one(st, slot=0, region_first=0, region_outer=512, dest_tier=R.TIER_CXL, cause=R.CAUSE_EVICTION,
                               flags=R.MR_EVICT_GPU_PAGES), None, None, st)
    res = r["results"][0]
    assert list(res["migrated"]) == [0, 0, 3] and res["first_touch"] == 0 and res["unmapped"] == 2
    assert _bits(r["state"].cpu_resident1[0]) == {4, 5, 9, 6, 7, 8} and _bits(r["state"].cpu_resident0[0]) == {1, 2, 3}
    assert _bits(r["state"].gpu_evicted[0]) == {6, 7, 8} and not r["state"].gpu_pte0[0].any()
    x = r["runs"][0]
    assert (int(x["first_page"]), int(x["bits"]) & 1023, (int(x["bits"]) >> 10) & 7, int(x["bits"]) >> 14) == (6, 3, 2, 2)
    # (c) DRAM -> CXL demotion with a page mask: CPU to CPU copy, the other node loses every page of the mask
    pm = T.masks_from_bool(np.isin(np.arange(512), [2, 3, 4, 7, 300]).reshape(1, 512))
    r = R.run_oracle(orc, _one(st, slot=0, region_first=0, region_outer=256, dest_tier=R.TIER_CXL, cause=R.CAUSE_API_MIGRATE,
                               flags=R.MR_HAS_PAGE_MASK), pm, None, st)
    res = r["results"][0]
    assert list(res["migrated"]) == [2, 0, 1] and res["first_touch"] == 0          # 2, 3 from DRAM, 7 from the GPU; 4 is there
    assert _bits(r["state"].cpu_resident0[0]) == {1} and _bits(r["state"].gpu_resident[0]) == {6, 8}
    assert not r["state"].gpu_evicted[0].any()                                     # not an eviction
    # (d) coming back to the GPU clears the eviction mark, and a prefetch mask splits the runs by cause
    st2 = st.copy()
    setb(st2.gpu_evicted, 0, [1, 2, 3, 4])
    pf = T.masks_from_bool(np.isin(np.arange(512), [2]).reshape(1, 512))
    r = R.run_oracle(orc, _one(st2, slot=0, region_first=1, region_outer=4, dest_tier=R.TIER_HBM, cause=R.CAUSE_REPLAYABLE_FAULT,
                               flags=R.MR_HAS_PREFETCH_MASK), None, pf, st2)
    assert _bits(r["state"].gpu_evicted[0]) == {4}
    runs = r["runs"][:r["results"]["run_count"][0]]
    assert [(int(x["first_page"]), int(x["bits"]) & 1023, (int(x["bits"]) >> 13) & 1) for x in runs] == [(1, 1, 0), (2, 1, 1), (3, 1, 0)]


def test_promote_evict_round_trip_restores_residency_on_one_tier():
    """Everything to the GPU, then everything evicted to DRAM: every page that was resident anywhere (or was touched)
    ends on node 0, the GPU is empty, and the eviction marks are exactly the pages that left the GPU."""
    n_slots = 64
    st0 = _state(n_slots, 7)
    sim = T.load_simt()
    up = np.zeros(n_slots, R.MAKE_RESIDENT_REQ_DT)
    up["slot"] = np.arange(n_slots); up["region_outer"] = 512; up["dest_tier"] = R.TIER_HBM; up["cause"] = R.CAUSE_API_MIGRATE
    a = R.run_simt(sim, up, None, None, st0)
    assert (T.bool_from_masks(a["state"].gpu_resident)).all() and not a["state"].gpu_evicted.any()
    down = up.copy()
    down["dest_tier"] = R.TIER_DRAM; down["cause"] = R.CAUSE_EVICTION; down["flags"] = R.MR_EVICT_GPU_PAGES
    b = R.run_simt(sim, down, None, None, a["state"])
    assert T.bool_from_masks(b["state"].cpu_resident0).all() and not b["state"].gpu_resident.any()
    assert T.bool_from_masks(b["state"].gpu_evicted).all() and not b["state"].gpu_pte0.any()
    assert int(b["results"]["migrated"][:, 2].sum()) == n_slots * 512 and b["n_runs"] == n_slots


# ------------------------------------------------------------------------------------------------
# The CPU restatement against the reference's own uvm_va_block_make_resident (compiled verbatim into
# oracle/_ref/libref_resident.so).
@pytest.mark.parametrize("seed,rdup", [(1, False), (2, False), (3, False), (4, False), (5, False),
                                       (11, True), (12, True), (13, True), (14, True), (15, True)])
def test_oracle_matches_the_reference_make_resident(seed, rdup):
    """rdup: the table carries read_duplicated pages and 40 % of the requests are
    uvm_va_block_make_resident_read_duplicate calls (the others break read duplication where they land)."""
    ref = T.load_ref("resident")
    if ref is None:
        pytest.skip("oracle/_ref/libref_resident.so not built (needs /root/reference)")
    n_slots = 128
    st0 = _state(n_slots, seed, rdup)
    rq, pm, pf = R.random_requests(n_slots, seed, frac=0.9, bad=3, rdup_frac=0.4 if rdup else 0.0, evict_cause_frac=0.3)
    want = R.run_ref(ref, rq, pm, pf, st0)
    got = R.run_oracle(T.load_oracle(), rq, pm, pf, st0)
    R.assert_equal(want, got, "reference vs restatement, seed %d" % seed)
    res = want["results"]
    assert int(res["migrated"][:, 0].sum()) > 500 and int(res["migrated"][:, 1].sum()) > 200 and int(res["migrated"][:, 2].sum()) > 300
    assert int(res["first_touch"].sum()) > 100 and int(res["unmapped"].sum()) > 50


@pytest.mark.parametrize("seed,gpu_id,peers", [(21, 1, ((3, True),)), (22, 4, ((2, True), (6, True))), (23, 2, ((1, False), (5, True))),
                                               (24, 3, ((1, True), (2, False), (7, False), (9, True))), (25, 9, ((4, False),))])
def test_oracle_matches_the_reference_with_other_gpus(seed, gpu_id, peers):
    """M2 with several GPUs in the table: sources closest first (the other GPUs in ascending id before the CPU for a GPU
    destination, uvm_va_space.c:1879-1924), pages held only by a GPU without direct access staged through CPU pages
    (pages_staged uvm_va_block.c:3140-3170 -- block_populate_pages is compiled verbatim too -- and the two hops
    :4754-4829), CPU destinations fed by every GPU, every other holder losing the pages afterwards."""
    ref = T.load_ref("resident")
    if ref is None:
        pytest.skip("oracle/_ref/libref_resident.so not built (needs /root/reference)")
    n_slots = 160
    st0 = R.with_peers(_state(n_slots, seed, False), seed, peers)
    rq, pm, pf = R.random_requests(n_slots, seed, frac=0.9, bad=3, evict_cause_frac=0.3)
    want = R.run_ref(ref, rq, pm, pf, st0, gpu_id=gpu_id)
    got = R.run_oracle(T.load_oracle(), rq, pm, pf, st0, gpu_id=gpu_id)
    R.assert_equal(want, got, "reference vs restatement with peers, seed %d" % seed)
    res = want["results"]
    assert int(res["migrated_peer"].sum()) > 300
    staged = any(not d for _, d in peers)
    src = (want["runs"]["bits"][:want["n_runs"]] >> 10) & 7
    dst = want["runs"]["bits"][:want["n_runs"]] >> 14
    # a staged page makes two runs: peer -> DRAM, then DRAM -> HBM
    assert (int(((src >= 3) & (src < 7) & (dst == 1)).sum()) > 20) or not staged


@pytest.mark.parametrize("seed,gpu_id,peers", [(31, 1, ((3, True),)), (32, 4, ((2, True), (6, True))), (33, 2, ((1, False), (5, True))),
                                               (34, 3, ((1, True), (2, False), (7, False), (9, True))), (35, 9, ((4, False),)),
                                               (36, 5, ((1, False), (2, False), (3, False), (4, False)))])
def test_simt_make_resident_with_other_gpus_matches_oracle(seed, gpu_id, peers):
    """The product kernel (under the emulator) on tables that name other GPUs: same masks, counts and copy runs as the
    restatement, which the test above holds against the reference's code."""
    n_slots = 128
    st0 = R.with_peers(_state(n_slots, seed, False), seed, peers)
    rq, pm, pf = R.random_requests(n_slots, seed, frac=0.9, bad=3, evict_cause_frac=0.3)
    want = R.run_oracle(T.load_oracle(), rq, pm, pf, st0, gpu_id=gpu_id)
    got = R.run_simt(T.load_simt(), rq, pm, pf, st0, gpu_id=gpu_id)
    R.assert_equal(want, got, "emulator vs restatement with peers, seed %d" % seed)
    assert int(want["results"]["migrated_peer"].sum()) > 300
    # read duplication is not defined for such tables: the request is rejected, nothing changes
    st1 = st0.copy().enable_read_duplication()
    rq1 = _one(st1, slot=0, region_first=0, region_outer=512, dest_tier=R.TIER_HBM, cause=R.CAUSE_API_MIGRATE, flags=R.MR_READ_DUPLICATE)
    for run in (R.run_oracle(T.load_oracle(), rq1, None, None, st1, gpu_id=gpu_id), R.run_simt(T.load_simt(), rq1, None, None, st1, gpu_id=gpu_id)):
        assert run["results"]["slot"][0] == 0xFFFFFFFF
        run["state"].assert_equal(st1, "rejected request")


def test_reference_cpu_page_copy_loop_is_byte_exact():
    """libref_copy.so = block_copy_pages (uvm_va_block.c:3946-3991) compiled verbatim with kmap -> identity: copies the
    pages of a region of every va_block, one memcpy(PAGE_SIZE) per page, and marks them dirty at the destination."""
    import ctypes
    rc = T.load_ref("copy")
    if rc is None:
        pytest.skip("reference libraries not built (no /root/reference)")
    rc.ref_copy_pages_cpu.restype = ctypes.c_uint64
    rc.ref_copy_pages_cpu.argtypes = [ctypes.c_void_p, ctypes.c_void_p] + [ctypes.c_uint32] * 4
    blocks = 5
    src = np.random.default_rng(3).integers(0, 256, blocks << 21, dtype=np.uint8)
    dst = np.full(blocks << 21, 0xEE, np.uint8)
    marked = rc.ref_copy_pages_cpu(dst.ctypes.data, src.ctypes.data, 1, 4, 7, 300)
    assert marked == 3 * (300 - 7)
    want = np.full(blocks << 21, 0xEE, np.uint8)
    for b in range(1, 4):
        lo, hi = (b << 21) + 7 * 4096, (b << 21) + 300 * 4096
        want[lo:hi] = src[lo:hi]
    assert np.array_equal(dst, want)
