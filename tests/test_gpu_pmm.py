"""GPU tests of section 1d: servicing under an HBM capacity (uvmtier_service_fault_batches_tiered_device).  The whole
loop -- chunks needed, victims in the reference's eviction order, eviction, servicing, list moves -- is replayed on the
CPU from the oracle's pieces and must leave the same state, flags, records and stamps; with real data behind the
masks every resident page still holds its bytes after being evicted and promoted again."""
import ctypes

import numpy as np
import pytest

import faultlib as F
import residentlib as R
import servicelib as S
import uvmtier
import uvmtier_testlib as T
from uvmtier import TIER_CXL, TIER_DRAM, TIER_HBM, tracegen as TG

pytestmark = pytest.mark.gpu
P = T.P
USED = 1 << 63


def _select_weighted(key, weight, target):
    """The smallest threshold T with sum(weight[key != 0 and key <= T]) >= target, and the indices it selects (every
    non-zero key when the weights do not add up): the selection rule of the tiered loop."""
    if target <= 0:
        return np.zeros(0, np.int64), False
    live = np.nonzero(key != 0)[0]
    order = live[np.argsort(key[live], kind="stable")]
    csum = np.cumsum(weight[order].astype(np.int64))
    if len(order) == 0 or csum[-1] < target:
        return order, True
    T = key[order[int(np.searchsorted(csum, target))]]
    return live[key[live] <= T], False


def cpu_tiered(orc, k1, gp, tbl, lay, layout, st0, capacity, chunk_batches, dram_cap_pages, demote=False):
    """The tiered loop from the oracle's pieces (oracle_pmm.c, oracle_resident.c, oracle_service.c): per round the
    va_blocks it faults on, the victims (ascending stamp until enough of them are va_blocks the round leaves alone),
    optionally the demotion of the coldest DRAM va_blocks, the evictions, the servicing pass, the list moves."""
    B = int(gp["max_batch_size"][0])
    nb = len(k1["batch"])
    n_slots = st0.n
    st = st0.copy()
    stamp = np.zeros(n_slots, "<u8")
    dram_stamp = np.zeros(n_slots, "<u8")
    clock = 0
    flags = np.zeros(nb * B, "<u2")
    batch = k1["batch"].copy()
    recs = np.zeros(nb * B, S.SERVICE_GROUP_DT)
    has = np.zeros(nb * B, bool)
    stats = dict(waves=0, dram=0, cxl=0, peak=0, demoted=0, over=0)
    rnd = 0

    def make_resident(rq, pm=None):
        nonlocal st
        res = R.run_oracle(orc, rq, pm, None, st, run_cap=1 << 20)
        st = res["state"]
        return res

    for b0 in range(0, nb, chunk_batches):
        n = min(chunk_batches, nb - b0)
        e0 = b0 * B
        rnd += 1
        touched = np.zeros(n_slots, bool)
        for b in range(b0, b0 + n):
            for g in range(int(batch["num_block_groups"][b])):
                grp = k1["groups"][b * B + g]
                if grp["flags"] & 1:
                    continue
                sp = int(tbl["va_space"][grp["ctx"]])
                blk = (int(grp["block_start"]) - lay["va_base"]) >> 21
                if 0 <= blk < lay["blocks_per_space"] and sp < lay["n_spaces"]:
                    touched[sp * lay["blocks_per_space"] + blk] = True
        gp_pages = T.bool_from_masks(st.gpu_resident).sum(axis=1)
        dp_pages = T.bool_from_masks(st.cpu_resident0).sum(axis=1)
        needed = int((touched & (stamp == 0)).sum())
        in_use = int((stamp != 0).sum())
        victims, short = _select_weighted(stamp, ((stamp != 0) & ~touched).astype(np.int64), in_use + needed - capacity)
        k = len(victims)
        if k:                                       # the same va_blocks pick_root_chunk_to_evict hands out, k times (pinned by libref_pmm)
            picked = np.zeros(k, "<u4")
            assert orc.oracle_pmm_pick(P(stamp), ctypes.c_uint32(n_slots), ctypes.c_uint32(k), P(picked)) == k
            assert set(picked.tolist()) == set(victims.tolist())
        arriving = int(gp_pages[victims].sum())
        dram_pages = int(dp_pages.sum())
        demoted = np.zeros(0, np.int64)
        if demote and k and dram_pages + arriving > dram_cap_pages:
            cand = (dp_pages != 0) & (stamp == 0) & ~touched
            demoted, _ = _select_weighted(np.where(cand, dram_stamp + 1, 0).astype("<u8"), dp_pages, dram_pages + arriving - dram_cap_pages)
        dest = R.TIER_DRAM if dram_pages - int(dp_pages[demoted].sum()) + arriving <= dram_cap_pages else R.TIER_CXL
        if len(demoted):
            rq = np.zeros(len(demoted), R.MAKE_RESIDENT_REQ_DT)
            rq["slot"] = demoted; rq["region_outer"] = 512; rq["dest_tier"] = R.TIER_CXL; rq["cause"] = R.CAUSE_API_MIGRATE
            rq["flags"] = R.MR_HAS_PAGE_MASK; rq["mask_row"] = np.arange(len(demoted))
            make_resident(rq, st.cpu_resident0[demoted].copy())
            dram_stamp[demoted] = 0
            stats["demoted"] += len(demoted)
        if k:
            rq = np.zeros(k, R.MAKE_RESIDENT_REQ_DT)
            rq["slot"] = victims; rq["region_outer"] = 512; rq["dest_tier"] = dest; rq["cause"] = R.CAUSE_EVICTION
            rq["flags"] = R.MR_EVICT_GPU_PAGES
            res = make_resident(rq)
            ptrs = st.ptrs()
            orc.oracle_pmm_note_make_resident(P(stamp), ctypes.c_uint32(n_slots), ctypes.c_uint64(clock), ctypes.c_uint32(k), P(rq),
                                              P(res["results"]), ctypes.byref(ptrs))
            on_dram = T.bool_from_masks(st.cpu_resident0).any(axis=1)
            moved = res["results"]["migrated"][:, 2] != 0
            if dest == R.TIER_DRAM:
                dram_stamp[victims[moved]] = rnd
            dram_stamp[victims[~on_dram[victims]]] = 0
            stats["waves"] += 1
            stats["dram" if dest == R.TIER_DRAM else "cxl"] += k
        after = in_use - k + needed + int(touched[victims].sum())
        stats["over"] += after > capacity
        stats["peak"] = max(stats["peak"], after)
        sub = dict(ordered_rec=k1["ordered_rec"][e0:e0 + n * B], groups=k1["groups"][e0:e0 + n * B], batch=batch[b0:b0 + n])
        want = S.run_service_oracle(orc, sub, gp, tbl, layout, st, run_cap=1 << 21)
        st = want["state"]
        flags[e0:e0 + n * B] = want["flags"]
        batch[b0:b0 + n] = want["batch"]
        pos = np.zeros(n * B, S.SERVICE_GROUP_DT)
        r = want["records"]
        pos[r["batch"].astype(np.int64) * B + r["group"].astype(np.int64)] = r
        orc.oracle_pmm_note_service(P(stamp), ctypes.c_uint32(n_slots), ctypes.c_uint64(clock), ctypes.c_uint32(n), ctypes.c_uint32(B),
                                    P(batch[b0:b0 + n].copy()), P(pos))
        clock += n * B
        r = r.copy()
        r["batch"] += b0
        recs[r["batch"].astype(np.int64) * B + r["group"].astype(np.int64)] = r
        has[r["batch"].astype(np.int64) * B + r["group"].astype(np.int64)] = True
    return dict(state=st, stamp=stamp, dram_stamp=dram_stamp, flags=flags, batch=batch, records=recs, has=has, stats=stats)


def preprocess(ctx, pk, gp, tbl):
    n = len(pk)
    B = int(gp["max_batch_size"][0])
    ring = np.zeros((n + 1, 8), np.uint32)
    ring[:n] = pk
    r = ctx.fetch_preprocess_fault_batches(ring, 0, n)
    k1 = F.product_to_checker_view(n, B, tbl, r["disp"], r["ordered"], r["groups"], r["batch"], r["entries"], ring[:n].copy())
    pad = len(r["batch"]) * B - n
    if pad:
        r["ordered"] = np.concatenate([r["ordered"], np.zeros(pad, r["ordered"].dtype)])
        r["groups"] = np.concatenate([r["groups"], np.zeros(pad, r["groups"].dtype)])
    k1["ordered_rec"] = r["ordered"]; k1["groups"] = r["groups"]
    return r, k1


@pytest.mark.parametrize("dist,capacity,chunk_batches,dram_cap_blocks,demote",
                         [("hotset", 40, 8, 30, False), ("uniform", 24, 3, 1000, False), ("stream", 3, 1, 0, False),
                          ("uniform", 500, 16, 10, False), ("hotset", 40, 8, 30, True), ("uniform", 64, 1, 60, True),
                          ("uniform", 24, 3, 100, True)])
def test_tiered_service_matches_the_cpu_loop(dist, capacity, chunk_batches, dram_cap_blocks, demote):
    orc = T.load_oracle()
    B = 64
    pk, tbl, lay = TG.fault_trace(12000, 31, dist, 2, total_va_bytes=128 << 21)
    gp = TG.gpu_params(B)
    n_slots = lay["blocks_per_space"] * 2
    st0 = S.random_state(n_slots, 11, populated=0.8)
    st0.gpu_resident[:] = 0
    st0.gpu_pte0[:] = 0; st0.gpu_pte1[:] = 0; st0.gpu_pte2[:] = 0
    layout = S.make_layout(lay)
    with uvmtier.Context(0) as ctx:
        ctx.fault_buffer_init(gp, tbl)
        r, k1 = preprocess(ctx, pk, gp, tbl)
        ds = ctx.block_state_alloc(layout)
        ds.upload(st0)
        pmm = ctx.pmm_create(layout, capacity)
        got = ctx.service_fault_batches_tiered(r, layout, ds, pmm, chunk_batches, dram_cap_blocks * 512, execute_copies=False,
                                               demote=demote)
        ctx.pmm_destroy(pmm)
        ds.free()
    want = cpu_tiered(orc, k1, gp, tbl, lay, layout, st0, capacity, chunk_batches, dram_cap_blocks * 512, demote)
    want["state"].assert_equal(got["state"], dist)
    assert np.array_equal(want["stamp"], got["stamps"])
    assert np.array_equal(want["dram_stamp"], got["dram_stamps"])
    assert np.array_equal(want["flags"], got["flags"])
    for f in ("num_duplicate_faults", "num_invalid_prefetch_faults"):
        assert np.array_equal(want["batch"][f], got["batch"][f])
    valid = want["has"]
    for f in S.RECORD_FIELDS:
        assert np.array_equal(want["records"][f][valid], got["service"][f][valid]), f
    s = got["stats"]
    assert (s.eviction_waves, s.evicted_to_dram, s.evicted_to_cxl, s.peak_chunks_in_use, s.blocks_demoted, s.overcommitted_passes) == \
        (want["stats"]["waves"], want["stats"]["dram"], want["stats"]["cxl"], want["stats"]["peak"], want["stats"]["demoted"],
         want["stats"]["over"])
    if demote:
        assert s.blocks_demoted > 0
    in_use = int((got["stamps"] != 0).sum())
    assert s.chunks_in_use == in_use and np.array_equal(T.bool_from_masks(got["state"].gpu_resident).any(axis=1), got["stamps"] != 0)
    if capacity < 100:
        # the capacity holds unless one chunk of batches alone touches more va_blocks than HBM has room for
        assert s.eviction_waves > 0 and in_use <= max(capacity, chunk_batches * B)
        if chunk_batches * B <= capacity:
            assert s.peak_chunks_in_use <= capacity
    else:
        assert s.eviction_waves == 0


def test_tiered_service_moves_the_right_bytes():
    """Real data behind the masks: 64 va_blocks, room for 20 on the GPU.  After the trace every page that was resident
    somewhere still holds its bytes -- in HBM, or in the tier it was evicted to -- and first-touched pages are zero."""
    n_spaces, blocks = 1, 64
    B = 64
    pk, tbl, lay = TG.fault_trace(6000, 9, "hotset", n_spaces, total_va_bytes=blocks << 21)
    gp = TG.gpu_params(B)
    n = lay["blocks_per_space"] * n_spaces
    st0 = S.random_state(n, 4, populated=0.7)
    st0.gpu_resident[:] = 0
    st0.gpu_pte0[:] = 0; st0.gpu_pte1[:] = 0; st0.gpu_pte2[:] = 0
    layout = S.make_layout(lay)
    with uvmtier.Context(0) as c:
        c.fault_buffer_init(gp, tbl)
        info = {t: c.tier_create(t, n << 21) for t in (TIER_HBM, TIER_DRAM, TIER_CXL)}
        views = {t: info[t].host_view().reshape(n * 512, 4096) for t in (TIER_DRAM, TIER_CXL)}
        truth = TG.rng_stream(123, n * 512 * 1024).view(np.uint8).reshape(n * 512, 4096)
        where = {TIER_DRAM: T.bool_from_masks(st0.cpu_resident0).ravel(), TIER_CXL: T.bool_from_masks(st0.cpu_resident1).ravel()}
        for t in (TIER_DRAM, TIER_CXL):
            views[t][:] = 0xEE
            views[t][where[t]] = truth[where[t]]
        uvmtier._check(uvmtier.lib().uvmtier_memset_device(c.handle, ctypes.c_void_p(info[TIER_HBM].device_ptr), 0xEE,
                                                           ctypes.c_size_t(n << 21)), "memset")
        resident0 = where[TIER_DRAM] | where[TIER_CXL]
        r, _ = preprocess(c, pk, gp, tbl)
        ds = c.block_state_alloc(layout)
        ds.upload(st0)
        pmm = c.pmm_create(layout, 20)
        got = c.service_fault_batches_tiered(r, layout, ds, pmm, 4, 15 * 512, execute_copies=True)
        s = got["stats"]
        assert s.eviction_waves > 3 and s.evicted_to_dram > 0 and s.evicted_to_cxl > 0 and s.pages_evicted > 0
        assert s.pages_promoted > s.pages_evicted
        state = got["state"]
        hbm = c.from_device(info[TIER_HBM].device_ptr, n << 21, np.uint8).reshape(n * 512, 4096)
        bufs = {TIER_HBM: hbm, TIER_DRAM: views[TIER_DRAM], TIER_CXL: views[TIER_CXL]}
        masks = {TIER_HBM: T.bool_from_masks(state.gpu_resident).ravel(), TIER_DRAM: T.bool_from_masks(state.cpu_resident0).ravel(),
                 TIER_CXL: T.bool_from_masks(state.cpu_resident1).ravel()}
        everywhere = masks[TIER_HBM] | masks[TIER_DRAM] | masks[TIER_CXL]
        assert not (resident0 & ~everywhere).any()
        for t, m in masks.items():
            keep = m & resident0
            assert np.array_equal(bufs[t][keep], truth[keep]), "bytes of pages resident on tier %d" % t
            assert not bufs[t][m & ~resident0].any(), "first-touched pages on tier %d are zero" % t
        assert int(masks[TIER_HBM].reshape(n, 512).any(axis=1).sum()) <= 20 + 4 * B
        c.pmm_destroy(pmm)
        ds.free()


@pytest.mark.parametrize("demote", [False, True])
def test_tiered_service_with_an_hbm_arena_smaller_than_the_va(demote):
    """uvmtier_pmm_enable_chunk_table: 96 va_blocks of VA over an HBM arena of 24 chunks.  va_blocks take a chunk when their
    first page arrives and give it back when they are evicted; after the trace every page that was resident somewhere
    still holds its bytes -- in its va_block's chunk, or in the tier it was evicted (or demoted) to -- first-touched pages
    are zero, no two va_blocks share a chunk, and masks, stamps and records equal the CPU loop's (the table changes where
    bytes live, not what the loop decides)."""
    orc = T.load_oracle()
    blocks, cap, B, chunk_batches, dram_cap = 96, 24, 8, 3, 30 * 512
    pk, tbl, lay = TG.fault_trace(5000, 17, "uniform", 1, total_va_bytes=blocks << 21)
    gp = TG.gpu_params(B)
    n = lay["blocks_per_space"]
    st0 = S.random_state(n, 6, populated=0.7)
    st0.gpu_resident[:] = 0
    st0.gpu_pte0[:] = 0; st0.gpu_pte1[:] = 0; st0.gpu_pte2[:] = 0
    layout = S.make_layout(lay)
    with uvmtier.Context(0) as c:
        c.fault_buffer_init(gp, tbl)
        info = {TIER_HBM: c.tier_create(TIER_HBM, cap << 21), TIER_DRAM: c.tier_create(TIER_DRAM, n << 21),
                TIER_CXL: c.tier_create(TIER_CXL, n << 21)}
        views = {t: info[t].host_view().reshape(n * 512, 4096) for t in (TIER_DRAM, TIER_CXL)}
        truth = TG.rng_stream(321, n * 512 * 1024).view(np.uint8).reshape(n * 512, 4096)
        where = {TIER_DRAM: T.bool_from_masks(st0.cpu_resident0).ravel(), TIER_CXL: T.bool_from_masks(st0.cpu_resident1).ravel()}
        for t in (TIER_DRAM, TIER_CXL):
            views[t][:] = 0xEE
            views[t][where[t]] = truth[where[t]]
        uvmtier._check(uvmtier.lib().uvmtier_memset_device(c.handle, ctypes.c_void_p(info[TIER_HBM].device_ptr), 0xEE,
                                                           ctypes.c_size_t(cap << 21)), "memset")
        resident0 = where[TIER_DRAM] | where[TIER_CXL]
        r, k1 = preprocess(c, pk, gp, tbl)
        ds = c.block_state_alloc(layout)
        ds.upload(st0)
        pmm = c.pmm_create(layout, cap)
        c.pmm_enable_chunk_table(pmm)
        # a round may not fault on more va_blocks than there are chunks
        st_code = uvmtier.lib().uvmtier_service_fault_batches_tiered_device(
            c.handle, ctypes.byref(pmm), ctypes.c_uint32(1), uvmtier.P(1), uvmtier.P(1), uvmtier.P(1), uvmtier.P(layout), ctypes.byref(ds.ptrs),
            uvmtier.P(1), uvmtier.P(1), ctypes.c_uint32(4), ctypes.c_uint64(0), ctypes.c_int(1), None)
        assert st_code == 0x1F
        got = c.service_fault_batches_tiered(r, layout, ds, pmm, chunk_batches, dram_cap, execute_copies=True, demote=demote)
        s = got["stats"]
        assert s.eviction_waves > 10 and s.peak_chunks_in_use <= cap and s.overcommitted_passes == 0
        assert (s.blocks_demoted > 0) == demote
        state, table = got["state"], got["chunk_of_slot"]
        holders = np.nonzero(got["stamps"] != 0)[0]
        assert (table[holders] < cap).all() and len(set(table[holders].tolist())) == len(holders)
        assert (np.delete(table, holders) == 0xFFFFFFFF).all()
        hbm_raw = c.from_device(info[TIER_HBM].device_ptr, cap << 21, np.uint8).reshape(cap, 512, 4096)
        hbm = np.zeros((n, 512, 4096), np.uint8)
        hbm[holders] = hbm_raw[table[holders]]
        bufs = {TIER_HBM: hbm.reshape(n * 512, 4096), TIER_DRAM: views[TIER_DRAM], TIER_CXL: views[TIER_CXL]}
        masks = {TIER_HBM: T.bool_from_masks(state.gpu_resident).ravel(), TIER_DRAM: T.bool_from_masks(state.cpu_resident0).ravel(),
                 TIER_CXL: T.bool_from_masks(state.cpu_resident1).ravel()}
        assert not (resident0 & ~(masks[TIER_HBM] | masks[TIER_DRAM] | masks[TIER_CXL])).any()
        for t, m in masks.items():
            keep = m & resident0
            assert np.array_equal(bufs[t][keep], truth[keep]), "bytes of pages resident on tier %d" % t
            assert not bufs[t][m & ~resident0].any(), "first-touched pages on tier %d are zero" % t
        if demote:
            assert int(masks[TIER_DRAM].sum()) <= dram_cap + 512      # the budget holds up to what one va_block adds
        c.pmm_destroy(pmm)
        ds.free()
    want = cpu_tiered(orc, k1, gp, tbl, lay, layout, st0, cap, chunk_batches, dram_cap, demote)
    want["state"].assert_equal(got["state"], "chunk table")
    assert np.array_equal(want["stamp"], got["stamps"]) and np.array_equal(want["flags"], got["flags"])
