"""GPU tests of uvmtier_make_resident_device (section 1c): the masks, counts and copy runs are bit-exact against the
CPU restatement, and executing the runs moves the right bytes between HBM, host DRAM and the CXL node in every
direction (promotion, eviction, DRAM <-> CXL demotion / promotion, first-touch zero fill)."""
import numpy as np
import pytest

import residentlib as R
import servicelib as S
import uvmtier
import uvmtier_testlib as T
from uvmtier import TIER_CXL, TIER_DRAM, TIER_HBM, tracegen as TG

pytestmark = pytest.mark.gpu


def _state(n_slots, seed, rdup=False):
    st = S.random_state(n_slots, seed)
    rng = np.random.default_rng(seed + 100)
    cpu = T.bool_from_masks(st.cpu_resident0) | T.bool_from_masks(st.cpu_resident1)
    st.gpu_evicted = T.masks_from_bool(cpu & (rng.random(cpu.shape) < 0.3))
    return R.with_read_duplication(st, seed) if rdup else st


@pytest.mark.parametrize("seed,n_slots,rdup", [(1, 96, False), (5, 4096, False), (2, 96, True), (6, 4096, True)])
def test_make_resident_matches_oracle(seed, n_slots, rdup):
    """rdup: the table has read-duplicated pages and 40 % of the requests are read-duplicating copies
    (uvm_va_block_make_resident_read_duplicate); some moves to the CPU carry cause EVICTION without the flag."""
    st0 = _state(n_slots, seed, rdup)
    rq, pm, pf = R.random_requests(n_slots, seed, bad=3, rdup_frac=0.4 if rdup else 0.0, evict_cause_frac=0.3)
    want = R.run_oracle(T.load_oracle(), rq, pm, pf, st0, run_cap=1 << 20)
    lay = dict(va_base=TG.VA_BASE, blocks_per_space=n_slots, n_spaces=1)
    layout = S.make_layout(lay)
    with uvmtier.Context(0) as c:
        ds = c.block_state_alloc(layout)
        ds.upload(st0)
        got = c.make_resident(rq, pm, pf, layout, ds, run_cap=1 << 20)
        ds.free()
    R.assert_equal(want, got, "seed %d" % seed)
    R.check_invariants(st0, got)


def test_moves_between_all_tiers_are_byte_exact():
    """Three waves on 48 va_blocks with real data behind the masks: everything the requests name goes to the GPU,
    then half of the blocks are evicted to the CXL node and the other half to DRAM, then the CXL half is promoted to
    DRAM (CPU to CPU).  After each wave every resident page holds the bytes it started with (or zeros if it was
    first touched), wherever it lives now."""
    n = 48
    rng = np.random.default_rng(3)
    st0 = S.random_state(n, 9, populated=0.8)
    with uvmtier.Context(0) as c:
        info = {t: c.tier_create(t, n << 21) for t in (TIER_HBM, TIER_DRAM, TIER_CXL)}
        views = {t: info[t].host_view() for t in (TIER_DRAM, TIER_CXL)}
        # the truth: one 4 KB pattern per page, placed in the arena of the tier that holds the page
        truth = TG.rng_stream(77, n * 512 * 1024).view(np.uint8).reshape(n * 512, 4096)
        where = {TIER_HBM: T.bool_from_masks(st0.gpu_resident).ravel(), TIER_DRAM: T.bool_from_masks(st0.cpu_resident0).ravel(),
                 TIER_CXL: T.bool_from_masks(st0.cpu_resident1).ravel()}
        hbm0 = np.full((n * 512, 4096), 0xEE, np.uint8)
        hbm0[where[TIER_HBM]] = truth[where[TIER_HBM]]
        uvmtier._check(uvmtier.lib().uvmtier_memcpy_h2d(c.handle, uvmtier.P(info[TIER_HBM].device_ptr), uvmtier.P(hbm0),
                                                        __import__("ctypes").c_size_t(hbm0.nbytes)), "h2d")
        for t in (TIER_DRAM, TIER_CXL):
            v = views[t].reshape(n * 512, 4096)
            v[:] = 0xEE
            v[where[t]] = truth[where[t]]
        resident0 = where[TIER_HBM] | where[TIER_DRAM] | where[TIER_CXL]

        lay = dict(va_base=TG.VA_BASE, blocks_per_space=n, n_spaces=1)
        layout = S.make_layout(lay)
        ds = c.block_state_alloc(layout)
        ds.upload(st0)

        def check(state, touched, what):
            hbm = c.from_device(info[TIER_HBM].device_ptr, n << 21, np.uint8).reshape(n * 512, 4096)
            bufs = {TIER_HBM: hbm, TIER_DRAM: views[TIER_DRAM].reshape(n * 512, 4096), TIER_CXL: views[TIER_CXL].reshape(n * 512, 4096)}
            masks = {TIER_HBM: T.bool_from_masks(state.gpu_resident).ravel(), TIER_DRAM: T.bool_from_masks(state.cpu_resident0).ravel(),
                     TIER_CXL: T.bool_from_masks(state.cpu_resident1).ravel()}
            for t, m in masks.items():
                keep = m & resident0
                assert np.array_equal(bufs[t][keep], truth[keep]), "%s: bytes of pages resident on tier %d" % (what, t)
                assert not bufs[t][m & touched].any(), "%s: first-touched pages on tier %d are zero" % (what, t)
            return masks

        def all_blocks(dest, cause, flags, slots):
            rq = np.zeros(len(slots), R.MAKE_RESIDENT_REQ_DT)
            rq["slot"] = slots; rq["region_outer"] = 512; rq["dest_tier"] = dest; rq["cause"] = cause; rq["flags"] = flags
            return rq

        # wave 1: promotion of a random half of every block's pages to the GPU (page masks)
        pm = T.masks_from_bool(rng.random((n, 512)) < 0.5)
        rq = all_blocks(TIER_HBM, R.CAUSE_API_MIGRATE, R.MR_HAS_PAGE_MASK, np.arange(n))
        rq["mask_row"] = np.arange(n)
        a = c.make_resident(rq, pm, None, layout, ds, execute=True)
        touched = T.bool_from_masks(pm).ravel() & ~resident0
        m1 = check(a["state"], touched, "promotion")
        assert m1[TIER_HBM][T.bool_from_masks(pm).ravel()].all() and int(a["results"]["first_touch"].sum()) == int(touched.sum())

        # wave 2: eviction of whatever the GPU holds: even blocks to the CXL node, odd blocks to DRAM
        rq = np.concatenate([all_blocks(TIER_CXL, R.CAUSE_EVICTION, R.MR_EVICT_GPU_PAGES, np.arange(0, n, 2)),
                             all_blocks(TIER_DRAM, R.CAUSE_EVICTION, R.MR_EVICT_GPU_PAGES, np.arange(1, n, 2))])
        b = c.make_resident(rq, None, None, layout, ds, execute=True)
        m2 = check(b["state"], touched, "eviction")
        assert not m2[TIER_HBM].any() and np.array_equal(T.bool_from_masks(b["state"].gpu_evicted).ravel(), m1[TIER_HBM])
        assert int(b["results"]["migrated"][:, 2].sum()) == int(m1[TIER_HBM].sum())

        # wave 3: the even blocks' CXL pages move to DRAM (CPU to CPU through the GPU's copy kernel)
        rq = all_blocks(TIER_DRAM, R.CAUSE_API_MIGRATE, 0, np.arange(0, n, 2))
        d = c.make_resident(rq, None, None, layout, ds, execute=True)
        m3 = check(d["state"], touched | ~(resident0 | touched), "CXL -> DRAM")
        even = np.repeat(np.arange(n) % 2 == 0, 512)
        assert not m3[TIER_CXL][even].any() and m3[TIER_DRAM][even].all()         # whole-block region: the rest is first touch
        assert int(d["results"]["migrated"][:, 1].sum()) == int(m2[TIER_CXL][even].sum())

        # wave 4: read-duplicate the first half of every block onto the GPU: both copies hold the same bytes, the CPU
        # keeps its pages, and the GPU's copy is marked read_duplicated
        ds.enable_read_duplication()
        rq = all_blocks(TIER_HBM, R.CAUSE_API_MIGRATE, R.MR_READ_DUPLICATE, np.arange(n))
        rq["region_outer"] = 256
        e = c.make_resident(rq, None, None, layout, ds, execute=True)
        m4 = check(e["state"], touched | ~(resident0 | touched), "read duplication")
        first_half = np.tile(np.arange(512) < 256, n)
        rd = T.bool_from_masks(e["state"].read_duplicated).ravel()
        assert np.array_equal(m4[TIER_DRAM], m3[TIER_DRAM]) and np.array_equal(m4[TIER_CXL], m3[TIER_CXL])
        assert m4[TIER_HBM][first_half].all() and not m4[TIER_HBM][~first_half].any()
        assert np.array_equal(rd, first_half & (m3[TIER_DRAM] | m3[TIER_CXL]))
        # wave 5: moving the odd blocks to the CXL node for real breaks the duplication again
        rq = all_blocks(TIER_CXL, R.CAUSE_API_MIGRATE, 0, np.arange(1, n, 2))
        f = c.make_resident(rq, None, None, layout, ds, execute=True)
        m5 = check(f["state"], touched | ~(resident0 | touched), "move after read duplication")
        odd = ~even
        assert m5[TIER_CXL][odd].all() and not m5[TIER_HBM][odd].any() and not m5[TIER_DRAM][odd].any()
        assert not T.bool_from_masks(f["state"].read_duplicated).ravel()[odd].any()
        ds.free()


@pytest.mark.parametrize("seed,gpu_id,peers", [(41, 1, ((3, True),)), (42, 3, ((1, True), (2, False), (7, False), (9, True))),
                                               (43, 5, ((1, False), (2, False), (3, False), (4, False)))])
def test_make_resident_with_other_gpus_matches_oracle(seed, gpu_id, peers):
    """Tables that name other GPUs of the box (uvmtier_block_state_add_peer): sources closest first, staging through
    DRAM for the GPUs without direct access, every other holder loses the pages (tests/test_oracle_resident.py holds
    the restatement against the reference's own code for the same configurations)."""
    n_slots = 2048
    st0 = R.with_peers(_state(n_slots, seed, False), seed, peers)
    rq, pm, pf = R.random_requests(n_slots, seed, frac=0.9, bad=3, evict_cause_frac=0.3)
    want = R.run_oracle(T.load_oracle(), rq, pm, pf, st0, run_cap=1 << 20, gpu_id=gpu_id)
    layout = S.make_layout(dict(va_base=TG.VA_BASE, blocks_per_space=n_slots, n_spaces=1), gpu_id=gpu_id)
    with uvmtier.Context(0) as c:
        ds = c.block_state_alloc(layout)
        ds.upload(st0)
        got = c.make_resident(rq, pm, pf, layout, ds, run_cap=1 << 20)
        ds.free()
    R.assert_equal(want, got, "peers, seed %d" % seed)
    assert int(got["results"]["migrated_peer"].sum()) > 5000


def _moves_with_peers(c, peer_ptrs, peer_reads, peers, gpu_id, n=32):
    """Real bytes behind a table with other GPUs: every page's 4 KB pattern sits in the arena of its holder; after
    'everything to this GPU' and then 'everything to DRAM' each page holds its pattern where the table says it lives."""
    import ctypes
    lib = uvmtier.lib()
    st0 = R.with_peers(S.random_state(n, 19, populated=0.8), 19, peers, share=0.5)
    info = {t: c.tier_create(t, n << 21) for t in (TIER_HBM, TIER_DRAM, TIER_CXL)}
    views = {t: info[t].host_view().reshape(n * 512, 4096) for t in (TIER_DRAM, TIER_CXL)}
    truth = TG.rng_stream(78, n * 512 * 1024).view(np.uint8).reshape(n * 512, 4096)
    where = {TIER_HBM: T.bool_from_masks(st0.gpu_resident).ravel(), TIER_DRAM: T.bool_from_masks(st0.cpu_resident0).ravel(),
             TIER_CXL: T.bool_from_masks(st0.cpu_resident1).ravel()}
    on_peer = [T.bool_from_masks(m).ravel() for _, _, m in st0.peers]
    resident0 = where[TIER_HBM] | where[TIER_DRAM] | where[TIER_CXL]
    for m in on_peer:
        assert not (m & resident0).any()
        resident0 = resident0 | m

    def put(dptr, mask):
        buf = np.full((n * 512, 4096), 0xEE, np.uint8)
        buf[mask] = truth[mask]
        uvmtier._check(lib.uvmtier_memcpy_h2d(c.handle, uvmtier.P(dptr), uvmtier.P(buf), ctypes.c_size_t(buf.nbytes)), "h2d")

    put(info[TIER_HBM].device_ptr, where[TIER_HBM])
    for k, p in enumerate(peer_ptrs):
        put(p, on_peer[k])
    for t in (TIER_DRAM, TIER_CXL):
        views[t][:] = 0xEE
        views[t][where[t]] = truth[where[t]]
    layout = S.make_layout(dict(va_base=TG.VA_BASE, blocks_per_space=n, n_spaces=1), gpu_id=gpu_id)
    ds = c.block_state_alloc(layout)
    ds.upload(st0)
    rq = np.zeros(n, R.MAKE_RESIDENT_REQ_DT)
    rq["slot"] = np.arange(n); rq["region_outer"] = 512; rq["dest_tier"] = TIER_HBM; rq["cause"] = R.CAUSE_API_MIGRATE
    a = c.make_resident(rq, None, None, layout, ds, execute=True)
    assert [int(x) for x in a["results"]["migrated_peer"].sum(axis=0)[:len(peers)]] == [int(m.sum()) for m in on_peer]
    assert T.bool_from_masks(a["state"].gpu_resident).all() and not any(m.any() for _, _, m in a["state"].peers)
    hbm = c.from_device(info[TIER_HBM].device_ptr, n << 21, np.uint8).reshape(n * 512, 4096)
    assert np.array_equal(hbm[resident0], truth[resident0]), "bytes after gathering everything on this GPU"
    assert not hbm[~resident0].any(), "first-touched pages are zero"
    staged = np.zeros(n * 512, bool)
    for k, (_, direct) in enumerate(peers):
        if not direct:
            staged |= on_peer[k]
    assert np.array_equal(views[TIER_DRAM][staged], truth[staged]), "the staging pages on node 0 hold the first hop"
    # and out again: everything to DRAM (the source is this GPU now)
    rq["dest_tier"] = TIER_DRAM
    b = c.make_resident(rq, None, None, layout, ds, execute=True)
    assert np.array_equal(views[TIER_DRAM][resident0], truth[resident0]) and not views[TIER_DRAM][~resident0].any()
    assert int(b["results"]["migrated"][:, 2].sum()) == n * 512
    # a page on a peer moves to the CXL node without touching this GPU
    st1 = ds.download()
    st1.cpu_resident0[:] = 0
    st1.peers[0][2][:] = np.uint64(0xFFFFFFFFFFFFFFFF)
    ds.upload(st1)
    put(peer_ptrs[0], np.ones(n * 512, bool))
    rq["dest_tier"] = TIER_CXL
    d = c.make_resident(rq, None, None, layout, ds, execute=True)
    assert int(d["results"]["migrated_peer"][:, 0].sum()) == n * 512 and np.array_equal(views[TIER_CXL], truth)
    ds.free()


def test_moves_from_other_gpus_are_byte_exact_alias_arenas():
    """Two more arenas of this GPU registered as UVMTIER_TIER_PEER0 / PEER0 + 1 stand in for two other GPUs (one with
    direct copy access, one without): the same mechanism uvmtier_tier_map_peer sets up with a real peer's arena."""
    import ctypes
    n = 32
    with uvmtier.Context(0) as c:
        ptrs = [c.alloc(n << 21), c.alloc(n << 21)]
        try:
            for k, p in enumerate(ptrs):
                uvmtier._check(uvmtier.lib().uvmtier_tier_register(c.handle, ctypes.c_uint32(uvmtier.TIER_PEER0 + k), ctypes.c_void_p(p),
                                                                  None, ctypes.c_uint64(n << 21)), "tier_register")
            _moves_with_peers(c, ptrs, None, ((2, True), (6, False)), gpu_id=4, n=n)
        finally:
            for k, p in enumerate(ptrs):
                c.tier_destroy(uvmtier.TIER_PEER0 + k)
                c.free(p)


def test_moves_from_a_second_gpu_are_byte_exact():
    """The same with GPU 1's HBM arena mapped through uvmtier_tier_map_peer (pages cross NVLink).  Skipped on a one-GPU box."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    n = 32
    with uvmtier.Context(0) as c, uvmtier.Context(1) as peer:
        pinfo = peer.tier_create(TIER_HBM, n << 21)
        c.tier_map_peer(0, peer)
        try:
            _moves_with_peers(c, [pinfo.device_ptr], None, ((2, True),), gpu_id=1, n=n)
        finally:
            c.tier_destroy(uvmtier.TIER_PEER0)


def test_servicing_passes_refuse_tables_with_other_gpus():
    import ctypes
    lib = uvmtier.lib()
    pk, tbl, lay = TG.fault_trace(512, 3, "uniform", 1, total_va_bytes=64 << 21)
    gp = TG.gpu_params(256)
    layout = S.make_layout(lay)
    with uvmtier.Context(0) as c:
        c.fault_buffer_init(gp, tbl)
        ds = c.block_state_alloc(layout)
        ds.add_peer(3, True)
        one = c.alloc(4096)
        try:
            st = lib.uvmtier_service_fault_batches_device(c.handle, ctypes.c_uint32(2), uvmtier.P(one), uvmtier.P(one), uvmtier.P(one),
                                                          uvmtier.P(layout), ctypes.byref(ds.ptrs), uvmtier.P(one), uvmtier.P(one),
                                                          uvmtier.P(one), ctypes.c_uint32(1), uvmtier.P(one))
            assert st == 0x56                                    # NV_ERR_NOT_SUPPORTED
            # ids ascend, differ from this GPU's, and four is the limit
            assert lib.uvmtier_block_state_add_peer(c.handle, uvmtier.P(layout), ctypes.byref(ds.ptrs), ctypes.c_uint32(2), ctypes.c_int(1)) == 0x1F
            assert lib.uvmtier_block_state_add_peer(c.handle, uvmtier.P(layout), ctypes.byref(ds.ptrs), ctypes.c_uint32(int(layout["gpu_id"][0])),
                                                    ctypes.c_int(1)) == 0x1F
            for g in (5, 6, 7):
                ds.add_peer(g, False)
            assert lib.uvmtier_block_state_add_peer(c.handle, uvmtier.P(layout), ctypes.byref(ds.ptrs), ctypes.c_uint32(9), ctypes.c_int(1)) != 0
            assert ds.ptrs.n_peers == 4 and ds.ptrs.peer_direct == 1
        finally:
            c.free(one)
            ds.free()
