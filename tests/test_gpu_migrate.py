"""GPU tests of the migration leg (K3): tiers, SM-driven page-list copies, byte-exactness.
Contents are opaque bytes: the oracle for a copy is numpy fancy indexing (the reference's CPU path is a
memcpy page loop, UVM/uvm_va_block.c:3972-3984); checks are full compares plus per-page checksums."""
import numpy as np
import pytest

import uvmtier
from uvmtier import TIER_CXL, TIER_DRAM, TIER_HBM, tracegen as TG

pytestmark = pytest.mark.gpu

ARENA = 64 << 20


def np_checksum(buf, offs, page_size):
    """numpy mirror of k_checksum_pages."""
    out = np.zeros(len(offs), np.uint64)
    idx = np.arange(1, page_size // 8 + 1, dtype=np.uint64)
    with np.errstate(over="ignore"):
        for k, o in enumerate(offs):
            w = buf[int(o):int(o) + page_size].view("<u8")
            v = w + np.uint64(0x9E3779B97F4A7C15) * idx
            v ^= v >> np.uint64(29)
            v *= np.uint64(0xBF58476D1CE4E5B9)
            v ^= v >> np.uint64(32)
            out[k] = v.sum(dtype=np.uint64)
    return out


@pytest.fixture(scope="module")
def tiers():
    c = uvmtier.Context(0)
    info = {t: c.tier_create(t, ARENA) for t in (TIER_HBM, TIER_DRAM, TIER_CXL)}
    yield c, info
    c.close()


def fill(info, tier, seed, pattern="rng"):
    v = info[tier].host_view()
    if pattern == "rng":
        v[:] = TG.rng_stream(seed, len(v) // 4).view(np.uint8)
    else:                       # the fork's byte pattern, tests/cxl_p2p_test.c:467-476: p[i] = (i + seed) & 0xFF
        v[:] = ((np.arange(len(v), dtype=np.uint64) + np.uint64(seed)) & np.uint64(0xFF)).astype(np.uint8)
    return v


def hbm_fill(c, info, value, nbytes):
    import ctypes
    uvmtier._check(uvmtier.lib().uvmtier_memset_device(c.handle, ctypes.c_void_p(info[TIER_HBM].device_ptr), ctypes.c_int(value),
                                                       ctypes.c_size_t(nbytes)), "memset")
    c.synchronize()


def hbm_read(c, info, nbytes=ARENA):
    return c.from_device(info[TIER_HBM].device_ptr, nbytes, np.uint8)


def test_numa_probe_and_cxl_labelling(tiers):
    c, info = tiers
    n, dram, cxl = uvmtier.numa_probe()
    assert n >= 1 and dram >= 0
    emulated = bool(info[TIER_CXL].flags & uvmtier.TIERF_CXL_EMULATED)
    assert emulated == (cxl < 0)       # a machine without a CPU-less node must say "CXL-emulated"
    assert info[TIER_DRAM].host_ptr and info[TIER_CXL].host_ptr and not info[TIER_HBM].host_ptr


@pytest.mark.parametrize("page_size", [4096, 65536, 2 << 20])
@pytest.mark.parametrize("order", ["sequential", "permutation", "sparse"])
@pytest.mark.parametrize("method", [0, 1])
def test_host_to_hbm_and_back_byte_exact(tiers, page_size, order, method):
    c, info = tiers
    n_pages = ARENA // page_size
    rng = np.random.default_rng(page_size + len(order))
    pages = np.arange(n_pages)
    if order == "permutation":
        dst_pages = rng.permutation(n_pages)
    elif order == "sparse":
        pages = np.sort(rng.choice(n_pages, max(1, n_pages // 2), replace=False))
        dst_pages = pages
    else:
        dst_pages = pages
    src_off = pages.astype(np.uint64) * np.uint64(page_size)
    dst_off = dst_pages.astype(np.uint64) * np.uint64(page_size)
    for src_tier, pat in ((TIER_DRAM, "rng"), (TIER_CXL, "fork")):
        src = fill(info, src_tier, 42 + page_size, pat)
        hbm_fill(c, info, 0xEE, ARENA)
        c.copy_page_list(TIER_HBM, src_tier, dst_off, src_off, page_size, method)
        got = hbm_read(c, info)
        want = np.full(ARENA, 0xEE, np.uint8)
        sv = src.reshape(-1, page_size)
        want.reshape(-1, page_size)[dst_pages] = sv[pages]
        assert np.array_equal(got, want), "host->HBM %s" % order
        # checksums of the migrated pages agree with the source pages (and with the numpy mirror on a sample)
        cs_dst = c.checksum_pages(TIER_HBM, dst_off, page_size)
        cs_src = c.checksum_pages(src_tier, src_off, page_size)
        assert np.array_equal(cs_dst, cs_src)
        assert np.array_equal(cs_src[:3], np_checksum(src, src_off[:3], page_size))
        # and back into the other host tier: HBM -> host, byte exact
        back_tier = TIER_CXL if src_tier == TIER_DRAM else TIER_DRAM
        back = info[back_tier].host_view()
        back[:] = 0x11
        c.copy_page_list(back_tier, TIER_HBM, src_off, dst_off, page_size, method)
        wantb = np.full(ARENA, 0x11, np.uint8)
        wantb.reshape(-1, page_size)[pages] = sv[pages]
        assert np.array_equal(back, wantb), "HBM->host %s" % order


def test_dram_to_cxl_copy_through_the_gpu(tiers):
    """Config A: 64 MB DRAM -> CXL NUMA node, 4 KB pages (16384 pages)."""
    c, info = tiers
    src = fill(info, TIER_DRAM, 7, "fork")
    dst = info[TIER_CXL].host_view()
    dst[:] = 0
    off = np.arange(ARENA // 4096, dtype=np.uint64) * np.uint64(4096)
    c.copy_page_list(TIER_CXL, TIER_DRAM, off, off, 4096)
    assert np.array_equal(dst, src)


def test_edge_cases(tiers):
    c, info = tiers
    # empty list
    c.copy_page_list(TIER_HBM, TIER_DRAM, np.zeros(0, "<u8"), np.zeros(0, "<u8"), 4096)
    # one page, and a count that is not a multiple of the stage size
    src = fill(info, TIER_DRAM, 3)
    for n in (1, 5, 4 * 12 * 148 + 3):
        off = np.arange(n, dtype=np.uint64) * np.uint64(4096)
        hbm_fill(c, info, 0, ARENA)
        c.copy_page_list(TIER_HBM, TIER_DRAM, off, off, 4096)
        got = hbm_read(c, info, (n + 1) * 4096)
        assert np.array_equal(got[:n * 4096], src[:n * 4096]) and not got[n * 4096:].any()
    # bad page size / unknown tier
    import ctypes
    z = np.zeros(1, "<u8")
    lib = uvmtier.lib()
    assert lib.uvmtier_copy_page_list(c.handle, 0, 1, ctypes.c_size_t(1), uvmtier.P(z), uvmtier.P(z), 8192, 0) == 0x1F
    assert lib.uvmtier_copy_page_list(c.handle, 5, 1, ctypes.c_size_t(1), uvmtier.P(z), uvmtier.P(z), 4096, 0) == 0x40


def test_copy_runs_from_the_service_pass_move_the_right_bytes():
    """End to end on one GPU: faults -> service -> copy runs -> bytes in HBM.  Every page the service pass made
    resident on the GPU holds the bytes of the CPU tier it came from (or zeros for first touch)."""
    import faultlib as F
    import servicelib as S
    import uvmtier_testlib as T
    n_blocks = 32
    with uvmtier.Context(0) as c:
        info = {t: c.tier_create(t, n_blocks << 21) for t in (TIER_HBM, TIER_DRAM, TIER_CXL)}
        dram = fill(info, TIER_DRAM, 11); cxl = fill(info, TIER_CXL, 12, "fork")
        hbm_fill(c, info, 0xAB, n_blocks << 21)
        pk, tbl, lay = TG.fault_trace(6000, 31, "uniform", 1, total_va_bytes=n_blocks << 21)
        gp = TG.gpu_params(256)
        st0 = S.random_state(n_blocks, 2, populated=0.8, ro_frac=0, none_frac=0)
        st0.gpu_resident[:] = 0; st0.gpu_pte0[:] = 0; st0.gpu_pte1[:] = 0; st0.gpu_pte2[:] = 0
        c.fault_buffer_init(gp, tbl)
        ring = np.zeros((len(pk) + 1, 8), np.uint32); ring[:len(pk)] = pk
        r = c.fetch_preprocess_fault_batches(ring, 0, len(pk))
        nb = len(r["batch"]); pad = nb * 256 - len(pk)
        r["ordered"] = np.concatenate([r["ordered"], np.zeros(pad, r["ordered"].dtype)])
        r["groups"] = np.concatenate([r["groups"], np.zeros(pad, r["groups"].dtype)])
        layout = S.make_layout(lay)
        ds = c.block_state_alloc(layout); ds.upload(st0)
        run_cap = 1 << 16
        got = c.service_fault_batches(r, layout, ds, run_cap)
        assert got["n_runs"] <= run_cap
        d_runs = c.to_device(got["runs"]); d_n = c.to_device(np.array([got["n_runs"]], "<u4"))
        import ctypes
        uvmtier._check(uvmtier.lib().uvmtier_copy_runs_device(c.handle, uvmtier.P(d_runs), uvmtier.P(d_n), ctypes.c_uint32(run_cap), ctypes.c_int(0)), "copy_runs")
        c.synchronize()
        hbm = c.from_device(info[TIER_HBM].device_ptr, n_blocks << 21, np.uint8).reshape(n_blocks * 512, 4096)
        from0 = T.bool_from_masks(st0.cpu_resident0) & T.bool_from_masks(got["state"].gpu_resident)
        from1 = T.bool_from_masks(st0.cpu_resident1) & T.bool_from_masks(got["state"].gpu_resident)
        touched = T.bool_from_masks(got["state"].gpu_resident) & ~from0 & ~from1
        untouched = ~T.bool_from_masks(got["state"].gpu_resident)
        d = dram.reshape(-1, 4096); x = cxl.reshape(-1, 4096)
        assert np.array_equal(hbm[from0.ravel()], d[from0.ravel()])
        assert np.array_equal(hbm[from1.ravel()], x[from1.ravel()])
        assert not hbm[touched.ravel()].any()
        assert (hbm[untouched.ravel()] == 0xAB).all()
        assert from0.sum() > 100 and from1.sum() > 30 and touched.sum() > 30
        ds.free()


def _peer_roundtrip(c, info, peer_tier, peer_read, page_size, method):
    """HBM -> peer tier (push: this GPU's SMs store into the peer arena, uvm_va_block.c:3427-3457) with the pages
    shuffled, then peer -> HBM (pull) back into place; both arenas compared byte for byte with numpy indexing."""
    import ctypes
    n = ARENA // page_size
    rng = np.random.default_rng(page_size + method)
    src = TG.rng_stream(77 + method, ARENA // 4).view(np.uint8)
    uvmtier._check(uvmtier.lib().uvmtier_memcpy_h2d(c.handle, ctypes.c_void_p(info[TIER_HBM].device_ptr), T_P(src),
                                                   ctypes.c_size_t(ARENA)), "h2d")
    perm = rng.permutation(n).astype(np.uint64) * np.uint64(page_size)
    lin = np.arange(n, dtype=np.uint64) * np.uint64(page_size)
    c.copy_page_list(peer_tier, TIER_HBM, perm, lin, page_size, method)          # push
    landed = peer_read()
    want = np.empty_like(src)
    want.reshape(n, page_size)[(perm // np.uint64(page_size)).astype(np.int64)] = src.reshape(n, page_size)
    assert np.array_equal(landed, want), "push moved wrong bytes"
    assert np.array_equal(c.checksum_pages(peer_tier, perm, page_size), np_checksum(src, lin, page_size))
    hbm_fill(c, info, 0xEE, ARENA)
    c.copy_page_list(TIER_HBM, peer_tier, lin, perm, page_size, method)          # pull, back into place
    assert np.array_equal(hbm_read(c, info), src), "pull moved wrong bytes"


def T_P(a):
    import uvmtier_testlib as T
    return T.P(a)


@pytest.mark.parametrize("page_size", [4096, 65536, 2 << 20])
@pytest.mark.parametrize("method", [0, 1])
def test_peer_tier_on_one_gpu_byte_exact(tiers, page_size, method):
    """The peer-tier mechanism on a single GPU: a second arena of the same GPU registered as UVMTIER_TIER_PEER0
    through uvmtier_tier_register (what uvmtier_tier_map_peer / _ipc_import do with another GPU's arena)."""
    import ctypes
    c, info = tiers
    d_peer = c.alloc(ARENA)
    try:
        uvmtier._check(uvmtier.lib().uvmtier_tier_register(c.handle, ctypes.c_uint32(uvmtier.TIER_PEER0), ctypes.c_void_p(d_peer), None,
                                                          ctypes.c_uint64(ARENA)), "tier_register")
        _peer_roundtrip(c, info, uvmtier.TIER_PEER0, lambda: c.from_device(d_peer, ARENA, np.uint8), page_size, method)
    finally:
        c.tier_destroy(uvmtier.TIER_PEER0)
        c.free(d_peer)


@pytest.mark.parametrize("page_size", [4096, 2 << 20])
def test_peer_tier_on_a_second_gpu_byte_exact(tiers, page_size):
    """Two GPUs in one process: uvmtier_tier_map_peer maps GPU 1's HBM arena as this context's peer tier; pages
    cross NVLink by SM stores (push) and SM loads (pull).  Skipped on a one-GPU box."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    c, info = tiers
    peer = uvmtier.Context(1)
    try:
        pinfo = peer.tier_create(TIER_HBM, ARENA)
        c.tier_map_peer(0, peer)
        _peer_roundtrip(c, info, uvmtier.TIER_PEER0, lambda: peer.from_device(pinfo.device_ptr, ARENA, np.uint8), page_size, 0)
    finally:
        c.tier_destroy(uvmtier.TIER_PEER0)
        peer.close()
