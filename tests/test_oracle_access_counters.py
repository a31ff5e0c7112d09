"""CPU tests of the access-counter notification path (section 1e of include/uvmtier.h): the product's kernels under
the SIMT emulator vs the CPU restatement (oracle/oracle_access_counters.c), and directed cases for the reference's
rules (fetch stops at the first entry that is not ready, untranslated notifications are ignored, expansion to
sub-granularity regions vs single pages, stale notifications, blocks the GPU no longer maps)."""
import numpy as np
import pytest

import aclib as A
import servicelib as S
import uvmtier_testlib as T
from uvmtier import tracegen as TG

MB2 = 2 << 20


def notifications(n, seed, n_spaces, blocks, hot=0.5):
    """Random notifications over `blocks` va_blocks per VA space: 2 MB-aligned region addresses (the tracking
    granularity), random sub-granularity masks, some instance pointers that belong to nobody."""
    rng = np.random.default_rng(seed)
    tbl = TG.instance_ptr_table(n_spaces)
    lay = dict(va_base=TG.VA_BASE, blocks_per_space=blocks, n_spaces=n_spaces)
    sp = rng.integers(0, n_spaces, n)
    blk = np.where(rng.random(n) < hot, rng.integers(0, max(blocks // 8, 1), n), rng.integers(0, blocks + 2, n))   # a few outside
    addr = (TG.VA_BASE + blk.astype(np.uint64) * np.uint64(MB2)).astype(np.uint64)
    addr += (rng.integers(0, 4, n).astype(np.uint64) * np.uint64(64 << 10)) * (rng.random(n) < 0.2)   # not always the block base
    inst = tbl["instance_ptr"][sp].copy()
    inst[rng.random(n) < 0.03] = 0xABCD000                       # stale channel: does not translate
    sub = rng.integers(0, 1 << 32, n, dtype=np.uint64).astype(np.uint32) & rng.integers(0, 1 << 32, n, dtype=np.uint64).astype(np.uint32)
    pk = A.encode_notifications(inst, addr, sub, ve_id=rng.integers(0, 4, n), counter=rng.integers(1, 65536, n))
    return pk, tbl, lay


def state_with_mappings(n_slots, seed):
    """Random residency in which the GPU maps pages of most blocks, some of them remotely (pages that live on the CPU):
    those are what access counters report."""
    st = S.random_state(n_slots, seed, populated=0.8)
    rng = np.random.default_rng(seed + 9)
    cpu = T.bool_from_masks(st.cpu_resident0) | T.bool_from_masks(st.cpu_resident1)
    remote = cpu & (rng.random(cpu.shape) < 0.3) & (st.logical_prot != 0)[:, None]
    st.gpu_pte0 |= T.masks_from_bool(remote)
    st.gpu_pte0[rng.random(n_slots) < 0.1] = 0                   # blocks the GPU does not map at all any more
    st.gpu_pte1 &= st.gpu_pte0
    st.gpu_pte2 &= st.gpu_pte1
    return st


@pytest.mark.parametrize("B,n,get,wrap", [(64, 1000, 0, 0), (256, 700, 0, 0), (32, 333, 50, 1), (128, 128, 0, 0)])
def test_simt_fetch_preprocess_matches_oracle(B, n, get, wrap):
    pk, tbl, _ = notifications(n, 3 + B, 3, 40)
    pk[n // 2 + 7, 7] &= 0x7FFFFFFF                              # one entry is not ready: its batch stops there
    ring_entries = n + 100 if not wrap else n + 20
    ring = np.zeros((ring_entries, 8), np.uint32)
    ring[(get + np.arange(n)) % ring_entries] = pk
    put = (get + n) % ring_entries
    want = A.run_preprocess(T.load_oracle(), "oracle_access_counters_fetch_preprocess", ring, get, put, B, tbl)
    got = A.run_preprocess(T.load_simt(), "simt_access_counters_fetch_preprocess", ring, get, put, B, tbl)
    A.assert_preprocess_equal(want, got, B)
    assert int(want["batch"]["num_cached_faults"].sum()) < n and (want["ordered"]["ctx"] == 0xFFFF).any()
    # sortedness inside every batch: untranslated first, then (rank, address), ties in cache order
    for b in range(len(want["batch"])):
        o = want["ordered"][b * B:b * B + int(want["batch"]["num_cached_faults"][b])]
        rank = np.where(o["ctx"] == 0xFFFF, 0, tbl["sort_rank"][np.minimum(o["ctx"], len(tbl) - 1)] + 1)
        key = list(zip(rank.tolist(), (o["address"] >> 12).tolist(), o["index"].tolist()))
        assert key == sorted(key)


@pytest.mark.parametrize("cpu_chunk_2m,granularity_2m,heavy_min", [(True, True, 96), (False, True, 96), (True, False, 96), (True, True, 2)])
def test_simt_service_matches_oracle(cpu_chunk_2m, granularity_2m, heavy_min):
    B = 64
    pk, tbl, lay = notifications(1500, 11, 2, 48)
    n = len(pk)
    ring = np.zeros((n + 1, 8), np.uint32)
    ring[:n] = pk
    orc, sim = T.load_oracle(), T.load_simt()
    k = A.run_preprocess(orc, "oracle_access_counters_fetch_preprocess", ring, 0, n, B, tbl)
    layout = S.make_layout(lay)
    st0 = state_with_mappings(lay["blocks_per_space"] * 2, 21)
    want = A.run_service_oracle(orc, k, B, tbl, layout, st0, granularity_2m=granularity_2m, cpu_chunk_2m=cpu_chunk_2m)
    got = A.run_service_simt(sim, k, B, tbl, layout, st0, granularity_2m=granularity_2m, cpu_chunk_2m=cpu_chunk_2m, heavy_min=heavy_min)
    S.assert_service_equal(want, got, B, "access counters")
    moved = int(want["records"]["page_fault_count"].sum())
    if granularity_2m:
        assert moved > (2000 if cpu_chunk_2m else 50) and want["n_runs"] > 0
        gb, ga = T.bool_from_masks(st0.gpu_resident), T.bool_from_masks(want["state"].gpu_resident)
        assert not (gb & ~ga).any() and int((ga & ~gb).sum()) >= moved
    else:
        assert moved == 0                                        # other tracking granularities are ignored (:1303-1306)
        want["state"].assert_equal(st0)


# ------------------------------------------------------------------------------------------------
# The restatement against the reference's own expand_notification_block + service_va_block_locked (compiled verbatim
# into oracle/_ref/libref_access_counters.so).
@pytest.mark.parametrize("seed,cpu_chunk_2m,granularity_2m", [(1, True, True), (2, True, True), (3, False, True), (4, False, True),
                                                              (5, True, False)])
def test_oracle_matches_the_reference_expansion_and_service(seed, cpu_chunk_2m, granularity_2m):
    ref = T.load_ref("access_counters")
    if ref is None:
        pytest.skip("oracle/_ref/libref_access_counters.so not built (needs /root/reference)")
    B = 128
    pk, tbl, lay = notifications(2500, 40 + seed, 2, 64)
    n = len(pk)
    ring = np.zeros((n + 1, 8), np.uint32)
    ring[:n] = pk
    orc = T.load_oracle()
    k = A.run_preprocess(orc, "oracle_access_counters_fetch_preprocess", ring, 0, n, B, tbl)
    layout = S.make_layout(lay)
    st0 = state_with_mappings(lay["blocks_per_space"] * 2, 50 + seed)
    want = A.run_service_ref(ref, k, B, tbl, layout, st0, granularity_2m=granularity_2m, cpu_chunk_2m=cpu_chunk_2m)
    got = A.run_service_oracle(orc, k, B, tbl, layout, st0, granularity_2m=granularity_2m, cpu_chunk_2m=cpu_chunk_2m)
    want["state"].assert_equal(got["state"], "reference vs restatement")
    assert len(want["records"]) == len(got["records"]) and want["n_runs"] == got["n_runs"]
    for f in S.RECORD_FIELDS:
        assert np.array_equal(want["records"][f], got["records"][f]), f
    assert np.array_equal(want["runs"][:want["n_runs"]], got["runs"][:got["n_runs"]])
    if granularity_2m:
        assert int(want["records"]["page_fault_count"].sum()) > 50


@pytest.mark.parametrize("B,n,get", [(64, 1000, 0), (256, 700, 0), (32, 333, 50), (128, 1280, 0)])
def test_oracle_matches_the_reference_parse_fetch_and_sort(B, n, get):
    """The reference's own parser, fetch loop and preprocess_notifications (compiled verbatim into
    oracle/_ref/libref_access_counters_front.so) against the restatement: parsed entries, sorted order, runs, counters,
    cleared VALID bits."""
    ref = T.load_ref("access_counters_front")
    if ref is None:
        pytest.skip("oracle/_ref/libref_access_counters_front.so not built (needs /root/reference)")
    pk, tbl, _ = notifications(n, 70 + B, 3, 40)
    pk[n // 2 + 7, 7] &= 0x7FFFFFFF                              # an entry that is not ready
    pk[B * 2, 7] &= 0x7FFFFFFF                                   # and one at the very start of a batch
    # upper-half (sign-extended) addresses and a system-memory instance block, for the parser
    pk[5, 3] |= 1 << 24
    pk[9, 0] |= 2 << 10
    ring_entries = n + 20
    ring = np.zeros((ring_entries, 8), np.uint32)
    ring[(get + np.arange(n)) % ring_entries] = pk
    put = (get + n) % ring_entries
    for clear_valid in (True, False):
        want = A.run_preprocess(ref, "ref_access_counters_fetch_preprocess", ring, get, put, B, tbl, clear_valid=clear_valid)
        got = A.run_preprocess(T.load_oracle(), "oracle_access_counters_fetch_preprocess", ring, get, put, B, tbl, clear_valid=clear_valid)
        fetched = want["entries"]["valid"] == 1
        assert np.array_equal(fetched, got["entries"]["valid"] == 1) and fetched.sum() > n // 2
        got["entries"][~fetched] = 0                             # the reference harness fills fetched entries only
        # Notifications that compare equal (same va_space, gpu and address) have no defined order: the kernel's sort()
        # is not stable.  The reference harness's stable sort leaves them in the order of the first sort (by instance
        # pointer and ve_id, :983-987), the library keeps cache order; nothing downstream depends on it (a run's
        # notifications are OR-ed into one mask).  Compare with ties put in cache order on both sides.
        for k in (want, got):
            o = k["ordered"]
            for bi in range(len(k["batch"])):
                m = int(k["batch"]["num_cached_faults"][bi])
                seg = o[bi * B:bi * B + m]
                pos = np.arange(m)
                first = np.r_[True, (seg["ctx"][1:] != seg["ctx"][:-1]) | (seg["address"][1:] != seg["address"][:-1])] if m else pos
                grp = np.cumsum(first) if m else pos
                o[bi * B:bi * B + m] = seg[np.lexsort((seg["index"], grp))]
        A.assert_preprocess_equal(want, got, B)
