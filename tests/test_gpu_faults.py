"""GPU parity tests of the fault-preprocessing leg through the C ABI: vs the oracle, the reference-made
golden vectors, and size-independent properties at the BASELINE size (1M entries)."""
import numpy as np
import pytest

import faultlib as F
import uvmtier_testlib as T
from uvmtier import tracegen as TG

pytestmark = pytest.mark.gpu


def gpu_fetch(ctx, pk, gp, tbl, clear_valid=True):
    ctx.fault_buffer_init(gp, tbl)
    n = len(pk)
    B = int(gp["max_batch_size"][0])
    ring = np.zeros((n + 1, 8), np.uint32)
    ring[:n] = pk
    # utlbs[].num_pending_faults: optional output, one row of counters per batch
    nb, U = (n + B - 1) // B, F.utlb_count(gp)
    d_utlb = ctx.alloc(max(nb * U * 2, 2))
    ctx.fault_buffer_set_utlb_pending_output(d_utlb)
    r = ctx.fetch_preprocess_fault_batches(ring, 0, n, clear_valid=clear_valid)
    utlb = ctx.from_device(d_utlb, nb * U, "<u2").reshape(nb, U)
    ctx.fault_buffer_set_utlb_pending_output(None)
    ctx.free(d_utlb)
    assert r["get"] == n
    res = F.product_to_checker_view(n, B, tbl, r["disp"], r["ordered"], r["groups"], r["batch"], r["entries"], ring[:n].copy())
    res["utlb_pending"] = utlb
    return res


def test_gpu_matches_golden_vectors(ctx):
    import os
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "faults_golden.npz"))
    for k in range(int(g["n_cases"])):
        pre = "c%d_" % k
        pk, tbl, gp = g[pre + "packets"], g[pre + "table"], g[pre + "gpu_params"]
        want = dict(entries=g[pre + "entries"], ordered=g[pre + "ordered"], va_space=g[pre + "va_space"], gpu=g[pre + "gpu"],
                    batch=g[pre + "batch"], ring_after=g[pre + "ring_after"])
        got = gpu_fetch(ctx, pk, gp, tbl)
        F.assert_fetch_equal(want, got, "golden %d" % k)
        F.check_groups(got, tbl, int(gp["max_batch_size"][0]))


@pytest.mark.parametrize("dist", ["uniform", "stream", "hotset", "adversarial"])
@pytest.mark.parametrize("B,n_spaces", [(256, 4), (32, 1), (100, 16), (1000, 4), (1024, 2)])
def test_gpu_matches_oracle(ctx, dist, B, n_spaces):
    orc = T.load_oracle()
    pk, tbl, _ = TG.fault_trace(20011, 42, dist, n_spaces)      # ragged last batch
    gp = TG.gpu_params(B)
    want = F.run_fetch_preprocess(orc, "oracle", pk, gp, tbl)
    got = gpu_fetch(ctx, pk, gp, tbl)
    F.assert_fetch_equal(want, got, "%s B=%d" % (dist, B))
    F.check_groups(got, tbl, B)


def test_gpu_parse_matches_oracle(ctx):
    orc = T.load_oracle()
    pk, tbl, _ = TG.fault_trace(10000, 1, "uniform", 4)
    for va_bits in (57, 49):
        gp = TG.gpu_params(va_bits=va_bits)
        ctx.fault_buffer_init(gp, tbl)
        got = ctx.fault_buffer_parse_entries(pk)
        want = F.run_parse(orc, "oracle", pk, gp)
        for f in F.DECODE_FIELDS + ("flags",):
            assert np.array_equal(got[f], want[f]), f


def test_gpu_edge_cases(ctx):
    orc = T.load_oracle()
    pk, tbl, _ = TG.fault_trace(512, 9, "uniform", 4)
    gp = TG.gpu_params(256)
    # stale channel in batch 1
    bad = pk.copy(); bad[300, 1] ^= 0x5
    want = F.run_fetch_preprocess(orc, "oracle", bad, gp, tbl)
    got = gpu_fetch(ctx, bad, gp, tbl)
    assert list(got["batch"]["status"]) == [0, F.NV_WARN_MORE_PROCESSING_REQUIRED]
    F.assert_fetch_equal(want, got)
    # coalescing off
    gp0 = TG.gpu_params(256, coalesce=0)
    pk2, tbl2, _ = TG.fault_trace(700, 3, "stream", 1)
    F.assert_fetch_equal(F.run_fetch_preprocess(orc, "oracle", pk2, gp0, tbl2), gpu_fetch(ctx, pk2, gp0, tbl2))
    # fatal fault types
    pk3, tbl3, _ = TG.fault_trace(600, 5, "adversarial", 2)
    fatal = np.arange(600) % 7 == 3
    pk3[fatal, 7] = (pk3[fatal, 7] & ~np.uint32(0x1F)) | TG.HW_FAULT_TYPE_PDE_SIZE
    pk3[fatal, 3] |= 0x100
    got = gpu_fetch(ctx, pk3, gp, tbl3)
    F.assert_fetch_equal(F.run_fetch_preprocess(orc, "oracle", pk3, gp, tbl3), got)
    F.check_groups(got, tbl3, 256)
    # empty ring (get == put) and ring wrap-around
    ctx.fault_buffer_init(gp, tbl)
    ring = np.zeros((1024, 8), np.uint32)
    r = ctx.fetch_preprocess_fault_batches(ring, 5, 5)
    assert r["get"] == 5 and len(r["disp"]) == 0
    ring[900:] = pk[:124]; ring[:388] = pk[124:]
    r = ctx.fetch_preprocess_fault_batches(ring, 900, 388)
    lin = gpu_fetch(ctx, pk, gp, tbl)
    assert np.array_equal(r["batch"], lin["batch"]) and np.array_equal(r["ordered"], lin["ordered_rec"][:512])
    # valid bits of fetched entries are cleared, the rest of the ring is untouched
    assert (ring[900:, 7] >> 31 == 0).all() and (ring[:388, 7] >> 31 == 0).all()
    # bad arguments
    import ctypes
    import uvmtier
    st = uvmtier.lib().uvmtier_fault_buffer_init(ctx.handle, T.P(TG.gpu_params(20000)), T.P(tbl), ctypes.c_size_t(len(tbl)))
    assert st == 0x1F


@pytest.mark.parametrize("dist,n,B,n_spaces", [("uniform", 1 << 20, 4096, 4), ("stream", 300000, 4096, 2), ("hotset", 100000, 2048, 4),
                                               ("adversarial", 40000, 3000, 2), ("uniform", 15000, 16384, 1)])
def test_gpu_batches_above_1024_entries(ctx, dist, n, B, n_spaces):
    """batch = 4096 and the whole trace in one batch.  The strided kernel against the restatement (which
    tests/test_oracle_faults.py holds against the reference's fetch / preprocess code at these batch sizes)."""
    orc = T.load_oracle()
    pk, tbl, _ = TG.fault_trace(n, 3, dist, n_spaces)
    gp = TG.gpu_params(B)
    got = gpu_fetch(ctx, pk, gp, tbl)
    want = F.run_fetch_preprocess(orc, "oracle", pk, gp, tbl)
    F.assert_fetch_equal(want, got, "%s %d" % (dist, B))
    F.check_groups(got, tbl, B)
    assert int(got["batch"]["num_cached_faults"].sum()) == n


def test_gpu_baseline_size_1m_entries(ctx):
    """Config B: 1,048,576 entries in 256-entry batches.  Full compare with the oracle on the sorted view
    and counters, plus properties: every batch's sorted view is ordered by the reference comparator, the
    multiset of fault_cache indices equals the non-filtered set, num_instances sums to num_cached."""
    orc = T.load_oracle()
    n = 1 << 20
    for dist in ("uniform", "stream"):
        pk, tbl, _ = TG.fault_trace(n, 1, dist, 4)
        gp = TG.gpu_params(256)
        got = gpu_fetch(ctx, pk, gp, tbl)
        want = F.run_fetch_preprocess(orc, "oracle", pk, gp, tbl)
        F.assert_fetch_equal(want, got, dist)
        bt = got["batch"]
        assert int(bt["num_cached_faults"].sum()) == n
        rec = got["ordered_rec"]
        live = (got["entries"]["flags"] & F.FE_FILTERED) == 0
        assert int(live.sum()) == int(bt["num_coalesced_faults"].sum())
        # sortedness inside every batch: (va_space, gpu, address, access desc, index)
        pos = np.arange(n) % 256
        valid = pos < np.repeat(bt["num_coalesced_faults"], 256)[:n]
        k1 = (tbl["sort_rank"][np.where(valid, rec["ctx"], 0)].astype(np.uint64) << np.uint64(48)) | (rec["fault_address"] >> np.uint64(12))
        k2 = ((7 - (rec["fault_access_type"] & 7)).astype(np.uint64) << np.uint64(16)) | rec["fault_index"]
        nxt_same_batch = valid[1:] & valid[:-1] & (pos[1:] != 0)
        ok = (k1[1:] > k1[:-1]) | ((k1[1:] == k1[:-1]) & (k2[1:] > k2[:-1]))
        assert ok[nxt_same_batch].all()


@pytest.mark.parametrize("chunk_batches,get", [(0, 0), (3, 0), (1, 0), (7, 900)])
def test_streamed_fetch_from_a_pinned_host_ring_equals_the_device_form(ctx, chunk_batches, get):
    """uvmtier_fetch_preprocess_fault_batches_streamed: the ring stays in pinned host memory and crosses PCIe in
    chunks while earlier chunks are preprocessed.  Same results as the host-buffer form, for chunks that do not
    divide the trace, a single batch per chunk, and a ring segment that wraps."""
    n, B = 5000, 64
    pk, tbl, _ = TG.fault_trace(n, 17, "stream", 2)
    gp = TG.gpu_params(B)
    ctx.fault_buffer_init(gp, tbl)
    ring_entries = n + 300
    ring = np.zeros((ring_entries, 8), np.uint32)
    idx = (get + np.arange(n)) % ring_entries
    ring[idx] = pk
    put = int((get + n) % ring_entries)
    ring_a = ring.copy()
    a = ctx.fetch_preprocess_fault_batches(ring_a, get, put)
    b = ctx.fetch_preprocess_fault_batches_streamed(ring, get, put, chunk_batches=chunk_batches)
    assert a["get"] == b["get"] == put
    for k in ("disp", "batch", "entries"):
        assert np.array_equal(a[k], b[k]), k
    nb = len(a["batch"])
    for bi in range(nb):                      # slots past num_coalesced / num_block_groups are unspecified
        m, g = int(a["batch"]["num_coalesced_faults"][bi]), int(a["batch"]["num_block_groups"][bi])
        assert np.array_equal(a["ordered"][bi * B:bi * B + m], b["ordered"][bi * B:bi * B + m])
        assert np.array_equal(a["groups"][bi * B:bi * B + g], b["groups"][bi * B:bi * B + g])
    assert np.array_equal(ring_a, b["ring_device"])          # VALID bits cleared on the device copy, host ring untouched
