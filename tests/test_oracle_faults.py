"""CPU tests of the fault-preprocessing leg: the oracle restatement vs the reference's own functions
(oracle/_ref/libref_faults.so) vs golden vectors, and the product's kernel source under the SIMT emulator."""
import os

import numpy as np
import pytest

import faultlib as F
import uvmtier_testlib as T
from uvmtier import tracegen as TG

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "faults_golden.npz")


def _ref():
    ref = T.load_ref("faults")
    if ref is None:
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    return ref


def test_directed_fetch_merge_quirk():
    """READ@uTLB1, WRITE@uTLB1, READ@uTLB2 on one page.  The WRITE becomes the representative
    with num_instances 2; the batch-wide last_fault keeps pointing at the filtered first entry, which
    absorbs the third (UVM/uvm_gpu_replayable_faults.c:804-817)."""
    pk, tbl = F.directed_quirk_trace()
    gp = TG.gpu_params()
    for res in (F.run_fetch_preprocess(T.load_oracle(), "oracle", pk, gp, tbl),
                F.run_fetch_preprocess_simt(T.load_simt(), pk, gp, tbl)):
        e = res["entries"]
        assert list((e["flags"] & F.FE_FILTERED) != 0) == [True, False, True]
        assert e["num_instances"][1] == 2 and e["access_type_mask"][1] == 0b110
        assert res["batch"]["num_coalesced_faults"][0] == 1 and res["batch"]["num_cached_faults"][0] == 3
        assert res["ordered"][0] == 1


def test_oracle_matches_reference_build_directed():
    ref = _ref()
    pk, tbl = F.directed_quirk_trace()
    gp = TG.gpu_params()
    F.assert_fetch_equal(F.run_fetch_preprocess(ref, "ref", pk, gp, tbl),
                         F.run_fetch_preprocess(T.load_oracle(), "oracle", pk, gp, tbl))


@pytest.mark.parametrize("dist", ["uniform", "stream", "hotset", "adversarial"])
@pytest.mark.parametrize("n_spaces", [1, 4, 16])
def test_oracle_matches_reference_build(dist, n_spaces):
    ref, orc = _ref(), T.load_oracle()
    for B in (256, 32, 1000):
        pk, tbl, _ = TG.fault_trace(3000, 42, dist, n_spaces)
        gp = TG.gpu_params(B)
        F.assert_fetch_equal(F.run_fetch_preprocess(ref, "ref", pk, gp, tbl),
                             F.run_fetch_preprocess(orc, "oracle", pk, gp, tbl), "%s %d %d" % (dist, n_spaces, B))


@pytest.mark.parametrize("dist", ["uniform", "stream", "hotset", "adversarial"])
@pytest.mark.parametrize("B,cta", [(256, 0), (32, 0), (64, 0), (100, 0), (200, 0), (1000, 0)])
def test_simt_kernel_matches_oracle(dist, B, cta):
    """One CTA per batch, batch sizes that do and do not fill the CTA; results include utlbs[].num_pending_faults."""
    orc, sim = T.load_oracle(), T.load_simt()
    pk, tbl, _ = TG.fault_trace(1500, 0xC0FFEE, dist, 4)
    gp = TG.gpu_params(B)
    want = F.run_fetch_preprocess(orc, "oracle", pk, gp, tbl)
    got = F.run_fetch_preprocess_simt(sim, pk, gp, tbl, force_cta_kernel=cta)
    assert got["utlb_pending"].sum() == want["batch"]["num_coalesced_faults"].sum()
    F.assert_fetch_equal(want, got, "%s %d" % (dist, B))
    F.check_groups(got, tbl, B)


@pytest.mark.parametrize("dist,n,B", [("uniform", 9000, 4096), ("adversarial", 5000, 2048), ("hotset", 12000, 16384), ("stream", 7000, 3000)])
def test_batches_above_1024_entries(dist, n, B):
    """uvm_perf_fault_batch_count above a CTA's worth (the reference clamps it to the fault buffer's size, :153-165): the
    strided kernel with the sequential coalescing scan, against the restatement and against the reference's own fetch /
    preprocess code; 16384 > n is the whole trace in one batch."""
    orc, sim = T.load_oracle(), T.load_simt()
    pk, tbl, _ = TG.fault_trace(n, 5, dist, 2, total_va_bytes=64 << 21)
    gp = TG.gpu_params(B)
    want = F.run_fetch_preprocess(orc, "oracle", pk, gp, tbl)
    ref = T.load_ref("faults")
    if ref is not None:
        F.assert_fetch_equal(F.run_fetch_preprocess(ref, "ref", pk, gp, tbl), want, "reference vs restatement %s %d" % (dist, B))
    got = F.run_fetch_preprocess_simt(sim, pk, gp, tbl)
    F.assert_fetch_equal(want, got, "%s %d" % (dist, B))
    F.check_groups(got, tbl, B)
    assert len(want["batch"]) == -(-n // B)


def test_parse_decode_all_field_values():
    """F1: every client id with a uTLB, every fault type, access type, aperture; HUB clients; 49- and 57-bit VAs."""
    orc, sim, ref = T.load_oracle(), T.load_simt(), T.load_ref("faults")
    rng = np.random.default_rng(3)
    n = 4096
    client = rng.choice(TG.T1_CLIENTS.tolist() + list(range(0x08, 0x13)) + list(range(0x29, 0x35)), n)
    acc = rng.choice([0, 1, 2, 3, 4, 8, 9, 10, 11], n)
    ft = rng.integers(0, 16, n)
    addr = (rng.integers(0, 1 << 45, n, dtype=np.uint64) << np.uint64(12))
    pk = TG.encode_packets(rng.integers(1, 1 << 40, n, dtype=np.uint64) << np.uint64(12), addr, acc, rng.integers(0, 8, n),
                           client, rng.integers(0, 1 << 60, n, dtype=np.uint64), veid=rng.integers(0, 64, n), fault_type=ft,
                           inst_aperture_hw=rng.choice([0, 2, 3], n))
    hub = rng.random(n) < 0.2                       # HUB clients: CE / HOST ids with their engine ids
    pk[hub, 7] |= 1 << 20
    hub_client = rng.choice([1, 2, 6, 0x0C, 0x18, 0x20, 0x27, 0x63, 0x6E], int(hub.sum())).astype(np.uint32)
    pk[hub, 7] = (pk[hub, 7] & ~np.uint32(0x7F << 8)) | (hub_client << 8)
    # engine ids the reference asserts on: CE0..CE19 = 65..84 for CE clients, HOST0..HOST44 = 85..129 for HOST/ESC
    pk[hub, 6] = np.where(np.isin(hub_client, [6, 0x63, 0x6E]), 90, 70)
    for va_bits in (57, 49):
        gp = TG.gpu_params(va_bits=va_bits)
        o = F.run_parse(orc, "oracle", pk, gp)
        s = np.zeros(n, F.FAULT_ENTRY_DT)
        sim.simt_fault_parse(T.P(pk), n, T.P(gp), T.P(s))
        for f in F.DECODE_FIELDS + ("flags",):
            assert np.array_equal(o[f], s[f]), f
        if ref is not None:
            r = F.run_parse(ref, "ref", pk, gp)
            for f in F.DECODE_FIELDS + ("flags",):
                assert np.array_equal(o[f], r[f]), f


def test_fatal_entries_stand_alone_and_never_merge():
    orc, sim = T.load_oracle(), T.load_simt()
    pk, tbl, _ = TG.fault_trace(600, 5, "adversarial", 2)
    fatal = np.arange(600) % 7 == 3
    pk[fatal, 7] = (pk[fatal, 7] & ~np.uint32(0x1F)) | TG.HW_FAULT_TYPE_PDE_SIZE
    pk[fatal, 3] |= 0x100          # move the fatal ones to their own VA region (SURVEY: no serviceable fault shares their block)
    gp = TG.gpu_params(256)
    want = F.run_fetch_preprocess(orc, "oracle", pk, gp, tbl)
    got = F.run_fetch_preprocess_simt(sim, pk, gp, tbl)
    F.assert_fetch_equal(want, got)
    F.check_groups(got, tbl, 256)
    e = got["entries"]
    # a fatal entry is never merged INTO another one (:949); it can still be displaced by a later, more
    # intrusive fault of its uTLB on the same page (the reference does not look at the target's is_fatal)
    assert ((e["flags"][fatal] & F.FE_FATAL) != 0).all()
    rule = (got["disp"] >> 24) & 3
    assert (rule[fatal] == 0).all()
    ref = T.load_ref("faults")
    if ref is not None:
        F.assert_fetch_equal(F.run_fetch_preprocess(ref, "ref", pk, gp, tbl), want)


def test_unknown_instance_ptr_requests_refetch():
    """translate_instance_ptrs: a stale channel -> NV_WARN_MORE_PROCESSING_REQUIRED for the batch (:1097-1107)."""
    orc, sim = T.load_oracle(), T.load_simt()
    pk, tbl, _ = TG.fault_trace(512, 9, "uniform", 4)
    pk[300, 1] ^= 0x5                      # corrupt one instance pointer in the second batch
    gp = TG.gpu_params(256)
    want = F.run_fetch_preprocess(orc, "oracle", pk, gp, tbl)
    got = F.run_fetch_preprocess_simt(sim, pk, gp, tbl)
    assert list(want["batch"]["status"]) == [0, F.NV_WARN_MORE_PROCESSING_REQUIRED]
    F.assert_fetch_equal(want, got)
    ref = T.load_ref("faults")
    if ref is not None:
        F.assert_fetch_equal(F.run_fetch_preprocess(ref, "ref", pk, gp, tbl), want)


def test_invalid_entry_ends_the_batch_and_coalesce_off():
    orc, sim = T.load_oracle(), T.load_simt()
    pk, tbl, _ = TG.fault_trace(256, 11, "stream", 1)
    gp = TG.gpu_params(256, coalesce=0)
    want = F.run_fetch_preprocess(orc, "oracle", pk, gp, tbl)
    got = F.run_fetch_preprocess_simt(sim, pk, gp, tbl)
    F.assert_fetch_equal(want, got)
    assert want["batch"]["num_coalesced_faults"][0] == 256
    # first batch stops at the first packet whose VALID bit is clear (BATCH_READY, :899-903)
    pk2 = pk[:200].copy()
    pk2[137, 7] &= 0x7FFFFFFF
    gp = TG.gpu_params(256)
    got = F.run_fetch_preprocess_simt(sim, pk2, gp, tbl)
    assert got["batch"]["num_cached_faults"][0] == 137
    want = F.run_fetch_preprocess(orc, "oracle", pk2[:137].copy(), gp, tbl)
    for f in ("num_cached_faults", "num_coalesced_faults", "is_single_instance_ptr"):
        assert got["batch"][f][0] == want["batch"][f][0]
    assert np.array_equal(got["ordered"][:137], want["ordered"][:137])


def test_oracle_and_simt_match_golden_vectors():
    g = np.load(GOLDEN)
    orc, sim = T.load_oracle(), T.load_simt()
    for k in range(int(g["n_cases"])):
        pre = "c%d_" % k
        pk, tbl, gp = g[pre + "packets"], g[pre + "table"], g[pre + "gpu_params"]
        want = dict(entries=g[pre + "entries"], ordered=g[pre + "ordered"], va_space=g[pre + "va_space"], gpu=g[pre + "gpu"],
                    batch=g[pre + "batch"], ring_after=g[pre + "ring_after"])
        F.assert_fetch_equal(want, F.run_fetch_preprocess(orc, "oracle", pk, gp, tbl), "golden %d oracle" % k)
        F.assert_fetch_equal(want, F.run_fetch_preprocess_simt(sim, pk, gp, tbl), "golden %d simt" % k)


def test_numpy_kiss_stream_is_the_reference_generator():
    """tracegen.rng_stream (numpy, lanes + jump-ahead) against a line-by-line loop of uvm_test_rng_32
    (UVM/uvm_test_rng.c:37-62) and against the library's C routine, lengths around the lane boundaries."""
    import ctypes
    import uvmtier

    def loop(seed, n):
        z, w, jcong, jsr, M = 362436069, 521288629, 380116160, seed & 0xFFFFFFFF, 0xFFFFFFFF
        out = []
        for _ in range(n):
            z = (36969 * (z & 65535) + (z >> 16)) & M
            w = (18000 * (w & 65535) + (w >> 16)) & M
            jcong = (69069 * jcong + 1234567) & M
            jsr ^= (jsr << 17) & M
            jsr ^= jsr >> 13
            jsr ^= (jsr << 5) & M
            out.append((((((z << 16) + w) & M) ^ jcong) + jsr) & M)
        return np.array(out, np.uint32)

    for seed, n in ((1, 1), (42, 3000), (0xC0FFEE, 70001)):
        assert np.array_equal(TG.rng_stream(seed, n), loop(seed, n)), (seed, n)
    for seed, n in ((7, 65536), (7, 65537), (123456789, 1 << 20)):
        c = np.empty(n, np.uint32)
        uvmtier.lib().uvmtier_test_rng_fill(ctypes.c_uint32(seed), ctypes.c_size_t(n), T.P(c))
        assert np.array_equal(TG.rng_stream(seed, n), c), (seed, n)
