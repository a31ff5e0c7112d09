"""CPU tests of the servicing pass (stage C/D/E): the product's kernels under the SIMT emulator vs the oracle,
plus directed cases for the reference's rules (duplicates, permissions, already-authorised, unmapped addresses,
first touch, batch order across runs of the same va_block)."""
import numpy as np
import pytest

import faultlib as F
import servicelib as S
import uvmtier_testlib as T
from uvmtier import tracegen as TG

SMALL_VA = 64 << 21      # 64 va_blocks in total: many batches hit the same block


def _k1(pk, gp, tbl):
    return F.run_fetch_preprocess_simt(T.load_simt(), pk, gp, tbl)


NONE_HEAVY, ALL_HEAVY = 0xFFFFFFFF, 0


def _both(pk, gp, tbl, lay, st0, pf=(1, 51, 1), **simt_args):
    k1 = _k1(pk, gp, tbl)
    layout = S.make_layout(lay)
    want = S.run_service_oracle(T.load_oracle(), k1, gp, tbl, layout, st0, pf)
    got = S.run_service_simt(T.load_simt(), k1, gp, tbl, layout, st0, pf, **simt_args)
    return k1, want, got


@pytest.mark.parametrize("dist", ["uniform", "stream", "hotset", "adversarial"])
@pytest.mark.parametrize("B", [256, 48])
def test_simt_service_matches_oracle(dist, B):
    """Default split: va_blocks with few entries on a group of four lanes each, the others on a warp each."""
    pk, tbl, lay = TG.fault_trace(1200, 21, dist, 2, total_va_bytes=SMALL_VA)
    gp = TG.gpu_params(B)
    st0 = S.random_state(lay["blocks_per_space"] * 2, 4)
    _, want, got = _both(pk, gp, tbl, lay, st0)
    S.assert_service_equal(want, got, B, "%s %d" % (dist, B))
    S.check_service_invariants(got, st0)


@pytest.mark.parametrize("dist", ["uniform", "hotset", "adversarial"])
@pytest.mark.parametrize("heavy_min", [NONE_HEAVY, ALL_HEAVY, 24])
def test_every_kernel_variant_matches_oracle(dist, heavy_min):
    """The lane-group kernel on its own, the warp-per-block kernel on its own (entries filtered 32 at a
    time), and a split in the middle of this trace's weights."""
    pk, tbl, lay = TG.fault_trace(900, 33, dist, 2, total_va_bytes=SMALL_VA)
    gp = TG.gpu_params(128)
    st0 = S.random_state(lay["blocks_per_space"] * 2, 5)
    _, want, got = _both(pk, gp, tbl, lay, st0, heavy_min=heavy_min)
    S.assert_service_equal(want, got, 128, "%s heavy_min %d" % (dist, heavy_min))


@pytest.mark.parametrize("pf", [(1, 0, 1), (1, 1, 1), (1, 25, 1), (1, 50, 1), (1, 75, 3), (1, 99, 1), (1, 100, 1), (1, 51, 20), (0, 51, 1)])
@pytest.mark.parametrize("heavy_min", [NONE_HEAVY, ALL_HEAVY])
def test_prefetch_tunables_on_both_kernels(pf, heavy_min):
    """Prefetch threshold 0..100, min_faults, prefetcher off: the lane-group kernel decides the levels inside a word from its
    page count (filled big pages pass every level below them unless the threshold is 100), the
    warp kernel evaluates every node; both must agree with the reference's leaf-to-root walk."""
    pk, tbl, lay = TG.fault_trace(1500, 57, "uniform", 2, total_va_bytes=SMALL_VA)
    gp = TG.gpu_params(64)
    st0 = S.random_state(lay["blocks_per_space"] * 2, 6, populated=0.55)
    _, want, got = _both(pk, gp, tbl, lay, st0, pf=pf, heavy_min=heavy_min)
    S.assert_service_equal(want, got, 64, "pf %s heavy_min %d" % (pf, heavy_min))
    S.check_service_invariants(got, st0)


def test_long_chains_of_duplicates_in_the_entry_parallel_filter():
    """Many uTLBs faulting on the same few pages, read-only VMAs and prefetch accesses mixed in: duplicates inherit
    'fatal' and 'invalid prefetch' along chains that cross the 32-entry chunks of the warp kernel."""
    rng = np.random.default_rng(77)
    n = 700
    tbl = TG.instance_ptr_table(1)
    lay = dict(va_base=TG.VA_BASE, blocks_per_space=4, n_spaces=1)
    page = rng.integers(0, 6, n) * 37
    blk = rng.integers(0, 4, n)
    addr = (TG.VA_BASE + (blk.astype(np.uint64) << np.uint64(21)) + (page.astype(np.uint64) << np.uint64(12))).astype(np.uint64)
    acc = rng.choice([TG.HW_VIRT_READ, TG.HW_VIRT_WRITE, TG.HW_VIRT_ATOMIC_STRONG, TG.HW_VIRT_PREFETCH], n)
    pk = TG.encode_packets(np.full(n, tbl["instance_ptr"][0]), addr, acc, rng.integers(0, 8, n),
                           [TG.T1_CLIENTS[i] for i in rng.integers(0, 8, n)], np.arange(n) + 1)
    gp = TG.gpu_params(512)
    st0 = S.random_state(4, 3, populated=0.4)
    st0.logical_prot[:] = [3, 1, 1, 0]
    st0.gpu_pte1[1:] = 0
    st0.gpu_pte2[1:] = 0
    st0.gpu_pte0[3] = 0
    for heavy_min in (ALL_HEAVY, NONE_HEAVY):
        k1, want, got = _both(pk, gp, tbl, lay, st0, heavy_min=heavy_min)
        S.assert_service_equal(want, got, 512, "heavy_min %d" % heavy_min)
    assert int((want["flags"] & 2 != 0).sum()) > 200 and int((want["flags"] & 4 != 0).sum()) > 20


def test_block_touched_by_more_than_32_batches_is_applied_in_batch_order():
    pk, tbl, lay = TG.fault_trace(1600, 3, "hotset", 1, total_va_bytes=8 << 21)
    gp = TG.gpu_params(32)                                   # 50 batches over 8 blocks
    st0 = S.random_state(8, 9, populated=0.5, ro_frac=0, none_frac=0)
    _, want, got = _both(pk, gp, tbl, lay, st0)
    per_slot = np.bincount(want["records"]["slot"], minlength=8)
    assert per_slot.max() > 32
    S.assert_service_equal(want, got, 32)


def test_directed_rules():
    """One va_block, hand-built faults: duplicate counting, read-only VMA, already mapped page, first touch."""
    tbl = TG.instance_ptr_table(1)
    lay = dict(va_base=TG.VA_BASE, blocks_per_space=2, n_spaces=1)
    A = lambda page, blk=0: TG.VA_BASE + (blk << 21) + (page << 12)
    R, W, AT, PF = TG.HW_VIRT_READ, TG.HW_VIRT_WRITE, TG.HW_VIRT_ATOMIC_STRONG, TG.HW_VIRT_PREFETCH
    #            page  access  gpc client
    rows = [(5, R, 0, 0), (5, W, 1, 0),          # same page, different uTLBs, W first in sort order; R is the duplicate
            (9, W, 0, 1),                        # already mapped RW -> authorised
            (40, R, 0, 2),                       # resident on CPU node 1 -> copy run from CXL
            (41, R, 0, 3),                       # resident on CPU node 0 -> copy run from DRAM
            (100, PF, 0, 4),                     # prefetch access, page resident nowhere -> first touch
            (3, W, 0, 5, 1), (3, R, 1, 5, 1), (7, PF, 0, 6, 1), (8, AT, 0, 7, 1)]   # block 1 is read-only
    addr = np.array([A(r[0], r[4] if len(r) > 4 else 0) for r in rows], np.uint64)
    pk = TG.encode_packets(np.full(len(rows), tbl["instance_ptr"][0]), addr, [r[1] for r in rows], [r[2] for r in rows],
                           [TG.T1_CLIENTS[r[3]] for r in rows], np.arange(len(rows)) + 100)
    gp = TG.gpu_params(256)
    st0 = S.HostBlockState(2)
    bit = lambda m, s, p: m.__setitem__((s, p // 64), m[s, p // 64] | np.uint64(1 << (p % 64)))
    for m in (st0.gpu_resident, st0.gpu_pte0, st0.gpu_pte1):
        bit(m, 0, 9)
    bit(st0.cpu_resident1, 0, 40); bit(st0.cpu_resident0, 0, 41)
    st0.logical_prot[1] = 1
    k1, want, got = _both(pk, gp, tbl, lay, st0, pf=(0, 51, 1))     # prefetch heuristic off: isolate stage C/E
    S.assert_service_equal(want, got, 256)
    o, fl = k1["ordered_rec"], want["flags"]
    page_of = lambda i: int((int(o["fault_address"][i]) - TG.VA_BASE) >> 12)
    byp = {page_of(i): [] for i in range(len(rows))}
    for i in range(len(rows)):
        byp[page_of(i)].append(int(fl[i]))
    assert byp[5] == [S_FL(serviced=1, acc=2), S_FL(dup=1)]
    assert byp[9] == [S_FL(auth=1)]
    assert byp[40] == [S_FL(serviced=1, acc=1)] and byp[41] == [S_FL(serviced=1, acc=1)]
    assert byp[100] == [S_FL(serviced=1, acc=0)]
    # read-only block 1, page 3: the READ was coalesced into the WRITE at fetch time (batch-wide rule), so ONE entry
    # with access_type_mask {READ, WRITE}: the write is fatal (cancel WRITE_AND_ATOMIC, reason permissions) but the
    # pending read instance is still serviced as a READ (check_fault_access_permissions :1326-1349)
    assert byp[512 + 3] == [S_FL(serviced=1, fatal=1, acc=1, reason=2, cancel=1)]
    assert byp[512 + 7] == [S_FL(serviced=1, acc=0)]
    assert byp[512 + 8] == [S_FL(fatal=1, reason=2, cancel=1)]
    assert int(want["batch"]["num_duplicate_faults"][0]) == 2
    rec = want["records"]
    assert len(rec) == 2 and list(rec["page_fault_count"]) == [4, 2]
    assert list(rec["migrated"][0]) == [1, 1] and rec["first_touch"][0] == 2 and rec["first_touch"][1] == 2
    st = want["state"]
    assert T.bool_from_masks(st.gpu_pte2)[0, [5, 40, 41, 100]].all()          # read-write VMA -> READ_WRITE_ATOMIC
    assert T.bool_from_masks(st.gpu_pte0)[1, [3, 7]].all() and not T.bool_from_masks(st.gpu_pte1)[1].any()
    assert not T.bool_from_masks(st.cpu_resident0)[0, 41] and not T.bool_from_masks(st.cpu_resident1)[0, 40]


def S_FL(serviced=0, dup=0, fatal=0, ip=0, auth=0, acc=0, reason=0, cancel=0):
    return serviced | dup << 1 | fatal << 2 | ip << 3 | auth << 4 | (acc << 5 if serviced else 0) | reason << 8 | cancel << 12


def test_addresses_outside_every_va_range():
    """service_fault_batch_dispatch -> service_fault_batch_fatal_notify: fatal (invalid address) or, for a
    prefetch access, invalid prefetch; nothing is serviced and no state changes."""
    pk, tbl, lay = TG.fault_trace(700, 8, "uniform", 2, total_va_bytes=SMALL_VA)
    out = np.arange(700) % 5 == 2
    pk[out, 3] |= 0x40                       # far above the managed range
    gp = TG.gpu_params(128)
    st0 = S.random_state(lay["blocks_per_space"] * 2, 6)
    k1, want, got = _both(pk, gp, tbl, lay, st0)
    S.assert_service_equal(want, got, 128)
    fl = want["flags"]
    assert ((fl >> 8) & 7 == 1).sum() > 50                     # UvmEventFatalReasonInvalidAddress
    assert int(want["batch"]["num_invalid_prefetch_faults"].sum()) > 0
    assert (want["records"]["slot"] == 0xFFFFFFFF).sum() > 50


def test_fatal_fault_types_are_skipped_and_failed_batches_not_serviced():
    pk, tbl, lay = TG.fault_trace(600, 5, "adversarial", 2, total_va_bytes=SMALL_VA)
    fatal = np.arange(600) % 7 == 3
    pk[fatal, 7] = (pk[fatal, 7] & ~np.uint32(0x1F)) | TG.HW_FAULT_TYPE_PDE_SIZE
    pk[fatal, 3] |= 0x100
    pk[400, 1] ^= 0x5                        # stale channel in the second batch -> whole batch refetched
    gp = TG.gpu_params(256)
    st0 = S.random_state(lay["blocks_per_space"] * 2, 2)
    k1, want, got = _both(pk, gp, tbl, lay, st0)
    assert list(k1["batch"]["status"]) == [0, F.NV_WARN_MORE_PROCESSING_REQUIRED, 0]
    S.assert_service_equal(want, got, 256)
    assert (want["records"]["batch"] != 1).all()


@pytest.mark.parametrize("thr", [1, 75, 100])
def test_prefetch_threshold_flows_into_service(thr):
    pk, tbl, lay = TG.fault_trace(800, 13, "stream", 1, total_va_bytes=SMALL_VA)
    gp = TG.gpu_params(256)
    st0 = S.random_state(lay["blocks_per_space"], 8, ro_frac=0, none_frac=0)
    _, want, got = _both(pk, gp, tbl, lay, st0, pf=(1, thr, 1))
    S.assert_service_equal(want, got, 256, "thr %d" % thr)


# ------------------------------------------------------------------------------------------------
# The CPU restatement against the reference's own entry loop (service_fault_batch_block_locked and its helpers,
# compiled verbatim into oracle/_ref/libref_service.so).
def _ref_vs_oracle(pk, gp, tbl, lay, st0, what, pf=(1, 51, 1)):
    ref = T.load_ref("service")
    if ref is None:
        pytest.skip("oracle/_ref/libref_service.so not built (needs /root/reference)")
    k1 = _k1(pk, gp, tbl)
    layout = S.make_layout(lay)
    a = S.run_service_ref(ref, k1, gp, tbl, layout, st0, pf)
    b = S.run_service_oracle(T.load_oracle(), k1, gp, tbl, layout, st0, pf)
    S.assert_ref_equals_oracle(a, b, what)
    return a, b


@pytest.mark.parametrize("dist", ["uniform", "stream", "hotset", "adversarial"])
@pytest.mark.parametrize("B", [256, 48])
def test_oracle_matches_the_reference_entry_loop(dist, B):
    pk, tbl, lay = TG.fault_trace(3000, 41, dist, 2, total_va_bytes=SMALL_VA)
    st0 = S.random_state(lay["blocks_per_space"] * 2, 14)
    a, _ = _ref_vs_oracle(pk, TG.gpu_params(B), tbl, lay, st0, "%s %d" % (dist, B))
    assert len(a["records"]) > 5


def test_oracle_matches_the_reference_on_unmapped_addresses_fatal_types_and_read_only_ranges():
    pk, tbl, lay = TG.fault_trace(2500, 8, "adversarial", 2, total_va_bytes=SMALL_VA)
    out = np.arange(2500) % 5 == 2
    pk[out, 3] |= 0x40                                          # far above the managed range: no va_range
    fatal = np.arange(2500) % 11 == 3
    pk[fatal, 7] = (pk[fatal, 7] & ~np.uint32(0x1F)) | TG.HW_FAULT_TYPE_PDE_SIZE
    pk[fatal, 3] |= 0x80                                        # the reference asserts that a va_block never mixes fatal
    st0 = S.random_state(lay["blocks_per_space"] * 2, 6, ro_frac=0.4, none_frac=0.2)   # fault types with serviceable ones (:1449)
    a, b = _ref_vs_oracle(pk, TG.gpu_params(128), tbl, lay, st0, "unmapped + fatal + read-only")
    assert int((a["flags"] & 4 != 0).sum()) > 100 and int((a["flags"] & 8 != 0).sum()) > 10
    assert int(a["batch"]["num_duplicate_faults"].sum()) > 100


def test_runs_that_need_no_service_are_evaluated_32_at_a_time():
    """Two va_blocks, almost fully resident and mapped, touched by ~150 batches: most runs need no service (the warp
    kernel evaluates them one per lane, speculatively), a few hit one of the unmapped pages and must be applied in
    order, in between."""
    rng = np.random.default_rng(5)
    n = 2400
    tbl = TG.instance_ptr_table(1)
    lay = dict(va_base=TG.VA_BASE, blocks_per_space=2, n_spaces=1)
    page = rng.integers(0, 512, n)
    blk = rng.integers(0, 2, n)
    addr = (TG.VA_BASE + (blk.astype(np.uint64) << np.uint64(21)) + (page.astype(np.uint64) << np.uint64(12))).astype(np.uint64)
    acc = rng.choice([TG.HW_VIRT_READ, TG.HW_VIRT_WRITE, TG.HW_VIRT_ATOMIC_STRONG, TG.HW_VIRT_PREFETCH], n)
    pk = TG.encode_packets(np.full(n, tbl["instance_ptr"][0]), addr, acc, rng.integers(0, 8, n),
                           [TG.T1_CLIENTS[i] for i in rng.integers(0, 8, n)], np.arange(n) + 1)
    gp = TG.gpu_params(16)
    st0 = S.HostBlockState(2)
    full = np.ones((2, 512), bool)
    holes = np.zeros((2, 512), bool)
    holes[0, [7, 130, 131, 400]] = True
    holes[1, 32:48] = True                                    # one 64 KB big page of block 1 still lives on the CXL node
    st0.gpu_resident = T.masks_from_bool(full & ~holes)
    st0.cpu_resident1 = T.masks_from_bool(holes)
    for m in ("gpu_pte0", "gpu_pte1", "gpu_pte2"):
        setattr(st0, m, T.masks_from_bool(full & ~holes))
    for pf in ((1, 51, 1), (0, 51, 1)):                      # with the prefetcher off every hole needs its own service
        _, want, got = _both(pk, gp, tbl, lay, st0, pf=pf, heavy_min=ALL_HEAVY)
        S.assert_service_equal(want, got, 16, "speculative runs, prefetch %d" % pf[0])
    rec = want["records"]
    assert len(rec) > 250 and 8 <= int((rec["page_fault_count"] > 0).sum()) <= 40


@pytest.mark.parametrize("heavy_min", [96, NONE_HEAVY, ALL_HEAVY])
def test_servicing_breaks_read_duplication(heavy_min):
    """A table with read-duplicated pages (uvm_va_block_make_resident_read_duplicate made them): a fault the GPU
    raises on such a page -- a write, since its write access was revoked -- goes through make_resident, which
    drops the CPU copy and the read_duplicated mark of every page it is handed (uvm_va_block.c:4538-4579)."""
    import residentlib as R
    pk, tbl, lay = TG.fault_trace(1500, 41, "hotset", 2, total_va_bytes=SMALL_VA)
    gp = TG.gpu_params(128)
    st0 = R.with_read_duplication(S.random_state(lay["blocks_per_space"] * 2, 9), 9, dup=0.3)
    k1, want, got = _both(pk, gp, tbl, lay, st0, heavy_min=heavy_min)
    S.assert_service_equal(want, got, 128, "read duplication, heavy_min %d" % heavy_min)
    before, after = T.bool_from_masks(st0.read_duplicated), T.bool_from_masks(got["state"].read_duplicated)
    assert not (after & ~before).any() and int((before & ~after).sum()) > 200
    cpu = T.bool_from_masks(got["state"].cpu_resident0) | T.bool_from_masks(got["state"].cpu_resident1)
    assert not (after & ~cpu).any() and not (T.bool_from_masks(got["state"].gpu_resident) & cpu & ~after).any()
    ref = T.load_ref("service")
    if ref is not None and heavy_min == 96:
        layout = S.make_layout(lay)
        S.assert_ref_equals_oracle(S.run_service_ref(ref, k1, gp, tbl, layout, st0), want, "read duplication")


@pytest.mark.parametrize("dist", ["uniform", "hotset", "adversarial"])
def test_reference_code_end_to_end_equals_the_restatement(dist):
    """The CPU arm with NO restated step on the data path except the mapping of the serviced pages: the reference's
    entry loop (libref_service.so) calling the reference's prefetch file (libref_prefetch.so) and the reference's
    uvm_va_block_make_resident (libref_resident.so).  Same records, copy runs, counters and state as the restatement."""
    rs, rp, rr = T.load_ref("service"), T.load_ref("prefetch"), T.load_ref("resident")
    if rs is None or rp is None or rr is None:
        pytest.skip("reference libraries not built (no /root/reference)")
    pk, tbl, lay = TG.fault_trace(2500, 91, dist, 2, total_va_bytes=SMALL_VA)
    gp = TG.gpu_params(128)
    st0 = S.random_state(lay["blocks_per_space"] * 2, 12)
    k1 = _k1(pk, gp, tbl)
    layout = S.make_layout(lay)
    want = S.run_service_oracle(T.load_oracle(), k1, gp, tbl, layout, st0)
    assert S.use_reference_back_half(rs, rp, rr)
    try:
        got = S.run_service_ref(rs, k1, gp, tbl, layout, st0)
    finally:
        S.use_reference_back_half(rs, None, None)
    S.assert_ref_equals_oracle(got, want, "reference end to end, %s" % dist)
    assert int((want["records"]["page_fault_count"] > 0).sum()) >= 4


@pytest.mark.parametrize("B", [2048, 4096])
def test_service_pass_on_batches_above_1024_entries(B):
    """The servicing pass is indifferent to the batch size (16-bit fields hold up to 65535 entries per batch)."""
    pk, tbl, lay = TG.fault_trace(6000, 71, "uniform", 2, total_va_bytes=SMALL_VA)
    gp = TG.gpu_params(B)
    st0 = S.random_state(lay["blocks_per_space"] * 2, 9)
    _, want, got = _both(pk, gp, tbl, lay, st0)
    S.assert_service_equal(want, got, B, "B %d" % B)
    S.check_service_invariants(got, st0)
