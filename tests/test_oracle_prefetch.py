"""CPU tests of the prefetch leg: reference KAT, oracle vs reference-compiled vs golden vectors,
and the product's kernel source under the SIMT emulator."""
import ctypes
import os

import numpy as np
import pytest

import uvmtier_testlib as T

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "prefetch_golden.npz")


def _kat(lib, prefix):
    # UVM/uvm_perf_utils_test.c:682-715: leaf_count 9 (level_count 5), bits {1,2,4,7,8}, walk from page 6
    pages = np.zeros(8, "<u8")
    for b in (1, 2, 4, 7, 8):
        pages[0] |= np.uint64(1 << b)
    c = np.zeros(16, "<u2"); f = np.zeros(16, "<u2"); o = np.zeros(16, "<u2")
    n = getattr(lib, prefix + "_prefetch_tree_traverse")(T.P(pages), 9, 5, 0, 6, T.P(c), T.P(f), T.P(o))
    assert n == 5
    assert list(c[:5]) == [0, 1, 2, 4, 5]          # level 4 (leaf) .. level 0 (root)
    assert list(f[:5]) == [6, 6, 4, 0, 0] and list(o[:5]) == [7, 8, 8, 8, 9]


def test_reference_kat_bitmap_tree_traversal_oracle():
    _kat(T.load_oracle(), "oracle")


def test_reference_kat_bitmap_tree_traversal_reference_build():
    ref = T.load_ref("prefetch")
    if ref is None:
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    _kat(ref, "ref")


def test_tunable_validation_matches_reference_rules():
    # UVM/uvm_perf_prefetch.c:504-535
    orc = T.load_oracle()
    orc.oracle_prefetch_init(1, 101, 0)
    assert orc.oracle_prefetch_threshold() == 51 and orc.oracle_prefetch_min_faults() == 1
    orc.oracle_prefetch_init(1, 100, 20)
    assert orc.oracle_prefetch_threshold() == 100 and orc.oracle_prefetch_min_faults() == 20
    orc.oracle_prefetch_init(1, 51, 21)
    assert orc.oracle_prefetch_min_faults() == 1
    ref = T.load_ref("prefetch")
    if ref is not None:
        for args in ((1, 101, 0), (1, 100, 20), (1, 7, 21)):
            ref.ref_prefetch_init(*args); orc.oracle_prefetch_init(*args)
            assert ref.ref_prefetch_threshold() == orc.oracle_prefetch_threshold()
            assert ref.ref_prefetch_min_faults() == orc.oracle_prefetch_min_faults()
    orc.oracle_prefetch_init(1, 51, 1)


def _golden_hint_cases():
    g = np.load(GOLDEN)
    for k in range(int(g["n_hint"])):
        pre = "hint%d_" % k
        inp = dict(blocks=g[pre + "blocks"], resident=g[pre + "resident"], faulted=g[pre + "faulted"], info=g[pre + "info"],
                   thrashing=g[pre + "thrashing"] if len(g[pre + "thrashing"]) else None,
                   cpu_read=g[pre + "cpu_read"] if len(g[pre + "cpu_read"]) else None)
        yield int(g[pre + "threshold"]), inp, (g[pre + "out_mask"], g[pre + "out_residency"], g[pre + "out_info"])


def _golden_ats_cases():
    g = np.load(GOLDEN)
    for k in range(int(g["n_ats"])):
        pre = "ats%d_" % k
        inp = dict(faulted=g[pre + "faulted"], residency=g[pre + "residency"], fregion=g[pre + "fregion"], mregion=g[pre + "mregion"])
        yield int(g[pre + "threshold"]), bool(g[pre + "with_residency"]), inp, g[pre + "out_mask"]


def test_oracle_matches_golden_vectors():
    orc = T.load_oracle()
    for thr, inp, want in _golden_hint_cases():
        got = T.run_prefetch_get_hint(orc, "oracle", inp, (1, thr, 1))
        for g_, w_ in zip(got, want):
            assert np.array_equal(g_, w_)
    for thr, with_res, inp, want in _golden_ats_cases():
        assert np.array_equal(T.run_compute_ats(orc, "oracle", inp, with_res, (1, thr, 1)), want)


def test_simt_kernel_matches_golden_vectors():
    sm = T.load_simt()
    for thr, inp, want in _golden_hint_cases():
        got = T.run_prefetch_get_hint_simt(sm, inp, (1, thr, 1))
        for g_, w_ in zip(got, want):
            assert np.array_equal(g_, w_)
    for thr, with_res, inp, want in _golden_ats_cases():
        assert np.array_equal(T.run_compute_ats_simt(sm, inp, with_res, (1, thr, 1)), want)


@pytest.mark.parametrize("cfg", [dict(), dict(full_blocks=False, aligned=False),
                                 dict(full_blocks=False, aligned=False, thrashing=True),
                                 dict(dst="cpu", full_blocks=False, aligned=False, thrashing=True)])
@pytest.mark.parametrize("thr", [51, 1, 100, 75, 0])
def test_oracle_reference_simt_agree_random(cfg, thr):
    orc, ref, sm = T.load_oracle(), T.load_ref("prefetch"), T.load_simt()
    for seed in (11, 12):
        inp = T.gen_prefetch_blocks(128, seed, **cfg)
        o = T.run_prefetch_get_hint(orc, "oracle", inp, (1, thr, 1))
        s = T.run_prefetch_get_hint_simt(sm, inp, (1, thr, 1))
        for a, b in zip(o, s):
            assert np.array_equal(a, b)
        if ref is not None:
            r = T.run_prefetch_get_hint(ref, "ref", inp, (1, thr, 1))
            for a, b in zip(o, r):
                assert np.array_equal(a, b)


def test_prefetch_disabled_returns_empty():
    orc, sm = T.load_oracle(), T.load_simt()
    inp = T.gen_prefetch_blocks(32, 5)
    for got in (T.run_prefetch_get_hint(orc, "oracle", inp, (0, 51, 1)), T.run_prefetch_get_hint_simt(sm, inp, (0, 51, 1))):
        assert not got[0].any() and (got[1] == T.ID_INVALID).all()
        assert np.array_equal(got[2], inp["info"])
    orc.oracle_prefetch_init(1, 51, 1)


def test_prefetch_edge_cases():
    """empty fault mask, single-page block, odd tail (n not a multiple of the 2 blocks a warp takes)"""
    orc, sm = T.load_oracle(), T.load_simt()
    inp = T.gen_prefetch_blocks(5, 9, full_blocks=False, aligned=False)
    inp["blocks"]["num_pages"][0] = 1
    inp["blocks"]["faulted_first"][0] = 0; inp["blocks"]["faulted_outer"][0] = 1
    inp["faulted"][0] = 0; inp["faulted"][0, 0] = 1
    inp["resident"][0] = 0
    inp["faulted"][1] = 0                       # no faulted page at all
    inp["blocks"]["faulted_first"][1] = 0; inp["blocks"]["faulted_outer"][1] = 0
    o = T.run_prefetch_get_hint(orc, "oracle", inp)
    s = T.run_prefetch_get_hint_simt(sm, inp)
    for a, b in zip(o, s):
        assert np.array_equal(a, b)
    assert T.bool_from_masks(o[0])[1].sum() == 0
