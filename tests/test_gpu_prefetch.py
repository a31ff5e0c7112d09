"""GPU parity tests of the prefetch leg: the CUDA path through the C ABI vs the oracle, the golden
vectors produced by the reference's own C, and size-independent properties at BASELINE sizes."""
import numpy as np
import pytest

import uvmtier_testlib as T
from test_oracle_prefetch import _golden_ats_cases, _golden_hint_cases

pytestmark = pytest.mark.gpu


def _ctx_with(thr):
    import uvmtier
    return uvmtier.Context(0, uvm_perf_prefetch_threshold=thr)


def test_gpu_matches_golden_vectors():
    for thr, inp, want in _golden_hint_cases():
        with _ctx_with(thr) as c:
            got = c.perf_prefetch_get_hint_va_block(inp["blocks"], inp["resident"], inp["faulted"], inp["thrashing"],
                                                    inp["cpu_read"], inp["info"])
        for g_, w_ in zip(got, want):
            assert np.array_equal(g_, w_)
    for thr, with_res, inp, want in _golden_ats_cases():
        with _ctx_with(thr) as c:
            got = c.perf_prefetch_compute_ats(inp["faulted"], inp["fregion"], inp["mregion"], inp["residency"] if with_res else None)
        assert np.array_equal(got, want)


@pytest.mark.parametrize("cfg", [dict(), dict(full_blocks=False, aligned=False),
                                 dict(full_blocks=False, aligned=False, thrashing=True),
                                 dict(dst="cpu", full_blocks=False, aligned=False, thrashing=True)])
@pytest.mark.parametrize("thr", [51, 1, 100, 30])
def test_gpu_matches_oracle_random(cfg, thr):
    orc = T.load_oracle()
    inp = T.gen_prefetch_blocks(4099, 21, **cfg)       # odd count: exercises the ragged tail
    want = T.run_prefetch_get_hint(orc, "oracle", inp, (1, thr, 1))
    with _ctx_with(thr) as c:
        got = c.perf_prefetch_get_hint_va_block(inp["blocks"], inp["resident"], inp["faulted"], inp["thrashing"],
                                                inp["cpu_read"], inp["info"])
    for g_, w_ in zip(got, want):
        assert np.array_equal(g_, w_)


def test_gpu_compute_ats_matches_oracle(ctx):
    orc = T.load_oracle()
    inp = T.gen_compute_ats(3001, 5)
    for with_res in (True, False):
        want = T.run_compute_ats(orc, "oracle", inp, with_res)
        got = ctx.perf_prefetch_compute_ats(inp["faulted"], inp["fregion"], inp["mregion"], inp["residency"] if with_res else None)
        assert np.array_equal(got, want)


def test_gpu_baseline_size_32768_blocks(ctx):
    """Config B: 64 GB VA = 32768 va_blocks.  Full compare with the oracle (it takes ~1 s) plus
    properties: the hint never contains a faulted page, stays inside the block, is idempotent."""
    orc = T.load_oracle()
    inp = T.gen_prefetch_blocks(32768, 99)
    want = T.run_prefetch_get_hint(orc, "oracle", inp)
    got = ctx.perf_prefetch_get_hint_va_block(inp["blocks"], inp["resident"], inp["faulted"], None, None, inp["info"])
    for g_, w_ in zip(got, want):
        assert np.array_equal(g_, w_)
    assert not (got[0] & inp["faulted"]).any()
    again = ctx.perf_prefetch_get_hint_va_block(inp["blocks"], inp["resident"], inp["faulted"], None, None, inp["info"])
    assert np.array_equal(again[0], got[0])


def test_gpu_empty_and_invalid_args(ctx):
    import ctypes
    import uvmtier
    z = np.zeros((0, 8), "<u8")
    out = ctx.perf_prefetch_compute_ats(z, np.zeros((0, 2), "<u2"), np.zeros((0, 2), "<u2"), None)
    assert out.shape == (0, 8)
    st = uvmtier.lib().uvmtier_perf_prefetch_compute_ats(ctx.handle, ctypes.c_size_t(4), None, None, None, None, None)
    assert st == 0x1F   # NV_ERR_INVALID_ARGUMENT
