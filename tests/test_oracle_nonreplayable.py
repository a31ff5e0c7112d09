"""Non-replayable fault servicing on the CPU: the restatement (oracle/oracle_nonreplayable.c) against the reference's own
service_managed_fault_in_block_locked compiled verbatim (oracle/_ref/libref_nonreplayable.so), and the product's kernels
under the SIMT emulator against the restatement."""
import numpy as np
import pytest

import nonreplaylib as N
import servicelib as S
import uvmtier_testlib as T
from uvmtier import tracegen as TG


def _problem(n, seed, blocks=24, n_spaces=2, B=64):
    tbl = TG.instance_ptr_table(n_spaces)
    lay = dict(va_base=TG.VA_BASE, blocks_per_space=blocks, n_spaces=n_spaces, gpu_id=1)
    gp = TG.gpu_params(B)
    st0 = S.random_state(blocks * n_spaces, seed, populated=0.6)
    pk = N.shadow_buffer(n, seed, lay, tbl)
    return pk, tbl, lay, gp, st0


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_restatement_matches_the_reference_function(seed):
    ref = T.load_ref("nonreplayable")
    if ref is None:
        pytest.skip("reference libraries not built (no /root/reference)")
    pk, tbl, lay, gp, st0 = _problem(3000, seed)
    layout = S.make_layout(lay)
    want = N.run_checker(ref, "ref", pk, gp, tbl, layout, st0)
    got = N.run_checker(T.load_oracle(), "oracle", pk, gp, tbl, layout, st0)
    N.assert_equal(want, got, "reference vs restatement, seed %d" % seed)
    fl = want["flags"]
    assert int((fl & 1).sum()) > 1500 and int(((fl & 4) != 0).sum()) > 100 and int((fl == 0).sum()) > 30
    reasons = set(((fl[(fl & 4) != 0] >> 8) & 7).tolist())
    assert reasons == {1, 2, 3}


def test_reference_end_to_end_with_the_reference_back_half():
    rn, rp, rr = T.load_ref("nonreplayable"), T.load_ref("prefetch"), T.load_ref("resident")
    if rn is None or rp is None or rr is None:
        pytest.skip("reference libraries not built (no /root/reference)")
    pk, tbl, lay, gp, st0 = _problem(2000, 9)
    layout = S.make_layout(lay)
    want = N.run_checker(T.load_oracle(), "oracle", pk, gp, tbl, layout, st0)
    assert S.use_reference_back_half(rn, rp, rr)
    try:
        got = N.run_checker(rn, "ref", pk, gp, tbl, layout, st0)
    finally:
        S.use_reference_back_half(rn, None, None)
    N.assert_equal(want, got, "reference front and back")


@pytest.mark.parametrize("seed,heavy_min,B", [(4, 96, 64), (5, 0, 32), (6, 0xFFFFFFFF, 256)])
def test_simt_kernels_match_the_restatement(seed, heavy_min, B):
    """k_nonreplayable_prepare + the servicing pass in non-replayable mode (no already-authorised shortcut, no
    duplicates, no cancel mode), light va_blocks on lane groups / heavy ones on a warp."""
    pk, tbl, lay, gp, st0 = _problem(1500, seed, blocks=8, B=B)
    layout = S.make_layout(lay)
    want = N.run_checker(T.load_oracle(), "oracle", pk, gp, tbl, layout, st0)
    got = N.run_simt(T.load_simt(), pk, gp, tbl, layout, st0, heavy_min=heavy_min)
    N.assert_equal(want, got, "simt seed %d heavy_min %d" % (seed, heavy_min))
