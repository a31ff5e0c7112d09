"""CPU tests of the servicing pass with thrashing hints inside the fault loop: the product's kernels under the emulator
against the restatement (oracle_service.c asking oracle_thrashing.c for every hint, where the reference's loop asks
uvm_perf_thrashing_get_hint, uvm_gpu_replayable_faults.c:1524-1545)."""
import numpy as np
import pytest

import faultlib as F
import servicelib as S
import thrashsvclib as TS
import uvmtier_testlib as T


@pytest.mark.parametrize("seed,topology,kw", [(1, "pcie", {}), (2, "c2c", {}), (3, "pcie", dict(pin_threshold=1, pin=2)),
                                              (4, "c2c", dict(threshold=1, epoch=3, max_resets=1))])
def test_simt_service_with_hints_matches_restatement(seed, topology, kw):
    orc, sim = T.load_oracle(), T.load_simt()
    sc = TS.scenario(seed, topology, n_faults=1500, **kw)
    k1 = F.run_fetch_preprocess_simt(sim, sc["pk"], sc["gp"], sc["tbl"])
    want = TS.run_oracle(orc, sc, k1)
    got = TS.run_simt(sim, sc, k1)
    throttled, pinned, remote = TS.check(want, got, sc["B"], "seed %d %s" % (seed, topology))
    assert np.array_equal(want["thrash"].counters[3:6], got["thrash"].counters[3:6])
    assert throttled + pinned + remote > 20, (throttled, pinned, remote)
    if topology == "c2c":
        assert remote == 0                                   # everybody reaches everybody: pins land where the page is asked for
    # va_blocks the detector does not track are serviced exactly as without hints
    plain = S.run_service_oracle(orc, k1, sc["gp"], sc["tbl"], sc["layout"], sc["st0"])
    untracked = sc["th"].blocks["tracked"] == 0
    rec_w, rec_p = want["records"], plain["records"]
    sel = untracked[rec_w["slot"].clip(0, sc["n_slots"] - 1)] & (rec_w["slot"] < sc["n_slots"])
    for f in S.RECORD_FIELDS:
        assert np.array_equal(rec_w[f][sel], rec_p[f][sel]), f


def test_hints_off_is_the_plain_pass():
    """With no tracked va_block the thrashing entry point's result is the plain pass's."""
    orc, sim = T.load_oracle(), T.load_simt()
    sc = TS.scenario(5, "pcie", n_faults=800)
    import thrashlib as TH
    sc["th"] = TH.HintState(sc["n_slots"])
    k1 = F.run_fetch_preprocess_simt(sim, sc["pk"], sc["gp"], sc["tbl"])
    got = TS.run_simt(sim, sc, k1)
    plain = S.run_service_oracle(orc, k1, sc["gp"], sc["tbl"], sc["layout"], sc["st0"])
    S.assert_service_equal(plain, got, sc["B"], "no history")
    assert not (got["flags"] & 0xE000).any()


@pytest.mark.parametrize("seed,kw", [(11, {}), (12, dict(pin_threshold=1, pin=2)), (13, dict(threshold=1, epoch=3, max_resets=1))])
def test_restated_loop_with_hints_matches_the_reference_loop(seed, kw):
    """The reference's service_fault_batch_block_locked compiled verbatim (libref_service.so), its hint call answered by the
    thrashing restatement, against the restated loop: throttled entries (mark_fault_throttled), the pages of
    thrashing_pin_mask, batch counters, records, copy runs, block state and the hint tables.  Topology "c2c": every pin
    lands on the faulting GPU (a pin elsewhere is where this library's pass and the reference part ways, see uvmtier.h)."""
    orc, ref, sim = T.load_oracle(), T.load_ref("service"), T.load_simt()
    if ref is None:
        pytest.skip("oracle/_ref/libref_service.so not built (needs /root/reference)")
    sc = TS.scenario(seed, "c2c", n_faults=2000, **kw)
    k1 = F.run_fetch_preprocess_simt(sim, sc["pk"], sc["gp"], sc["tbl"])
    want = TS.run_ref(ref, sc, k1)
    got = TS.run_oracle(orc, sc, k1)
    S.assert_ref_equals_oracle(want, got, "seed %d" % seed)
    import thrashlib as TH
    TH.assert_hint_state_equal(want["thrash"], got["thrash"], "seed %d" % seed)
    for bit, name in ((0x2000, "throttled"), (0x4000, "pinned"), (0x8000, "pinned elsewhere")):
        assert np.array_equal(want["flags"] & bit, got["flags"] & bit), name
    assert int((want["flags"] & 0x2000 != 0).sum()) + int((want["flags"] & 0x4000 != 0).sum()) > 20
    assert not (want["flags"] & 0x8000).any()
