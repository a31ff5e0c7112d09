"""CPU tests of the thrashing detector (section 1f of include/uvmtier.h): the restatement against the reference's own
event callback compiled verbatim (oracle/_ref/libref_thrashing.so), on directed sequences and on random event
streams; per-va_block processing order (what the device entry point uses) gives the time-order result."""
import numpy as np
import pytest

import thrashlib as TH
import uvmtier_testlib as T


def events(rows):
    ev = np.zeros(len(rows), TH.EVENT_DT)
    for i, r in enumerate(rows):
        ev[i] = tuple(r) + (0,) * (8 - len(r))
    return ev


def test_ping_pong_on_one_page_is_detected_at_the_threshold():
    orc = T.load_oracle()
    # (time, slot, first, n, processor, kind, flags): two processors take page 7 of va_block 2 from each other every 100 us
    ev = events([(1000 + 100000 * i, 2, 7, 1, i & 1, 0, 0) for i in range(6)])
    st = TH.run_oracle(orc, ev, 4)
    # event 0 only stamps the block; event 1 opens the tracking (other processor inside the window) and stamps the page;
    # events 2, 3, 4 are thrashing events 1, 2, 3 -> detected by event 4; event 5 is the fourth
    assert st.blocks["tracked"].tolist() == [0, 0, 1, 0] and st.blocks["num_thrashing_pages"][2] == 1
    assert st.thrashing_pages[2 * 16] == 1 << 7 and st.counters[0] == 1
    w = int(st.page_word[2 * 512 + 7])
    assert (w >> 60) & 7 == 4 and (w >> 58) & 1 == 1 and (w & ((1 << 58) - 1)) == (1000 + 500000) >> 6
    assert st.page_procs[2 * 512 + 7] == 3 and st.blocks["last_thrashing_time_stamp"][2] == 1000 + 500000
    # one processor alone never opens the tracking, however fast it faults
    ev = events([(1000 + 100 * i, 1, 0, 512, 1, 0, 0) for i in range(50)])
    st = TH.run_oracle(orc, ev, 2)
    assert st.blocks["tracked"].sum() == 0 and st.counters[0] == 0 and not st.page_word.any()


def test_window_expiry_resets_a_thrashing_page_and_user_migrations_clear_a_region():
    orc = T.load_oracle()
    rows = [(1000 + 1000 * i, 0, 0, 4, i & 1, i & 1, 0) for i in range(8)]      # migrations and revocations alternate
    rows.append((1000 + 8000 + 2 * TH.LAPSE_NS, 0, 0, 2, 0, 0, 0))               # after the window: pages 0, 1 reset
    rows.append((1000 + 9000 + 2 * TH.LAPSE_NS, 0, 3, 1, 0, 0, TH.EV_RESET))     # cudaMemPrefetchAsync-like: page 3 cleared
    st = TH.run_oracle(orc, events(rows), 1)
    assert st.thrashing_pages[0] == 0b0100 and st.blocks["num_thrashing_pages"][0] == 1 and st.counters.tolist()[:2] == [4, 3]
    assert st.page_word[0] == 0 and st.page_word[1] == 0 and st.page_word[3] == 0 and st.page_procs[3] == 0
    assert (int(st.page_word[2]) >> 58) & 3 == 3 and (int(st.page_word[2]) >> 60) & 7 == 6
    # the reset event does not stamp the block (it returns before the epilogue)
    assert st.blocks["last_time_stamp"][0] == 1000 + 8000 + 2 * TH.LAPSE_NS


@pytest.mark.parametrize("seed,n_slots,n_events,threshold,lapse", [(1, 64, 20000, 3, TH.LAPSE_NS), (42, 8, 15000, 1, TH.LAPSE_NS),
                                                                   (7, 200, 30000, 7, 100000), (0xC0FFEE, 32, 20000, 5, 4096)])
def test_restatement_matches_reference_event_callback(seed, n_slots, n_events, threshold, lapse):
    orc, ref = T.load_oracle(), T.load_ref("thrashing")
    ev = TH.ping_pong_events(n_slots, n_events, seed, lapse_ns=lapse)
    want = TH.run_ref(ref, ev, n_slots, lapse, threshold)
    got = TH.run_oracle(orc, ev, n_slots, lapse, threshold)
    TH.assert_state_equal(want, got, "seed %d" % seed)
    assert want.counters[0] > 20 and want.counters[1] > 5 and want.blocks["tracked"].sum() > 2
    # grouped by va_block (the device entry point's order): the va_blocks are independent, so nothing changes
    again = TH.run_oracle(orc, TH.by_slot(ev), n_slots, lapse, threshold)
    TH.assert_state_equal(want, again, "by slot, seed %d" % seed)


def test_directed_sequences_match_reference():
    orc, ref = T.load_oracle(), T.load_ref("thrashing")
    rows = [(1000 + 100000 * i, 2, 7, 1, i & 1, 0, 0) for i in range(6)]
    rows += [(10 ** 7 + 1000 * i, 0, 0, 4, i & 1, i & 1, 0) for i in range(8)]
    rows += [(10 ** 7 + 8000 + 2 * TH.LAPSE_NS, 0, 0, 2, 0, 0, 0), (10 ** 7 + 9000 + 2 * TH.LAPSE_NS, 0, 3, 1, 0, 0, TH.EV_RESET)]
    rows += [(10 ** 8 + 63 * i, 1, 500, 12, i % 3, 1, 0) for i in range(20)]     # time stamps closer than the 64 ns unit
    ev = events(rows)
    TH.assert_state_equal(TH.run_ref(ref, ev, 4), TH.run_oracle(orc, ev, 4), "directed")


# ---------------- the hint half: throttle, pin, un-pin, epoch reset ----------------
def _same_hints(a, b, what):
    assert len(a) == len(b)
    for i, (x, y) in enumerate(zip(a, b)):
        for f in ("type", "pin_residency", "pin_processors", "throttle_end"):
            assert np.array_equal(x[f], y[f]), "%s: hint batch %d field %s: want %s got %s" % (what, i, f, x[f], y[f])


@pytest.mark.parametrize("seed,topology,procs,kw", [
    (1, "pcie", (0, 1), {}), (2, "pcie", (0, 1, 9), {}), (3, "c2c", (0, 1), {}), (4, "nvlink", (1, 9), {}),
    (5, "nvlink", (0, 1, 9), dict(pref_frac=0.5)), (6, "pcie", (0, 1), dict(pin_threshold=2, pin=1)),
    (7, "c2c", (0, 1, 2), dict(threshold=1, epoch=30, max_resets=2)), (8, "pcie", (0, 1, 9), dict(pref_frac=1.0, pin_threshold=1, pin=0))])
def test_hint_restatement_matches_reference_on_driver_histories(seed, topology, procs, kw):
    """Events, hint queries and pin expiries interleaved the way a driver produces them, through the reference's own
    uvm_perf_thrashing_get_hint / thrashing_event_cb / thrashing_unpin_pages (compiled verbatim) and through the
    restatement: every hint and every field of the tracking state."""
    orc, ref = T.load_oracle(), T.load_ref("thrashing_hint")
    if ref is None:
        pytest.skip("oracle/_ref/libref_thrashing_hint.so not built (needs /root/reference)")
    pref_frac = kw.pop("pref_frac", 0.0)
    cfg = TH.make_cfg(topology, gpus=tuple(p for p in (1, 2, 9) if p in procs) or (1,), **kw)
    n_slots = 6
    steps, sim = TH.driver_history(n_slots, 6000, seed, cfg, procs=procs, pref_frac=pref_frac)
    want, want_h = TH.run_steps_ref(ref, steps, n_slots, cfg)
    got, got_h = TH.run_steps_oracle(orc, steps, n_slots, cfg)
    _same_hints(want_h, got_h, "seed %d" % seed)
    TH.assert_hint_state_equal(want, got, "seed %d" % seed)
    types = np.concatenate([h["type"] for h in want_h])
    assert (types == TH.HINT_THROTTLE).sum() > 5 and (types == TH.HINT_PIN).sum() > 5 and (types == TH.HINT_NONE).sum() > 5


@pytest.mark.parametrize("seed,topology", [(11, "pcie"), (12, "c2c"), (13, "nvlink")])
def test_hint_restatement_matches_reference_on_random_streams(seed, topology):
    """Not a history a driver could produce (queries with arbitrary residency and preferred location, events on pinned
    pages): every branch of the hint gets exercised, the two implementations must still agree."""
    orc, ref = T.load_oracle(), T.load_ref("thrashing_hint")
    if ref is None:
        pytest.skip("oracle/_ref/libref_thrashing_hint.so not built (needs /root/reference)")
    rng = np.random.default_rng(seed)
    procs = np.array([0, 1, 2, 9])
    cfg = TH.make_cfg(topology, gpus=(1, 2, 9), pin_threshold=3, epoch=50, pin=20)
    n_slots, now, steps = 4, 5000, []
    for _ in range(4000):
        now += int(rng.choice([100, 2000, TH.LAPSE_NS // 3, TH.LAPSE_NS * 2, TH.LAPSE_NS * 60], p=[0.4, 0.3, 0.2, 0.08, 0.02]))
        r = rng.random()
        if r < 0.5:
            n = int(rng.integers(1, 5))
            ev = np.zeros(n, TH.EVENT_DT)
            for j in range(n):
                kind = int(rng.random() < 0.2)
                flags = 0 if kind else int(rng.choice([0, 0, 0, TH.EV_RESET, TH.EV_PREFETCH]))
                ev[j] = (now + j, int(rng.integers(0, n_slots)), int(rng.integers(0, 8)), int(rng.integers(1, 4)), int(rng.choice(procs)), kind, flags, 0)
            steps.append(("events", ev))
        elif r < 0.97:
            n = int(rng.integers(1, 6))
            q = np.zeros(n, TH.QUERY_DT)
            for j in range(n):
                res = int(rng.integers(1, 16))
                resident = sum(1 << int(procs[b]) for b in range(4) if (res >> b) & 1)
                pref = int(rng.choice([TH.NO_PREF, TH.NO_PREF, 0, 1, 2, 9]))
                q[j] = (now + j, int(rng.integers(0, n_slots)), int(rng.integers(0, 10)), int(rng.choice(procs)), 0, resident, pref, 0)
            steps.append(("hints", q))
        else:
            steps.append(("expire", now))
    want, want_h = TH.run_steps_ref(ref, steps, n_slots, cfg)
    got, got_h = TH.run_steps_oracle(orc, steps, n_slots, cfg)
    _same_hints(want_h, got_h, "seed %d" % seed)
    TH.assert_hint_state_equal(want, got, "seed %d" % seed)
    types = np.concatenate([h["type"] for h in want_h])
    assert (types == TH.HINT_THROTTLE).sum() > 10 and (types == TH.HINT_PIN).sum() > 20


@pytest.mark.parametrize("seed,topology,procs", [(21, "pcie", (0, 1)), (22, "c2c", (0, 1, 2)), (23, "nvlink", (0, 1, 9))])
def test_simt_hint_kernels_match_restatement_on_driver_histories(seed, topology, procs):
    """The product's kernels (k_thrash_events with the hint tables, k_thrash_hints, k_thrash_unpin_expired) under the
    emulator against the restatement, on interleaved events / one-by-one hint queries / expiries."""
    orc, sim = T.load_oracle(), T.load_simt()
    cfg = TH.make_cfg(topology, gpus=tuple(p for p in (1, 2, 9) if p in procs), pin_threshold=3, pin=20, epoch=200)
    n_slots = 5
    steps, _ = TH.driver_history(n_slots, 400, seed, cfg, procs=procs, hot_pages=2)
    want, want_h = TH.run_steps_oracle(orc, steps, n_slots, cfg)
    got, got_h = TH.run_steps_simt(sim, steps, n_slots, cfg)
    _same_hints(want_h, got_h, "seed %d" % seed)
    TH.assert_hint_state_equal(want, got, "seed %d" % seed)
    types = np.concatenate([h["type"] for h in want_h])
    assert (types == TH.HINT_THROTTLE).sum() > 3 and (types == TH.HINT_PIN).sum() > 0


@pytest.mark.parametrize("seed,topology", [(31, "pcie"), (32, "nvlink")])
def test_simt_hint_kernels_match_restatement_on_random_streams(seed, topology):
    orc, sim = T.load_oracle(), T.load_simt()
    cfg = TH.make_cfg(topology, gpus=(1, 2, 9), pin_threshold=3, epoch=50, pin=20)
    n_slots = 4
    steps = TH.random_stream(seed, n_slots, 200)
    want, want_h = TH.run_steps_oracle(orc, steps, n_slots, cfg)
    got, got_h = TH.run_steps_simt(sim, steps, n_slots, cfg)
    _same_hints(want_h, got_h, "seed %d" % seed)
    TH.assert_hint_state_equal(want, got, "seed %d" % seed)
