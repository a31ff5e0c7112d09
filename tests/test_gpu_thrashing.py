"""GPU tests of the thrashing detector through the C ABI vs the CPU restatement: every table bit-exact, over several
calls (history carries over), plus what the entry point does with events it cannot place."""
import numpy as np
import pytest

import servicelib as S
import thrashlib as TH
import uvmtier
import uvmtier_testlib as T
from uvmtier import tracegen as TG

pytestmark = pytest.mark.gpu


def gpu_state(ctx, t, counters):
    tb = ctx.thrashing_tables(t)
    st = TH.State(t.n_slots)
    st.blocks = tb["blocks"].view(TH.BLOCK_DT)
    st.page_word, st.page_procs, st.thrashing_pages = tb["page_word"], tb["page_procs"], tb["thrashing_pages"]
    st.counters = counters
    return st


@pytest.mark.parametrize("seed,n_slots,n_events,threshold,lapse,calls", [(1, 64, 20000, 3, TH.LAPSE_NS, 1), (42, 8, 15000, 1, TH.LAPSE_NS, 3),
                                                                         (7, 2048, 60000, 7, 100000, 2), (9, 32, 20000, 5, 4096, 4)])
def test_gpu_thrashing_tables_match_oracle(seed, n_slots, n_events, threshold, lapse, calls):
    orc = T.load_oracle()
    ev = TH.ping_pong_events(n_slots, n_events, seed, lapse_ns=lapse)
    layout = S.make_layout(dict(va_base=TG.VA_BASE, blocks_per_space=n_slots, n_spaces=1))
    want = TH.State(n_slots)
    total = np.zeros(4, "<u4")
    with uvmtier.Context(0) as ctx:
        t = ctx.thrashing_create(layout, lapse, threshold)
        for part in np.array_split(ev, calls):                    # time slices: each call continues the history
            part = TH.by_slot(part)
            total += ctx.thrashing_note_events(t, part)
            TH.run_oracle(orc, part, n_slots, lapse, threshold, st=want)
        got = gpu_state(ctx, t, total)
        ctx.thrashing_destroy(t)
    TH.assert_state_equal(want, got, "seed %d" % seed)
    assert total[2] == 0 and want.counters[0] > 20 and int(want.blocks["num_thrashing_pages"].sum()) == int(
        sum(bin(int(w)).count("1") for w in want.thrashing_pages))


def test_events_that_cannot_be_placed_are_dropped_and_counted():
    layout = S.make_layout(dict(va_base=TG.VA_BASE, blocks_per_space=4, n_spaces=1))
    rows = [(1000, 1, 0, 8, 0, 0, 0), (2000, 1, 0, 8, 1, 0, 0), (3000, 9, 0, 8, 0, 0, 0), (4000, 2, 510, 8, 0, 0, 0),
            (5000, 2, 0, 8, 40, 0, 0), (6000, 1, 0, 8, 0, 0, 0), (7000, 1, 0, 8, 1, 0, 0)]   # slot 9, page 518, processor 40, slot 1 again
    ev = np.zeros(len(rows), TH.EVENT_DT)
    for i, r in enumerate(rows):
        ev[i] = tuple(r) + (0,)
    with uvmtier.Context(0) as ctx:
        t = ctx.thrashing_create(layout)
        assert t.lapse_ns == 500000 and t.threshold == 3
        cnt = ctx.thrashing_note_events(t, ev)
        tb = ctx.thrashing_tables(t)
        assert ctx.thrashing_note_events(t, ev[:0]).tolist() == [0, 0, 0, 0]
        ctx.thrashing_destroy(t)
    assert cnt.tolist() == [0, 0, 5, 0]
    blocks = tb["blocks"].view(TH.BLOCK_DT)
    assert blocks["last_time_stamp"].tolist() == [0, 2000, 0, 0] and blocks["tracked"].tolist() == [0, 1, 0, 0]


def test_thrashing_entry_points_reject_bad_calls():
    import ctypes
    lib = uvmtier.lib()
    layout = S.make_layout(dict(va_base=TG.VA_BASE, blocks_per_space=4, n_spaces=1))
    INVALID_ARGUMENT = 0x1F
    with uvmtier.Context(0) as ctx:
        h = ctx.handle
        launches = uvmtier.kernel_launch_count()
        t = uvmtier.ThrashingState()
        assert lib.uvmtier_thrashing_create(h, uvmtier.P(layout), ctypes.c_uint64(0), ctypes.c_uint32(8), ctypes.byref(t)) == INVALID_ARGUMENT
        assert lib.uvmtier_thrashing_create(h, None, ctypes.c_uint64(0), ctypes.c_uint32(3), ctypes.byref(t)) == INVALID_ARGUMENT
        d = ctx.alloc(64)
        # a table that was never created, missing counters, events announced but not given
        assert lib.uvmtier_thrashing_note_events_device(h, ctypes.byref(t), ctypes.c_uint32(1), uvmtier.P(d), uvmtier.P(d)) == INVALID_ARGUMENT
        t = ctx.thrashing_create(layout)
        assert lib.uvmtier_thrashing_note_events_device(h, ctypes.byref(t), ctypes.c_uint32(1), uvmtier.P(d), None) == INVALID_ARGUMENT
        assert lib.uvmtier_thrashing_note_events_device(h, ctypes.byref(t), ctypes.c_uint32(1), None, uvmtier.P(d)) == INVALID_ARGUMENT
        assert uvmtier.kernel_launch_count() == launches
        ctx.thrashing_destroy(t)
        assert t.d_blocks is None and t.n_slots == 0
        ctx.free(d)


# ---------------- the hint half on the GPU ----------------
def _same_hints(a, b, what):
    assert len(a) == len(b)
    for i, (x, y) in enumerate(zip(a, b)):
        for f in ("type", "pin_residency", "pin_processors", "throttle_end"):
            assert np.array_equal(x[f], y[f]), "%s: hint batch %d field %s: want %s got %s" % (what, i, f, x[f], y[f])


@pytest.mark.parametrize("seed,topology,procs,kw", [(41, "pcie", (0, 1), dict(pin_threshold=3)), (42, "c2c", (0, 1, 2), dict(threshold=1, epoch=30, max_resets=2)),
                                                    (43, "nvlink", (0, 1, 9), dict(pin_threshold=2, pin=1))])
def test_gpu_hints_match_restatement_on_driver_histories(seed, topology, procs, kw):
    """uvmtier_thrashing_get_hints_device / _note_events_device / _unpin_expired_device interleaved the way a driver would
    call them, against the restatement (held against the reference's own code in tests/test_oracle_thrashing.py): every
    hint, every field of the tracking state, the counters."""
    import uvmtier
    from uvmtier import blockstate as BS, tracegen as TG
    orc = T.load_oracle()
    cfg = TH.make_cfg(topology, gpus=tuple(p for p in (1, 2, 9) if p in procs), **kw)
    n_slots = 6
    steps, _ = TH.driver_history(n_slots, 2500, seed, cfg, procs=procs, hot_pages=4)
    want, want_h = TH.run_steps_oracle(orc, steps, n_slots, cfg)
    layout = BS.make_layout(dict(va_base=TG.VA_BASE, blocks_per_space=n_slots, n_spaces=1))
    with uvmtier.Context(0) as ctx:
        got, got_h = TH.run_steps_gpu(ctx, steps, layout, n_slots, cfg)
    _same_hints(want_h, got_h, "seed %d" % seed)
    TH.assert_hint_state_equal(want, got, "seed %d" % seed)
    types = np.concatenate([h["type"] for h in want_h])
    assert (types == TH.HINT_THROTTLE).sum() > 3 and (types == TH.HINT_PIN).sum() > 0


def test_gpu_hints_match_restatement_on_random_batches_and_many_blocks():
    """Batches of queries over 512 va_blocks (one thread per va_block's group), events and expiries in between."""
    import uvmtier
    from uvmtier import blockstate as BS, tracegen as TG
    orc = T.load_oracle()
    cfg = TH.make_cfg("pcie", gpus=(1, 2, 9), pin_threshold=3, epoch=50, pin=20)
    n_slots = 512
    steps = TH.random_stream(77, n_slots, 400, batch=3000)
    want, want_h = TH.run_steps_oracle(orc, steps, n_slots, cfg)
    layout = BS.make_layout(dict(va_base=TG.VA_BASE, blocks_per_space=n_slots, n_spaces=1))
    with uvmtier.Context(0) as ctx:
        got, got_h = TH.run_steps_gpu(ctx, steps, layout, n_slots, cfg)
        # the hint half has to be switched on first, and only once
        t = ctx.thrashing_create(layout)
        lib = uvmtier.lib()
        import ctypes
        one = ctx.alloc(64)
        assert lib.uvmtier_thrashing_get_hints_device(ctx.handle, ctypes.byref(t), ctypes.c_uint32(1), uvmtier.P(one), uvmtier.P(one), uvmtier.P(one)) == 0x40
        ctx.thrashing_enable_hints(t, cfg)
        assert lib.uvmtier_thrashing_enable_hints(ctx.handle, ctypes.byref(t), uvmtier.P(cfg)) == 0x40
        ctx.free(one)
        ctx.thrashing_destroy(t)
    _same_hints(want_h, got_h, "batches")
    TH.assert_hint_state_equal(want, got, "batches")
    types = np.concatenate([h["type"] for h in want_h])
    assert (types == TH.HINT_THROTTLE).sum() > 50 and (types == TH.HINT_PIN).sum() > 50


@pytest.mark.parametrize("seed,topology,kw", [(1, "pcie", {}), (2, "c2c", {}), (3, "pcie", dict(pin_threshold=1, pin=2))])
def test_gpu_service_with_hints_matches_restatement(seed, topology, kw):
    """uvmtier_service_fault_batches_thrashing_device: throttled / pinned / pinned-elsewhere entries, records, copy runs, block
    state and the hint tables against the restated loop (which tests/test_oracle_thrashing_service.py holds against the
    reference's loop compiled verbatim)."""
    import faultlib as F
    import thrashsvclib as TS
    import uvmtier
    orc = T.load_oracle()
    sc = TS.scenario(seed, topology, n_faults=6000, blocks=48, **kw)
    with uvmtier.Context(0) as ctx:
        ctx.fault_buffer_init(sc["gp"], sc["tbl"])
        n = len(sc["pk"])
        ring = np.zeros((n + 1, 8), np.uint32)
        ring[:n] = sc["pk"]
        r = ctx.fetch_preprocess_fault_batches(ring, 0, n)
        B = sc["B"]
        pad = len(r["batch"]) * B - n
        r["ordered"] = np.concatenate([r["ordered"], np.zeros(pad, r["ordered"].dtype)])
        r["groups"] = np.concatenate([r["groups"], np.zeros(pad, r["groups"].dtype)])
        k1 = dict(ordered_rec=r["ordered"], groups=r["groups"], batch=r["batch"])
        want = TS.run_oracle(orc, sc, k1)
        got = TS.run_gpu(ctx, sc, r)
    throttled, pinned, remote = TS.check(want, got, B, "seed %d %s" % (seed, topology))
    assert np.array_equal(want["thrash"].counters[3:6], got["thrash"].counters[3:6])
    assert throttled + pinned + remote > 10, (throttled, pinned, remote)


def test_a_va_block_in_two_groups_keeps_the_group_that_starts_first():
    """Every va_block shows up in two separate groups, far apart in the array (different CTAs take their heads): the first
    group is the one that is applied, the second is dropped and counted -- on every run, not by the luck of the schedule."""
    n_slots = 4096
    layout = S.make_layout(dict(va_base=TG.VA_BASE, blocks_per_space=n_slots, n_spaces=1))
    ev = np.zeros(2 * n_slots * 2, TH.EVENT_DT)
    half = n_slots * 2
    for part, (t0, proc) in enumerate(((1000, 0), (900000, 1))):
        sl = slice(part * half, (part + 1) * half)
        ev["slot"][sl] = np.repeat(np.arange(n_slots), 2)
        ev["time_stamp"][sl] = t0 + np.tile([0, 10], n_slots)
        ev["num_pages"][sl] = 4
        ev["processor"][sl] = proc
    with uvmtier.Context(0) as ctx:
        for _ in range(3):
            t = ctx.thrashing_create(layout)
            cnt = ctx.thrashing_note_events(t, ev)
            blocks = ctx.thrashing_tables(t)["blocks"].view(TH.BLOCK_DT)
            ctx.thrashing_destroy(t)
            assert cnt[2] == half
            assert (blocks["last_time_stamp"] == 1010).all() and (blocks["last_processor"] == 0).all()
