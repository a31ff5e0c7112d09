"""The servicing pass with the thrashing hint inside the fault loop (uvmtier_service_fault_batches_thrashing_device):
scenario builder and runners for the restatement, the emulator and the GPU."""
import ctypes

import numpy as np

import servicelib as S
import thrashlib as TH
import uvmtier_testlib as T
from uvmtier import tracegen as TG

P = T.P


def scenario(seed, topology="pcie", n_faults=2500, B=64, blocks=24, gpu_id=1, **cfg_kw):
    """A fault trace on a few va_blocks, a state table, and a thrashing history in which the CPU (0) and this GPU have been
    fighting over many of the trace's pages: tracked va_blocks with thrashing, throttled and pinned pages, next to
    va_blocks the detector has never looked at."""
    rng = np.random.default_rng(seed)
    orc = T.load_oracle()
    pk, tbl, lay = TG.fault_trace(n_faults, seed, "uniform", 1, total_va_bytes=blocks << 21)
    gp = TG.gpu_params(B)
    n_slots = lay["blocks_per_space"]
    st0 = S.random_state(n_slots, seed + 1, gpu_id=gpu_id)
    cfg = TH.make_cfg(topology, gpus=(gpu_id,), **cfg_kw)
    lapse = int(cfg["lapse_ns"][0])
    th = TH.HintState(n_slots)
    tb = th.tables()
    # history: on two thirds of the va_blocks the two processors take whole 64 KB regions from each other, close enough in
    # time for the pages to count as thrashing; in between both ask for hints (throttled and pinned pages stay behind)
    now = 10_000
    regions = {s: rng.choice(np.arange(0, 512, 16), 8, replace=False) for s in range(n_slots) if s % 3 != 2}
    for round_ in range(10):
        ev = []
        for s, firsts in regions.items():
            for first in firsts:
                ev.append((now + int(rng.integers(0, lapse // 16)), s, int(first), 16, gpu_id if (round_ + s) & 1 else 0, 0, 0, 0))
        ev = np.array(ev, TH.EVENT_DT)
        ev = ev[np.argsort(ev["slot"], kind="stable")]
        orc.oracle_thrashing_note_events_tables(ctypes.c_uint32(len(ev)), P(ev), P(cfg), tb)
        now += lapse // 8
        if round_ >= 4:
            thr = np.nonzero(np.unpackbits(th.thrashing_pages.view(np.uint8), bitorder="little"))[0]
            if len(thr):
                pick = rng.choice(thr, min(len(thr), 300), replace=False)
                q = np.zeros(len(pick), TH.QUERY_DT)
                q["time_stamp"] = now
                q["slot"] = pick // 512
                q["page"] = pick % 512
                q["requester"] = rng.choice([0, gpu_id], len(pick))
                q["resident_processors"] = rng.choice([1, 1 << gpu_id], len(pick))
                q["preferred_location"] = TH.NO_PREF
                q = q[np.argsort(q["slot"], kind="stable")]
                out = np.zeros(len(q), TH.HINT_DT)
                orc.oracle_thrashing_get_hints(ctypes.c_uint32(len(q)), P(q), P(cfg), tb, P(out))
                now += lapse // 16
    th.counters[:] = 0
    k1 = None
    layout = S.make_layout(lay, gpu_id=gpu_id)
    return dict(pk=pk, tbl=tbl, lay=lay, gp=gp, layout=layout, st0=st0, cfg=cfg, th=th, now=now, B=B, n_slots=n_slots)


def batch_times(sc, nb):
    """NV_GETTIME() per batch: mostly close together, now and then a pause longer than the nap or the window."""
    rng = np.random.default_rng(99)
    lapse = int(sc["cfg"]["lapse_ns"][0])
    steps = rng.choice([2000, lapse // 20, lapse // 4, 2 * lapse], nb, p=[0.55, 0.3, 0.13, 0.02])
    return (sc["now"] + np.cumsum(steps)).astype("<u8")


def copy_hint_state(th):
    c = TH.HintState(th.n)
    for k in TH.HintState.NAMES:
        setattr(c, k, getattr(th, k).copy())
    return c


def run_oracle(orc, sc, k1, pf=(1, 51, 1)):
    th = copy_hint_state(sc["th"])
    tb = th.tables()
    times = batch_times(sc, len(k1["batch"]))
    orc.oracle_service_set_thrashing(P(sc["cfg"]), tb, P(times))
    try:
        res = S.run_service_oracle(orc, k1, sc["gp"], sc["tbl"], sc["layout"], sc["st0"], pf)
    finally:
        orc.oracle_service_set_thrashing(None, None, None)
    res["thrash"] = th
    return res


def run_simt(sim, sc, k1, pf=(1, 51, 1), heavy_min=96, run_cap=1 << 20):
    th = copy_hint_state(sc["th"])
    tb = th.tables()
    times = batch_times(sc, len(k1["batch"]))
    B, nb = sc["B"], len(k1["batch"])
    flags, svc, runs, counters = S._alloc_outputs(nb, B, run_cap)
    batch = k1["batch"].copy()
    st = sc["st0"].copy()
    ptrs = st.ptrs()
    sim.simt_service_fault_batches_thrashing(ctypes.c_uint32(nb), ctypes.c_uint32(B), P(k1["ordered_rec"]), P(k1["groups"]), P(batch),
                                             P(sc["tbl"]), P(sc["layout"]), ctypes.byref(ptrs), ctypes.c_uint32(pf[0]), ctypes.c_uint32(pf[1]),
                                             ctypes.c_uint32(pf[2]), P(flags), P(svc), P(runs), ctypes.c_uint32(run_cap), P(counters),
                                             ctypes.c_uint32(heavy_min), tb, P(sc["cfg"]), P(times))
    return dict(flags=flags, service=svc, runs=runs, n_runs=int(counters[0]), batch=batch, state=st, thrash=th)


def check(want, got, B, what):
    S.assert_service_equal(want, got, B, what)
    TH.assert_hint_state_equal(want["thrash"], got["thrash"], what)
    fl = want["flags"]
    return int((fl & 0x2000 != 0).sum()), int((fl & 0x4000 != 0).sum()), int((fl & 0x8000 != 0).sum())


def run_ref(ref, sc, k1, pf=(1, 51, 1)):
    """libref_service.so: the reference's own service_fault_batch_block_locked (verbatim) with its call of
    uvm_perf_thrashing_get_hint answered by the thrashing restatement on a copy of the scenario's tables."""
    th = copy_hint_state(sc["th"])
    tb = th.tables()
    times = batch_times(sc, len(k1["batch"]))
    ref.ref_service_set_thrashing(P(sc["cfg"]), tb, P(times))
    try:
        res = S.run_service_ref(ref, k1, sc["gp"], sc["tbl"], sc["layout"], sc["st0"], pf)
    finally:
        ref.ref_service_set_thrashing(None, None, None)
    res["thrash"] = th
    return res


def run_gpu(ctx, sc, r, pf=None):
    """Through the C ABI: r = the context's own preprocess output for sc's trace (padded to whole batches)."""
    th0 = sc["th"]
    t = ctx.thrashing_create(sc["layout"], int(sc["cfg"]["lapse_ns"][0]), int(sc["cfg"]["threshold"][0]))
    ctx.thrashing_enable_hints(t, sc["cfg"])
    ctx.thrashing_upload(t, {k: getattr(th0, k) for k in TH.HintState.NAMES if k != "counters"})
    ds = ctx.block_state_alloc(sc["layout"])
    ds.upload(sc["st0"])
    try:
        got = ctx.service_fault_batches_thrashing(r, sc["layout"], ds, t, batch_times(sc, len(r["batch"])))
        tabs = ctx.thrashing_tables(t)
    finally:
        ds.free()
        ctx.thrashing_destroy(t)
    th = TH.HintState(th0.n)
    th.blocks = tabs["blocks"].view(TH.HBLOCK_DT)
    for k in ("page_word", "page_procs", "thrashing_pages", "page_throttle", "page_hint", "page_throttled", "page_deadline", "pinned_pages"):
        setattr(th, k, tabs[k])
    th.counters[:6] = got["thrashing_counters"][:6]
    got["thrash"] = th
    return got
