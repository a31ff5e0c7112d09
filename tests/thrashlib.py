"""Event streams and state arrays for the thrashing detector (section 1f of include/uvmtier.h)."""
import ctypes

import numpy as np

import uvmtier_testlib as T

P = T.P
EVENT_DT = np.dtype([("time_stamp", "<u8"), ("slot", "<u4"), ("first_page", "<u2"), ("num_pages", "<u2"), ("processor", "<u2"),
                     ("kind", "u1"), ("flags", "u1"), ("reserved", "<u4")])
BLOCK_DT = np.dtype([("last_time_stamp", "<u8"), ("last_thrashing_time_stamp", "<u8"), ("last_processor", "<u2"),
                     ("num_thrashing_pages", "<u2"), ("tracked", "<u4"), ("reserved", "<u8")])
assert EVENT_DT.itemsize == 24 and BLOCK_DT.itemsize == 32
KIND_MIGRATION, KIND_REVOCATION, EV_RESET = 0, 1, 1
LAPSE_NS, THRESHOLD = 500 * 1000, 3                     # uvm_perf_thrashing.c:246,264


def ping_pong_events(n_slots, n_events, seed, n_procs=3, hot_frac=0.3, reset_frac=0.02, revoke_frac=0.15, lapse_ns=LAPSE_NS):
    """Migrations and revocations in time order.  A share of the va_blocks is fought over by several processors with
    gaps around the window (so counters rise, saturate and get reset), the rest sees one processor or slow traffic;
    now and then a user-driven migration clears a region."""
    rng = np.random.default_rng(seed)
    ev = np.zeros(n_events, EVENT_DT)
    hot = rng.random(n_slots) < hot_frac
    clock = np.full(n_slots, 1000, np.int64)            # every va_block has its own pace; merged by time below
    for i in range(n_events):
        slot = int(rng.integers(0, n_slots))
        r = rng.random()
        if hot[slot]:
            clock[slot] += int(rng.choice([40, 300, 3000, lapse_ns // 8, lapse_ns + 64, 3 * lapse_ns], p=[0.2, 0.3, 0.3, 0.1, 0.05, 0.05]))
            proc = int(rng.integers(0, n_procs))
            first = int(rng.choice([0, 16, 100, 480])) + int(rng.integers(0, 4))
            num = int(rng.choice([1, 2, 16, 31, 64]))
        else:
            clock[slot] += int(rng.integers(1, 2 * lapse_ns))
            proc = int(slot % n_procs) if r < 0.8 else int(rng.integers(0, n_procs))
            first = int(rng.integers(0, 512))
            num = int(rng.integers(1, 40))
        num = min(num, 512 - first)
        kind = KIND_REVOCATION if rng.random() < revoke_frac else KIND_MIGRATION
        flags = EV_RESET if kind == KIND_MIGRATION and rng.random() < reset_frac else 0
        if flags:
            first, num = (0, 512) if rng.random() < 0.5 else (first, num)
        ev[i] = (int(clock[slot]), slot, first, num, proc, kind, flags, 0)
    return ev[np.argsort(ev["time_stamp"], kind="stable")]


def by_slot(ev):
    """The order the device entry point wants: grouped by va_block, time order kept inside a va_block."""
    return ev[np.argsort(ev["slot"], kind="stable")]


class State:
    def __init__(self, n_slots):
        self.n = n_slots
        self.blocks = np.zeros(n_slots, BLOCK_DT)
        self.page_word = np.zeros(n_slots * 512, "<u8")
        self.page_procs = np.zeros(n_slots * 512, "<u4")
        self.thrashing_pages = np.zeros(n_slots * 16, "<u4")
        self.counters = np.zeros(4, "<u4")

    def fields(self):
        return dict(blocks=self.blocks, page_word=self.page_word, page_procs=self.page_procs,
                    thrashing_pages=self.thrashing_pages, counters=self.counters[:2])


def run_oracle(orc, ev, n_slots, lapse_ns=LAPSE_NS, threshold=THRESHOLD, st=None):
    st = st or State(n_slots)
    orc.oracle_thrashing_note_events(ctypes.c_uint32(len(ev)), P(ev), ctypes.c_uint64(lapse_ns), ctypes.c_uint32(threshold),
                                     P(st.blocks), P(st.page_word), P(st.page_procs), P(st.thrashing_pages), P(st.counters))
    return st


def run_ref(ref, ev, n_slots, lapse_ns=LAPSE_NS, threshold=THRESHOLD):
    st = State(n_slots)
    ref.ref_thrashing_note_events(ctypes.c_uint32(n_slots), ctypes.c_uint32(len(ev)), P(ev), ctypes.c_uint64(lapse_ns),
                                  ctypes.c_uint32(threshold), P(st.blocks), P(st.page_word), P(st.page_procs),
                                  P(st.thrashing_pages), P(st.counters))
    return st


def assert_state_equal(want, got, what=""):
    w, g = want.fields(), got.fields()
    for k in w:
        if k == "blocks":
            for f in ("last_time_stamp", "last_thrashing_time_stamp", "last_processor", "num_thrashing_pages", "tracked"):
                bad = np.nonzero(w[k][f] != g[k][f])[0]
                assert len(bad) == 0, "%s blocks.%s differs at slot %d: want %s got %s" % (what, f, bad[0], w[k][f][bad[0]], g[k][f][bad[0]])
        else:
            bad = np.nonzero(w[k] != g[k])[0]
            assert len(bad) == 0, "%s %s differs at %d: want %#x got %#x" % (what, k, bad[0], int(w[k][bad[0]]), int(g[k][bad[0]]))


# ---------------- the hint half ----------------
CFG_DT = np.dtype([("lapse_ns", "<u8"), ("nap_ns", "<u8"), ("epoch_ns", "<u8"), ("pin_ns", "<u8"), ("threshold", "<u4"),
                   ("pin_threshold", "<u4"), ("max_resets", "<u4"), ("has_memory", "<u4"), ("accessible_from", "<u4", (32,)),
                   ("can_access", "<u4", (32,)), ("has_fast_link", "<u4", (32,)), ("has_native_atomics", "<u4", (32,))])
QUERY_DT = np.dtype([("time_stamp", "<u8"), ("slot", "<u4"), ("page", "<u2"), ("requester", "u1"), ("pad", "u1"),
                     ("resident_processors", "<u4"), ("preferred_location", "<u2"), ("pad2", "<u2")])
HINT_DT = np.dtype([("throttle_end", "<u8"), ("pin_processors", "<u4"), ("type", "u1"), ("pin_residency", "u1"), ("pad", "<u2")])
HBLOCK_DT = np.dtype([("last_time_stamp", "<u8"), ("last_thrashing_time_stamp", "<u8"), ("last_processor", "<u2"),
                      ("num_thrashing_pages", "<u2"), ("tracked", "<u4"), ("pinned_count", "<u2"), ("reset_count", "u1"), ("pad", "u1"),
                      ("throttling_count", "<u4")])
assert (CFG_DT.itemsize, QUERY_DT.itemsize, HINT_DT.itemsize, HBLOCK_DT.itemsize) == (560, 24, 16, 32)
HINT_NONE, HINT_PIN, HINT_THROTTLE = 0, 1, 2
EV_PREFETCH = 2
NO_PREF = 0xFFFF


def make_cfg(topology="pcie", gpus=(1, 9), lapse_ns=LAPSE_NS, threshold=THRESHOLD, pin_threshold=10, nap=1, epoch=2000, pin=300,
             max_resets=4):
    """uvm_perf_thrashing.c:246-310 defaults (nap, epoch, pin are multiples of the lapse).  Topologies: "pcie" -- the CPU
    reaches only its own memory, GPUs reach the CPU's and each other's, no fast links; "nvlink" -- the same with fast links
    and native atomics between the GPUs; "c2c" -- everybody reaches everybody over fast links with native atomics."""
    cfg = np.zeros(1, CFG_DT)
    cfg["lapse_ns"], cfg["threshold"], cfg["pin_threshold"], cfg["max_resets"] = lapse_ns, threshold, pin_threshold, max_resets
    cfg["nap_ns"], cfg["epoch_ns"], cfg["pin_ns"] = nap * lapse_ns, epoch * lapse_ns, pin * lapse_ns
    allp = 1 | sum(1 << g for g in gpus)
    gmask = allp & ~1
    cfg["has_memory"] = allp
    for p in (0,) + tuple(gpus):
        me = 1 << p
        if topology == "c2c":
            acc_from = can = fast = atom = allp
        else:
            acc_from = allp if p == 0 else gmask            # who can reach p's memory
            can = me if p == 0 else allp                    # whose memory p can reach
            fast = (gmask & ~me) if (topology == "nvlink" and p) else 0
            atom = me | fast
        cfg["accessible_from"][0, p], cfg["can_access"][0, p], cfg["has_fast_link"][0, p], cfg["has_native_atomics"][0, p] = acc_from, can, fast, atom
    return cfg


class HintState:
    """Detector + hint tables on the host."""
    def __init__(self, n_slots):
        self.n = n_slots
        self.blocks = np.zeros(n_slots, HBLOCK_DT)
        self.page_word = np.zeros(n_slots * 512, "<u8")
        self.page_procs = np.zeros(n_slots * 512, "<u4")
        self.thrashing_pages = np.zeros(n_slots * 16, "<u4")
        self.page_throttle = np.zeros(n_slots * 512, "<u8")
        self.page_hint = np.zeros(n_slots * 512, "<u4")
        self.page_throttled = np.zeros(n_slots * 512, "<u4")
        self.page_deadline = np.zeros(n_slots * 512, "<u8")
        self.pinned_pages = np.zeros(n_slots * 16, "<u4")
        self.counters = np.zeros(8, "<u4")

    NAMES = ("blocks", "page_word", "page_procs", "thrashing_pages", "page_throttle", "page_hint", "page_throttled", "page_deadline",
             "pinned_pages", "counters")

    def tables(self):
        return (ctypes.c_void_p * 10)(*[getattr(self, k).ctypes.data for k in self.NAMES])


def assert_hint_state_equal(want, got, what=""):
    for k in HintState.NAMES:
        a, b = getattr(want, k), getattr(got, k)
        if k == "blocks":
            for f in HBLOCK_DT.names:
                if f == "pad":
                    continue
                bad = np.nonzero(a[f] != b[f])[0]
                assert len(bad) == 0, "%s blocks.%s differs at slot %d: want %s got %s" % (what, f, bad[0], a[f][bad[0]], b[f][bad[0]])
        else:
            n = 6 if k == "counters" else len(a)
            bad = np.nonzero(a[:n] != b[:n])[0]
            assert len(bad) == 0, "%s %s differs at %d (slot %d page %d): want %#x got %#x" % (
                what, k, bad[0], bad[0] // 512, bad[0] % 512, int(a[bad[0]]), int(b[bad[0]]))


def run_steps_oracle(orc, steps, n_slots, cfg):
    """steps: ("events", array) | ("hints", array) | ("expire", now).  Returns (state, list of hint arrays)."""
    st, hints = HintState(n_slots), []
    tb = st.tables()
    for kind, arg in steps:
        if kind == "events":
            orc.oracle_thrashing_note_events_tables(ctypes.c_uint32(len(arg)), P(arg), P(cfg), tb)
        elif kind == "hints":
            out = np.zeros(len(arg), HINT_DT)
            orc.oracle_thrashing_get_hints(ctypes.c_uint32(len(arg)), P(arg), P(cfg), tb, P(out))
            hints.append(out)
        else:
            orc.oracle_thrashing_unpin_expired(ctypes.c_uint32(n_slots), ctypes.c_uint64(arg), tb)
    return st, hints


def run_steps_ref(ref, steps, n_slots, cfg):
    ref.ref_thrash_open.restype = ctypes.c_void_p
    h = ctypes.c_void_p(ref.ref_thrash_open(ctypes.c_uint32(n_slots), P(cfg)))
    hints = []
    for kind, arg in steps:
        if kind == "events":
            ref.ref_thrash_events(h, ctypes.c_uint32(len(arg)), P(arg))
        elif kind == "hints":
            out = np.zeros(len(arg), HINT_DT)
            ref.ref_thrash_hints(h, ctypes.c_uint32(len(arg)), P(arg), P(out))
            hints.append(out)
        else:
            ref.ref_thrash_unpin_expired(h, ctypes.c_uint64(arg))
    st = HintState(n_slots)
    ref.ref_thrash_export(h, *[P(getattr(st, k)) for k in HintState.NAMES])
    ref.ref_thrash_close(h)
    return st, hints


def driver_history(n_slots, n_faults, seed, cfg, procs=(0, 1), hot_pages=6, pref_frac=0.0, query_batch=1):
    """A coherent history, the way a driver produces it: processors take turns faulting on a few hot pages; every fault
    asks for a hint and acts on it -- NONE: the page moves to the requester (a MIGRATION event); PIN: it moves to the
    pin's residency if it is not there (an event on a pinned page: not counted); THROTTLE: nothing happens.  Now and
    then a revocation, a long pause (window / epoch), the expiry worker, a user migration.  The result is a list of
    steps for run_steps_*; hints are asked one by one (each depends on the previous one's effect)."""
    rng = np.random.default_rng(seed)
    lapse = int(cfg["lapse_ns"][0])
    now = 10_000
    where = {}                                            # (slot, page) -> processor holding it
    steps = []
    hot = [(int(rng.integers(0, n_slots)), int(rng.integers(0, 512))) for _ in range(hot_pages)]
    sim = HintState(n_slots)                              # the generator runs the restatement to learn each hint
    orc = T.load_oracle()
    tb = sim.tables()
    pref = {s: (int(rng.choice(procs)) if rng.random() < pref_frac else NO_PREF) for s, _ in hot}
    for i in range(n_faults):
        r = rng.random()
        now += int(rng.choice([200, 3000, lapse // 4, lapse + 100, 40 * lapse, 2500 * lapse], p=[0.45, 0.3, 0.15, 0.07, 0.025, 0.005]))
        slot, page = hot[int(rng.integers(0, len(hot)))]
        req = int(rng.choice(procs))
        cur = where.get((slot, page), 0)
        if r < 0.04:
            steps.append(("expire", now))
            orc.oracle_thrashing_unpin_expired(ctypes.c_uint32(n_slots), ctypes.c_uint64(now), tb)
            continue
        if r < 0.10:                                      # a permission revocation on the page
            ev = np.zeros(1, EVENT_DT); ev[0] = (now, slot, page, 1, req, KIND_REVOCATION, 0, 0)
            steps.append(("events", ev))
            orc.oracle_thrashing_note_events_tables(ctypes.c_uint32(1), P(ev), P(cfg), tb)
            continue
        if r < 0.115:                                     # cudaMemPrefetchAsync-like user migration of the whole va_block
            ev = np.zeros(1, EVENT_DT); ev[0] = (now, slot, 0, 512, req, KIND_MIGRATION, EV_RESET, 0)
            steps.append(("events", ev))
            orc.oracle_thrashing_note_events_tables(ctypes.c_uint32(1), P(ev), P(cfg), tb)
            for k in list(where):
                if k[0] == slot:
                    where[k] = req
            continue
        if req == cur and rng.random() < 0.7:
            continue                                      # no fault: the page is local
        q = np.zeros(1, QUERY_DT)
        q[0] = (now, slot, page, req, 0, 1 << cur, pref[slot], 0)
        out = np.zeros(1, HINT_DT)
        steps.append(("hints", q))
        orc.oracle_thrashing_get_hints(ctypes.c_uint32(1), P(q), P(cfg), tb, P(out))
        dst = None
        if out["type"][0] == HINT_NONE:
            dst = req
        elif out["type"][0] == HINT_PIN and int(out["pin_residency"][0]) != cur:
            dst = int(out["pin_residency"][0])
        if dst is not None and dst != cur:
            ev = np.zeros(1, EVENT_DT); ev[0] = (now + 50, slot, page, 1, dst, KIND_MIGRATION, 0, 0)
            steps.append(("events", ev))
            orc.oracle_thrashing_note_events_tables(ctypes.c_uint32(1), P(ev), P(cfg), tb)
            where[(slot, page)] = dst
    return steps, sim


def run_steps_simt(sim, steps, n_slots, cfg):
    st, hints = HintState(n_slots), []
    tb = st.tables()
    for kind, arg in steps:
        if kind == "events":
            sim.simt_thrash_events(ctypes.c_uint32(n_slots), ctypes.c_uint32(len(arg)), P(arg), ctypes.c_uint64(int(cfg["lapse_ns"][0])),
                                   ctypes.c_uint32(int(cfg["threshold"][0])), tb)
        elif kind == "hints":
            out = np.zeros(len(arg), HINT_DT)
            sim.simt_thrash_hints(ctypes.c_uint32(n_slots), ctypes.c_uint32(len(arg)), P(arg), P(cfg), tb, P(out))
            hints.append(out)
        else:
            sim.simt_thrash_unpin_expired(ctypes.c_uint32(n_slots), ctypes.c_uint64(arg), tb)
    return st, hints


def run_steps_gpu(ctx, steps, layout, n_slots, cfg):
    """The same steps through the C ABI on the GPU; returns (HintState downloaded, hints)."""
    t = ctx.thrashing_create(layout, int(cfg["lapse_ns"][0]), int(cfg["threshold"][0]))
    ctx.thrashing_enable_hints(t, cfg)
    hints = []
    cnt = np.zeros(8, np.int64)
    try:
        for kind, arg in steps:
            if kind == "events":
                c = ctx.thrashing_note_events(t, arg)
                cnt[:3] += c[:3]
            elif kind == "hints":
                h, c = ctx.thrashing_get_hints(t, arg)
                hints.append(h.view(HINT_DT))
                cnt[2:6] += c[2:6]
            else:
                cnt[1] += ctx.thrashing_unpin_expired(t, arg)[1]
        tabs = ctx.thrashing_tables(t)
    finally:
        ctx.thrashing_destroy(t)
    st = HintState(n_slots)
    st.blocks = tabs["blocks"].view(HBLOCK_DT)
    for k in ("page_word", "page_procs", "thrashing_pages", "page_throttle", "page_hint", "page_throttled", "page_deadline", "pinned_pages"):
        setattr(st, k, tabs[k])
    st.counters[:6] = cnt[:6]
    return st, hints


def random_stream(seed, n_slots, n_steps, procs=(0, 1, 2, 9), batch=6):
    """Events, query batches (grouped by slot, time order inside a slot) and expiries; not a coherent driver history."""
    rng = np.random.default_rng(seed)
    procs = np.array(procs)
    now, steps = 5000, []
    for _ in range(n_steps):
        now += int(rng.choice([100, 2000, LAPSE_NS // 3, LAPSE_NS * 2, LAPSE_NS * 60], p=[0.4, 0.3, 0.2, 0.08, 0.02]))
        r = rng.random()
        if r < 0.5:
            n = int(rng.integers(1, batch))
            ev = np.zeros(n, EVENT_DT)
            for j in range(n):
                kind = int(rng.random() < 0.2)
                flags = 0 if kind else int(rng.choice([0, 0, 0, EV_RESET, EV_PREFETCH]))
                ev[j] = (now + j, int(rng.integers(0, n_slots)), int(rng.integers(0, 8)), int(rng.integers(1, 4)), int(rng.choice(procs)), kind, flags, 0)
            steps.append(("events", by_slot(ev)))
        elif r < 0.97:
            n = int(rng.integers(1, batch + 1))
            q = np.zeros(n, QUERY_DT)
            for j in range(n):
                res = int(rng.integers(1, 1 << len(procs)))
                resident = sum(1 << int(procs[b]) for b in range(len(procs)) if (res >> b) & 1)
                pref = int(rng.choice([NO_PREF, NO_PREF] + [int(p) for p in procs]))
                q[j] = (now + j, int(rng.integers(0, n_slots)), int(rng.integers(0, 10)), int(rng.choice(procs)), 0, resident, pref, 0)
            steps.append(("hints", q[np.argsort(q["slot"], kind="stable")]))
        else:
            steps.append(("expire", now))
    return steps
