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
