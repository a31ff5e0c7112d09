"""CPU tests of the root-chunk eviction order (section 1d of include/uvmtier.h): the stamp restatement against the
reference's own list code compiled verbatim (oracle/_ref/libref_pmm.so), and the product's kernels under the SIMT
emulator against the restatement -- the list moves of a servicing pass and of make_resident calls, the count of
chunks a batch needs, and the selection of the k least recently used chunks."""
import ctypes

import numpy as np
import pytest

import faultlib as F
import residentlib as R
import servicelib as S
import uvmtier_testlib as T
from uvmtier import tracegen as TG

P = T.P
USED = 1 << 63
EV_ALLOC, EV_USED, EV_UNUSED, EV_PICK = 0, 1, 2, 3


def random_events(n_slots, n_events, seed):
    """A plausible life of the lists: chunks are allocated, their blocks serviced again and again, some lose their
    last GPU page, and every now and then the allocator asks for a victim."""
    rng = np.random.default_rng(seed)
    ev = np.zeros((n_events, 2), "<u4")
    order = {}                                          # slot -> (on the used list, sequence): only to know who is free
    for i in range(n_events):
        r = rng.random()
        slot = int(rng.integers(0, n_slots))
        if r < 0.12:
            ev[i] = (EV_PICK, 0)
            if order:
                del order[min(order, key=lambda s: order[s])]
        elif slot not in order:
            ev[i] = (EV_ALLOC, slot)
            order[slot] = (1, i)
        elif r < 0.17:
            ev[i] = (EV_USED if r < 0.145 else EV_UNUSED, (slot + 1) % n_slots)      # sometimes on a chunk that may be free
            if (slot + 1) % n_slots in order:
                order[(slot + 1) % n_slots] = (1 if r < 0.145 else 0, i)
        else:
            ev[i] = (EV_USED if r < 0.75 else EV_UNUSED, slot)
            order[slot] = (1 if r < 0.75 else 0, i)
    return ev


@pytest.mark.parametrize("seed,n_slots,n_events", [(1, 8, 400), (2, 64, 3000), (3, 500, 20000), (4, 3, 200)])
def test_stamp_model_matches_the_reference_lists(seed, n_slots, n_events):
    ref = T.load_ref("pmm")
    if ref is None:
        pytest.skip("oracle/_ref/libref_pmm.so not built (needs /root/reference)")
    orc = T.load_oracle()
    ev = random_events(n_slots, n_events, seed)
    # a victim's chunk is free afterwards: events on it are dropped until it is allocated again, as the caller would
    va, vb = np.zeros(n_events, "<u4"), np.zeros(n_events, "<u4")
    na = ref.ref_pmm_replay(ctypes.c_uint32(n_slots), ctypes.c_uint32(n_events), P(ev), P(va))
    nb = orc.oracle_pmm_replay(ctypes.c_uint32(n_slots), ctypes.c_uint32(n_events), P(ev), P(vb))
    assert na == nb == int((ev[:, 0] == EV_PICK).sum())
    assert np.array_equal(va[:na], vb[:nb])
    assert (va[:na] != 0xFFFFFFFF).sum() > na // 2


def _pick_oracle(orc, stamp, k):
    out = np.zeros(max(k, 1), "<u4")
    n = orc.oracle_pmm_pick(P(stamp), ctypes.c_uint32(len(stamp)), ctypes.c_uint32(k), P(out))
    return out[:n]


def _pick_simt(sim, stamp, k, dest=1):
    rq = np.zeros(max(k, 1), R.MAKE_RESIDENT_REQ_DT)
    sim.simt_pmm_pick.restype = ctypes.c_uint32
    n = sim.simt_pmm_pick(P(stamp), ctypes.c_uint32(len(stamp)), ctypes.c_uint32(k), ctypes.c_uint32(dest), P(rq))
    return rq[:n]


@pytest.mark.parametrize("n_slots,frac,k", [(1000, 0.5, 37), (1000, 0.5, 0), (1000, 0.5, 5000), (300, 1.0, 300), (64, 0.0, 9),
                                             (5000, 0.3, 1), (5000, 0.9, 4499)])
def test_selection_takes_the_k_smallest_stamps(n_slots, frac, k):
    """Unused chunks (bit 63 clear) go before used ones, each in the order of its last move; sequence numbers of
    every magnitude, so that all eight digits of the radix selection matter."""
    rng = np.random.default_rng(n_slots + k)
    seq = rng.permutation(n_slots).astype(np.uint64) * np.uint64(0x0001_0203_0507_0B0D % (1 << 50)) + np.uint64(1)
    stamp = np.where(rng.random(n_slots) < 0.7, seq | np.uint64(USED), seq)
    stamp[rng.random(n_slots) >= frac] = 0
    want = _pick_oracle(T.load_oracle(), stamp, k)
    got = _pick_simt(T.load_simt(), stamp, k, dest=2)
    assert sorted(got["slot"].tolist()) == sorted(want.tolist())
    assert len(got) == min(k, int((stamp != 0).sum()))
    if len(got):
        assert (got["dest_tier"] == 2).all() and (got["cause"] == R.CAUSE_EVICTION).all() and (got["flags"] == R.MR_EVICT_GPU_PAGES).all()
        assert (got["region_first"] == 0).all() and (got["region_outer"] == 512).all()
        # every unused chunk is taken before any used one
        n_unused = int(((stamp != 0) & (stamp < USED)).sum())
        taken_unused = int((stamp[got["slot"]] < USED).sum())
        assert taken_unused == min(n_unused, len(got))


def test_list_moves_of_service_and_make_resident_and_the_chunks_a_batch_needs():
    """A small tiering loop under the emulator: two servicing passes, an eviction wave picked from the order they left,
    a migration to the CPU that empties some blocks (they become 'unused' and go first), another pass.  After every
    step the product's stamps equal the restatement's, and the count of needed chunks equals a direct count."""
    orc, sim = T.load_oracle(), T.load_simt()
    B = 64
    gp = TG.gpu_params(B)
    pk1, tbl, lay = TG.fault_trace(1500, 5, "hotset", 2, total_va_bytes=96 << 21)
    pk2, _, _ = TG.fault_trace(1200, 6, "uniform", 2, total_va_bytes=96 << 21)
    n_slots = lay["blocks_per_space"] * 2
    layout = S.make_layout(lay)
    st = S.random_state(n_slots, 3, populated=0.8)
    st.gpu_resident[:] = 0
    st.gpu_pte0[:] = 0; st.gpu_pte1[:] = 0; st.gpu_pte2[:] = 0          # nothing on the GPU yet: no chunk anywhere
    stamp_o, stamp_p = np.zeros(n_slots, "<u8"), np.zeros(n_slots, "<u8")
    clock = 0

    def needed(k1, stamp):
        nb = len(k1["batch"])
        seen, out = np.zeros((n_slots + 31) // 32, "<u4"), np.zeros(2, "<u4")
        sim.simt_pmm_count_needed(ctypes.c_uint32(nb), ctypes.c_uint32(B), P(k1["groups"]), P(k1["batch"]), P(tbl), P(layout),
                                  P(stamp), P(seen), P(out))
        slots = set()
        for b in range(nb):
            for g in range(int(k1["batch"]["num_block_groups"][b])):
                grp = k1["groups"][b * B + g]
                if grp["flags"] & 1:
                    continue
                sp = int(tbl["va_space"][grp["ctx"]])
                blk = (int(grp["block_start"]) - lay["va_base"]) >> 21
                if 0 <= blk < lay["blocks_per_space"] and sp < 2:
                    slots.add(sp * lay["blocks_per_space"] + blk)
        want = sum(1 for s in slots if stamp[s] == 0)
        assert (int(out[0]), int(out[1])) == (want, int((stamp != 0).sum()))
        return want

    def service(pk, state):
        nonlocal clock
        k1 = F.run_fetch_preprocess_simt(sim, pk, gp, tbl)
        need = needed(k1, stamp_p)
        got = S.run_service_simt(sim, k1, gp, tbl, layout, state)
        nb = len(k1["batch"])
        orc.oracle_pmm_note_service(P(stamp_o), ctypes.c_uint32(n_slots), ctypes.c_uint64(clock), ctypes.c_uint32(nb),
                                    ctypes.c_uint32(B), P(got["batch"]), P(got["service"]))
        sim.simt_pmm_note_service(ctypes.c_uint32(nb), ctypes.c_uint32(B), P(got["batch"]), P(got["service"]), P(stamp_p),
                                  ctypes.c_uint32(n_slots), ctypes.c_uint64(clock))
        clock += nb * B
        assert np.array_equal(stamp_o, stamp_p)
        return got["state"], need

    def make_resident(rq, state):
        nonlocal clock
        res = R.run_simt(sim, rq, None, None, state)
        ptrs = res["state"].ptrs()
        orc.oracle_pmm_note_make_resident(P(stamp_o), ctypes.c_uint32(n_slots), ctypes.c_uint64(clock), ctypes.c_uint32(len(rq)),
                                          P(rq), P(res["results"]), ctypes.byref(ptrs))
        sim.simt_pmm_note_make_resident(ctypes.c_uint32(len(rq)), P(rq), P(res["results"]), P(res["state"].gpu_resident),
                                        P(stamp_p), ctypes.c_uint32(n_slots), ctypes.c_uint64(clock))
        clock += len(rq)
        assert np.array_equal(stamp_o, stamp_p)
        return res

    st, need1 = service(pk1, st)
    in_use = int((stamp_p != 0).sum())
    assert need1 >= in_use > 10 and (stamp_p[stamp_p != 0] >= USED).all()
    # every block with pages on the GPU holds a chunk, and only those
    gpu_any = T.bool_from_masks(st.gpu_resident).any(axis=1)
    assert np.array_equal(gpu_any, stamp_p != 0)
    # a migration to DRAM empties a few blocks without evicting them: they move to the unused list
    some = np.nonzero(gpu_any)[0][:6]
    rq = np.zeros(len(some), R.MAKE_RESIDENT_REQ_DT)
    rq["slot"] = some; rq["region_outer"] = 512; rq["dest_tier"] = R.TIER_DRAM; rq["cause"] = R.CAUSE_API_MIGRATE
    st = make_resident(rq, st)["state"]
    assert ((stamp_p[some] != 0) & (stamp_p[some] < USED)).all()
    # eviction wave: the unused six first, then the least recently serviced
    k = 20
    picked = _pick_simt(sim, stamp_p, k, dest=R.TIER_CXL)
    assert sorted(picked["slot"].tolist()) == sorted(_pick_oracle(orc, stamp_o, k).tolist()) and set(some) <= set(picked["slot"].tolist())
    res = make_resident(picked, st)
    st = res["state"]
    assert (stamp_p[picked["slot"]] == 0).all() and int((stamp_p != 0).sum()) == in_use - k
    assert not T.bool_from_masks(st.gpu_resident)[picked["slot"]].any()
    # the next pass brings some of them back: they need chunks again and rejoin the used list at its tail
    before = stamp_p.copy()
    st, need2 = service(pk2, st)
    assert need2 > 0 and (stamp_p[stamp_p != before] > before.max()).all()
    assert np.array_equal(T.bool_from_masks(st.gpu_resident).any(axis=1), stamp_p != 0)


@pytest.mark.parametrize("seed,n_slots,capacity,dram_cap_blocks,demote", [(1, 300, 120, 40, True), (2, 300, 120, 40, False),
                                                                          (3, 1000, 200, 10, True), (4, 64, 5, 1000, True),
                                                                          (5, 500, 450, 0, True)])
def test_round_planning_kernels_match_the_selection_rule(seed, n_slots, capacity, dram_cap_blocks, demote):
    """The device-side decisions of one round of the tiered loop (pmm_kernels.cuh, under the emulator) against the rule in
    plain numpy: victims = ascending stamp until `overflow` of them are va_blocks the round does not fault on; demotion =
    ascending DRAM stamp among the candidates until the DRAM pages that must go are covered; destination of the victims."""
    rng = np.random.default_rng(seed)
    sim = T.load_simt()
    in_use = rng.random(n_slots) < 0.6
    stamp = np.where(in_use, rng.permutation(n_slots).astype(np.uint64) + 1 + (rng.random(n_slots) < 0.7) * np.uint64(1 << 63), 0).astype("<u8")
    touched_b = rng.random(n_slots) < 0.3
    gres = T.masks_from_bool((rng.random((n_slots, 512)) < 0.3) & in_use[:, None])
    c0 = T.masks_from_bool((rng.random((n_slots, 512)) < 0.4) & (rng.random((n_slots, 1)) < 0.5))
    c1 = T.masks_from_bool(rng.random((n_slots, 512)) < 0.1)
    dram_stamp = (rng.integers(0, 6, n_slots)).astype("<u8")
    touched = np.zeros((n_slots + 31) // 32, "<u4")
    for s in np.nonzero(touched_b)[0]:
        touched[s >> 5] |= np.uint32(1 << (s & 31))
    e_w, e_a, d_w = np.zeros(n_slots, "<u4"), np.zeros(n_slots, "<u4"), np.zeros(n_slots, "<u4")
    d_k = np.zeros(n_slots, "<u8")
    plan = np.zeros(64, "<u4")
    ev, de = np.zeros(n_slots, R.MAKE_RESIDENT_REQ_DT), np.zeros(n_slots, R.MAKE_RESIDENT_REQ_DT)
    dram_cap = dram_cap_blocks * 512
    sim.simt_pmm_plan_round(ctypes.c_uint32(n_slots), P(stamp), P(dram_stamp), P(touched), P(gres), P(c0), P(c1), ctypes.c_uint32(capacity),
                            ctypes.c_uint64(dram_cap), ctypes.c_int(int(demote)), P(e_w), P(e_a), P(d_k), P(d_w), P(plan), P(ev), P(de))
    gp, dp = T.bool_from_masks(gres).sum(axis=1), T.bool_from_masks(c0).sum(axis=1)
    needed, used = int((touched_b & (stamp == 0)).sum()), int((stamp != 0).sum())
    assert (plan[0], plan[1], plan[2], plan[3]) == (needed, used, int(gp.sum()), int(dp.sum()))

    def select(key, weight, target):
        live = np.nonzero(key != 0)[0]
        order = live[np.argsort(key[live], kind="stable")]
        csum = np.cumsum(weight[order].astype(np.int64))
        if target <= 0:
            return order[:0]
        if len(order) == 0 or csum[-1] < target:
            return order
        return live[key[live] <= key[order[int(np.searchsorted(csum, target))]]]

    victims = select(stamp, ((stamp != 0) & ~touched_b).astype(np.int64), used + needed - capacity)
    assert sorted(ev["slot"][:plan[41]].tolist()) == sorted(victims.tolist()) and plan[16 + 2] == len(victims)
    arriving = int(gp[victims].sum())
    demoted = victims[:0]
    if demote and len(victims) and int(dp.sum()) + arriving > dram_cap:
        cand = (dp != 0) & (stamp == 0) & ~touched_b
        demoted = select(np.where(cand, dram_stamp + 1, 0).astype("<u8"), dp, int(dp.sum()) + arriving - dram_cap)
    assert sorted(de["slot"][:plan[40]].tolist()) == sorted(demoted.tolist())
    dest = R.TIER_DRAM if int(dp.sum()) - int(dp[demoted].sum()) + arriving <= dram_cap else R.TIER_CXL
    if len(victims):
        assert (ev["dest_tier"][:plan[41]] == dest).all() and (ev["flags"][:plan[41]] == R.MR_EVICT_GPU_PAGES).all()
    if len(demoted):
        assert (de["dest_tier"][:plan[40]] == R.TIER_CXL).all() and np.array_equal(de["mask_row"][:plan[40]], de["slot"][:plan[40]])
    assert plan[38] == used - len(victims) + needed + int(touched_b[victims].sum())
