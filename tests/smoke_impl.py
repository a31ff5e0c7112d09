"""smoke(): one small call per hot-path leg on cuda:0, compared with the CPU oracle."""
import ctypes

import numpy as np

import faultlib as F
import servicelib as S
import uvmtier
from uvmtier import TIER_DRAM, TIER_HBM, tracegen as TG
from uvmtier_testlib import gen_prefetch_blocks, load_oracle, run_prefetch_get_hint


def run():
    orc = load_oracle()
    with uvmtier.Context(0) as ctx:
        before = uvmtier.kernel_launch_count()

        # (2) prefetch hints, standalone
        inp = gen_prefetch_blocks(256, 3, full_blocks=False, aligned=False, thrashing=True)
        got = ctx.perf_prefetch_get_hint_va_block(inp["blocks"], inp["resident"], inp["faulted"], inp["thrashing"],
                                                  inp["cpu_read"], inp["info"])
        want = run_prefetch_get_hint(orc, "oracle", inp)
        for g, w, name in zip(got, want, ("prefetch mask", "hint residency", "prefetch_info")):
            assert np.array_equal(g, w), "smoke: %s differs from the oracle" % name

        # (1) fault batches: decode + coalesce + sort + va_block runs, then the per-block service pass
        n_blocks = 64
        pk, tbl, lay = TG.fault_trace(2000, 5, "stream", 2, total_va_bytes=n_blocks << 21)
        gp = TG.gpu_params(256)
        ctx.fault_buffer_init(gp, tbl)
        ring = np.zeros((len(pk) + 1, 8), np.uint32)
        ring[:len(pk)] = pk
        r = ctx.fetch_preprocess_fault_batches(ring, 0, len(pk))
        k1 = F.product_to_checker_view(len(pk), 256, tbl, r["disp"], r["ordered"], r["groups"], r["batch"], r["entries"],
                                       ring[:len(pk)].copy())
        F.assert_fetch_equal(F.run_fetch_preprocess(orc, "oracle", pk, gp, tbl), k1, "smoke fetch/preprocess")
        nb = len(r["batch"])
        pad = nb * 256 - len(pk)
        r["ordered"] = np.concatenate([r["ordered"], np.zeros(pad, r["ordered"].dtype)])
        r["groups"] = np.concatenate([r["groups"], np.zeros(pad, r["groups"].dtype)])
        k1["ordered_rec"], k1["groups"] = r["ordered"], r["groups"]
        layout = S.make_layout(lay)
        st0 = S.random_state(n_blocks, 4)
        ds = ctx.block_state_alloc(layout)
        ds.upload(st0)
        svc = ctx.service_fault_batches(r, layout, ds)
        S.assert_service_equal(S.run_service_oracle(orc, k1, gp, tbl, layout, st0), svc, 256, "smoke service")

        # (1c) make_resident as a call of its own: migrations and evictions on the state the faults left
        import residentlib as R
        st1 = svc["state"]
        rq, pm, pf = R.random_requests(st1.n, 6)
        mr = ctx.make_resident(rq, pm, pf, layout, ds)
        R.assert_equal(R.run_oracle(orc, rq, pm, pf, st1), mr, "smoke make_resident")
        ds.free()

        # (1g) non-replayable faults: parse + translate + one-by-one servicing of a shadow buffer
        import nonreplaylib as NR
        ctx.fault_buffer_init(gp, tbl)
        nrpk = NR.shadow_buffer(1500, 8, lay, tbl)
        nrds = ctx.block_state_alloc(layout)
        nrds.upload(st0)
        nrgot = ctx.service_non_replayable_faults(nrpk, layout, nrds)
        nrds.free()
        NR.assert_equal(NR.run_checker(orc, "oracle", nrpk, gp, tbl, layout, st0), nrgot, "smoke non-replayable faults")

        # (1e) access-counter notifications: fetch + sort + expansion + servicing on the same table shape
        import aclib as A
        from test_oracle_access_counters import notifications, state_with_mappings
        npk, ntbl, nlay = notifications(2000, 3, 2, 64)
        nring = np.zeros((len(npk) + 1, 8), np.uint32)
        nring[:len(npk)] = npk
        ctx.fault_buffer_init(gp, ntbl)
        ctx.access_counters_init()
        nlayout = S.make_layout(nlay)
        nst0 = state_with_mappings(128, 4)
        nds = ctx.block_state_alloc(nlayout)
        nds.upload(nst0)
        got = ctx.access_counters_service(nring, 0, len(npk), nlayout, nds)
        nds.free()
        nk = A.run_preprocess(orc, "oracle_access_counters_fetch_preprocess", nring, 0, len(npk), 256, ntbl)
        S.assert_service_equal(A.run_service_oracle(orc, nk, 256, ntbl, nlayout, nst0), got, 256, "smoke access counters")

        # (2c) thrashing detector: two processors fighting over a few va_blocks
        import thrashlib as TH
        tev = TH.by_slot(TH.ping_pong_events(16, 3000, 3))
        tt = ctx.thrashing_create(S.make_layout(dict(va_base=TG.VA_BASE, blocks_per_space=16, n_spaces=1)))
        tcnt = ctx.thrashing_note_events(tt, tev)
        ttab = ctx.thrashing_tables(tt)
        ctx.thrashing_destroy(tt)
        twant = TH.run_oracle(orc, tev, 16)
        assert np.array_equal(ttab["thrashing_pages"], twant.thrashing_pages) and np.array_equal(ttab["page_word"], twant.page_word)
        assert tcnt[0] == twant.counters[0] and tcnt[0] > 0, "smoke: thrashing detections differ"

        # (3) migration: 4 KB pages DRAM -> HBM in random order, byte exact
        nbytes = 8 << 20
        info = {t: ctx.tier_create(t, nbytes) for t in (TIER_HBM, TIER_DRAM)}
        src = info[TIER_DRAM].host_view()
        src[:] = TG.rng_stream(9, nbytes // 4).view(np.uint8)
        perm = np.random.default_rng(1).permutation(nbytes // 4096)
        off = np.arange(nbytes // 4096, dtype=np.uint64) * np.uint64(4096)
        ctx.copy_page_list(TIER_HBM, TIER_DRAM, perm.astype(np.uint64) * np.uint64(4096), off, 4096)
        hbm = ctx.from_device(info[TIER_HBM].device_ptr, nbytes, np.uint8).reshape(-1, 4096)
        assert np.array_equal(hbm[perm], src.reshape(-1, 4096)), "smoke: migrated pages differ"

        launches = uvmtier.kernel_launch_count() - before
        assert launches >= 10, "smoke: CUDA kernels were not launched"
    print("smoke ok: %s, %d kernel launches (prefetch, fault batches, service, make_resident, non-replayable faults, access counters, thrashing, migration)" %
          (uvmtier.lib().uvmtier_version().decode(), launches))
