"""GPU parity tests of the servicing pass (stage C/D/E) through the C ABI vs the oracle: flags, counters, records,
copy runs and the final va_block state, bit-exact; size-independent invariants at the BASELINE size."""
import numpy as np
import pytest

import faultlib as F
import servicelib as S
import uvmtier_testlib as T
from uvmtier import tracegen as TG

pytestmark = pytest.mark.gpu


def gpu_pipeline(ctx, pk, gp, tbl, lay, st0, run_cap=1 << 21):
    """fetch+preprocess then service on the GPU; returns (k1 in checker view, service result)."""
    ctx.fault_buffer_init(gp, tbl)
    n = len(pk)
    ring = np.zeros((n + 1, 8), np.uint32)
    ring[:n] = pk
    r = ctx.fetch_preprocess_fault_batches(ring, 0, n)
    B = int(gp["max_batch_size"][0])
    k1 = F.product_to_checker_view(n, B, tbl, r["disp"], r["ordered"], r["groups"], r["batch"], r["entries"], ring[:n].copy())
    # pad the per-entry arrays to whole batches (the device arrays are indexed b*B + s)
    nb = len(r["batch"])
    pad = nb * B - n
    if pad:
        r["ordered"] = np.concatenate([r["ordered"], np.zeros(pad, r["ordered"].dtype)])
        r["groups"] = np.concatenate([r["groups"], np.zeros(pad, r["groups"].dtype)])
        k1["ordered_rec"] = r["ordered"]; k1["groups"] = r["groups"]
    layout = S.make_layout(lay)
    ds = ctx.block_state_alloc(layout)
    try:
        ds.upload(st0)
        got = ctx.service_fault_batches(r, layout, ds, run_cap)
    finally:
        ds.free()
    return k1, layout, got


@pytest.mark.parametrize("dist", ["uniform", "stream", "hotset", "adversarial"])
@pytest.mark.parametrize("B,n_spaces,va_blocks", [(256, 4, 4096), (64, 2, 128), (1024, 1, 512), (4096, 2, 1024)])
def test_gpu_service_matches_oracle(dist, B, n_spaces, va_blocks):
    import uvmtier
    orc = T.load_oracle()
    pk, tbl, lay = TG.fault_trace(30000, 17, dist, n_spaces, total_va_bytes=va_blocks << 21)
    gp = TG.gpu_params(B)
    st0 = S.random_state(lay["blocks_per_space"] * n_spaces, 5)
    with uvmtier.Context(0) as ctx:
        k1, layout, got = gpu_pipeline(ctx, pk, gp, tbl, lay, st0)
    want = S.run_service_oracle(orc, k1, gp, tbl, layout, st0, run_cap=1 << 21)
    S.assert_service_equal(want, got, B, "%s B=%d" % (dist, B))
    S.check_service_invariants(got, st0)


@pytest.mark.parametrize("thr", [1, 100])
def test_gpu_service_prefetch_tunables_and_edge_cases(thr):
    import uvmtier
    orc = T.load_oracle()
    pk, tbl, lay = TG.fault_trace(5000, 8, "uniform", 2, total_va_bytes=64 << 21)
    out = np.arange(5000) % 5 == 2
    pk[out, 3] |= 0x40                        # outside every va_range
    fatal = np.arange(5000) % 11 == 3
    pk[fatal, 7] = (pk[fatal, 7] & ~np.uint32(0x1F)) | TG.HW_FAULT_TYPE_PDE_SIZE
    pk[fatal, 3] |= 0x100
    pk[700, 1] ^= 0x5                         # stale channel: that batch is not serviced
    gp = TG.gpu_params(128)
    st0 = S.random_state(lay["blocks_per_space"] * 2, 6)
    with uvmtier.Context(0, uvm_perf_prefetch_threshold=thr, uvm_perf_prefetch_min_faults=2) as ctx:
        k1, layout, got = gpu_pipeline(ctx, pk, gp, tbl, lay, st0)
    want = S.run_service_oracle(orc, k1, gp, tbl, layout, st0, pf=(1, thr, 2), run_cap=1 << 21)
    S.assert_service_equal(want, got, 128, "thr %d" % thr)


def test_gpu_service_run_buffer_too_small_is_reported(ctx):
    pk, tbl, lay = TG.fault_trace(4000, 3, "uniform", 1, total_va_bytes=64 << 21)
    gp = TG.gpu_params(256)
    st0 = S.random_state(lay["blocks_per_space"], 1)
    k1, layout, got = gpu_pipeline(ctx, pk, gp, tbl, lay, st0, run_cap=16)
    assert got["n_runs"] > 16                  # the call says how many runs it needs


def test_gpu_service_baseline_size(ctx):
    """Config B: 1M entries over 64 GB of VA (32768 va_blocks), 4 VA spaces, empty initial residency on the GPU,
    everything on the CPU nodes.  Full compare with the oracle + invariants + conservation of pages."""
    orc = T.load_oracle()
    n = 1 << 20
    pk, tbl, lay = TG.fault_trace(n, 1, "uniform", 4)
    gp = TG.gpu_params(256)
    n_slots = lay["blocks_per_space"] * 4
    st0 = S.random_state(n_slots, 7, populated=0.9, ro_frac=0.02, none_frac=0.0)
    k1, layout, got = gpu_pipeline(ctx, pk, gp, tbl, lay, st0, run_cap=1 << 23)
    want = S.run_service_oracle(orc, k1, gp, tbl, layout, st0, run_cap=1 << 23)
    S.assert_service_equal(want, got, 256, "1M")
    S.check_service_invariants(got, st0)
    # conservation: pages that left the CPU nodes are exactly the pages the copy runs move
    rec = got["service"][got["service"]["page_fault_count"] > 0]
    moved = int(rec["migrated"].sum())
    before = sum(int(T.bool_from_masks(m).sum()) for m in (st0.cpu_resident0, st0.cpu_resident1))
    after = sum(int(T.bool_from_masks(m).sum()) for m in (got["state"].cpu_resident0, got["state"].cpu_resident1))
    assert before - after == moved
    runs = got["runs"][:got["n_runs"]]
    cpu_runs = runs[((runs["bits"] >> 10) & 7) != 7]
    assert int((cpu_runs["bits"] & 0x3FF).sum()) == moved


@pytest.mark.parametrize("dist,chunk_batches,get,repeat", [("hotset", 0, 0, 1), ("uniform", 5, 0, 1), ("stream", 1, 0, 1),
                                                           ("hotset", 7, 700, 1), ("hotset", 0, 0, 4), ("uniform", 9, 700, 3)])
def test_streamed_fetch_preprocess_service_equals_the_one_pass_form(dist, chunk_batches, get, repeat):
    """uvmtier_service_replayable_faults_streamed: the ring stays in pinned host memory; every chunk is preprocessed
    and serviced while the next one is in flight.  Same flags, batch counters, records (with trace-wide batch
    numbers), copy runs and final state as servicing the whole trace in one pass -- for chunks that do not divide
    the trace, one batch per chunk, and a ring segment that wraps; many va_blocks are hit from several chunks.
    repeat > 1: the call is made again with the same buffers on a restored state table, which the library serves
    from a CUDA graph captured on the second call; the last call's results are the ones compared."""
    import uvmtier
    n, B = 20000, 64
    pk, tbl, lay = TG.fault_trace(n, 23, dist, 2, total_va_bytes=256 << 21)
    gp = TG.gpu_params(B)
    st0 = S.random_state(lay["blocks_per_space"] * 2, 8)
    layout = S.make_layout(lay)
    ring_entries = n + 1000
    ring = np.zeros((ring_entries, 8), np.uint32)
    ring[(get + np.arange(n)) % ring_entries] = pk
    put = int((get + n) % ring_entries)
    with uvmtier.Context(0) as ctx:
        ctx.fault_buffer_init(gp, tbl)
        ds = ctx.block_state_alloc(layout)
        ds.upload(st0)
        launches = []

        def restore():
            ds.upload(st0)
            launches.append(uvmtier.lib().uvmtier_kernel_launch_count())

        uvmtier.lib().uvmtier_kernel_launch_count.restype = __import__("ctypes").c_uint64
        got = ctx.service_replayable_faults_streamed(ring, get, put, layout, ds, run_cap=1 << 20, chunk_batches=chunk_batches,
                                                     repeat=repeat, before_each=restore)
        launches.append(uvmtier.lib().uvmtier_kernel_launch_count())
        per_call = np.diff(launches)
        assert (per_call == per_call[0]).all() and per_call[0] > 0     # replays count the kernels of the graph
        ds.free()
        k1, _, one = gpu_pipeline(ctx, pk, gp, tbl, lay, st0, run_cap=1 << 20)
    assert got["get"] == put
    # the copy-back through pinned memory carries what the device holds
    assert np.array_equal(got["flags"], got["device_copies"]["flags"]) and np.array_equal(got["batch"], got["device_copies"]["batch"])
    assert got["n_runs"] == int(got["device_copies"]["counters"][0])
    want = S.run_service_oracle(T.load_oracle(), k1, gp, tbl, layout, st0, run_cap=1 << 20)
    S.assert_service_equal(want, got, B, "streamed %s" % dist)
    S.assert_service_equal(want, one, B, "one pass %s" % dist)
