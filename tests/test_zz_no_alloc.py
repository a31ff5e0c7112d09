"""The servicing entry point for callers that may not allocate (uvmtier_service_fault_batches_noalloc_device).

Some users of this library run where asking the CUDA allocator for working space is not an option: the thread
that services GPU faults is exactly the code that has to make progress when device memory is full, and some hosts
fix their memory budget at start-up.  For them the library keeps the servicing pass's working space in a fixed
array inside the library image.  This test drives that entry point through the C ABI, twice in a row on different
inputs (the array is shared and keeps the previous call's contents, so the second call checks that stale contents
do no harm), and requires bit-identical results to the regular entry point; it also checks the size ceiling.

Kept in a file that sorts last so that a run stopping at the first failure has already run everything else.
"""
import ctypes

import numpy as np
import pytest

import servicelib as S
import uvmtier
from uvmtier import P, blockstate as BS, layouts as L, tracegen as TG

pytestmark = pytest.mark.gpu


def _preprocess(ctx, pk, gp, tbl):
    ctx.fault_buffer_init(gp, tbl)
    n = len(pk)
    ring = np.zeros((n + 1, 8), np.uint32)
    ring[:n] = pk
    r = ctx.fetch_preprocess_fault_batches(ring, 0, n)
    B = int(gp["max_batch_size"][0])
    pad = len(r["batch"]) * B - n
    r["ordered"] = np.concatenate([r["ordered"], np.zeros(pad, r["ordered"].dtype)])
    r["groups"] = np.concatenate([r["groups"], np.zeros(pad, r["groups"].dtype)])
    return r


def _service(ctx, fn_name, r, layout, st0, run_cap):
    """Runs one of the two servicing entry points on device copies of r; returns everything it produced."""
    lib = uvmtier.lib()
    nb, total = len(r["batch"]), len(r["ordered"])
    ds = ctx.block_state_alloc(layout)
    ds.upload(st0)
    d_ord, d_grp, d_bat = ctx.to_device(r["ordered"]), ctx.to_device(r["groups"]), ctx.to_device(r["batch"])
    d_flags, d_svc, d_runs, d_cnt = ctx.alloc(total * 2), ctx.alloc(total * 96), ctx.alloc(run_cap * 8), ctx.alloc(8)
    try:
        uvmtier._check(lib.uvmtier_memset_device(ctx.handle, ctypes.c_void_p(d_svc), ctypes.c_int(0), ctypes.c_size_t(total * 96)), "memset")
        uvmtier._check(lib.uvmtier_memset_device(ctx.handle, ctypes.c_void_p(d_runs), ctypes.c_int(0), ctypes.c_size_t(run_cap * 8)), "memset")
        st = getattr(lib, fn_name)(ctx.handle, ctypes.c_uint32(nb), P(d_ord), P(d_grp), P(d_bat), P(layout), ctypes.byref(ds.ptrs),
                                   P(d_flags), P(d_svc), P(d_runs), ctypes.c_uint32(run_cap), P(d_cnt))
        uvmtier._check(st, fn_name)
        ctx.synchronize()
        cnt = ctx.from_device(d_cnt, 2, "<u4")
        return dict(flags=ctx.from_device(d_flags, total, "<u2"), service=ctx.from_device(d_svc, total, BS.SERVICE_GROUP_DT),
                    runs=ctx.from_device(d_runs, run_cap, BS.COPY_RUN_DT), n_runs=int(cnt[0]),
                    batch=ctx.from_device(d_bat, nb, L.BATCH_DT), state=ds.download())
    finally:
        for p in (d_ord, d_grp, d_bat, d_flags, d_svc, d_runs, d_cnt):
            ctx.free(p)
        ds.free()


def _same(a, b, what):
    assert np.array_equal(a["flags"], b["flags"]), what + ": entry flags"
    assert np.array_equal(a["batch"], b["batch"]), what + ": batch counters"
    a["state"].assert_equal(b["state"], what)
    assert a["n_runs"] == b["n_runs"], what + ": number of copy runs"
    ra, rb = a["service"], b["service"]
    for f in S.RECORD_FIELDS:
        assert np.array_equal(ra[f], rb[f]), "%s: record.%s" % (what, f)
    # copy runs: same runs per record (where each record's runs sit in the array is decided by an atomic counter)
    has = np.nonzero(ra["run_count"] > 0)[0]
    if len(has):
        within = np.concatenate([np.arange(c) for c in ra["run_count"][has]])
        ia = np.repeat(ra["run_offset"][has].astype(np.int64), ra["run_count"][has]) + within
        ib = np.repeat(rb["run_offset"][has].astype(np.int64), rb["run_count"][has]) + within
        assert np.array_equal(a["runs"][ia], b["runs"][ib]), what + ": copy runs"


def test_no_alloc_service_matches_the_regular_entry_point_twice_in_a_row():
    run_cap = 1 << 20
    with uvmtier.Context(0) as ctx:
        # two different problems back to back: the second call finds the first call's leftovers in the shared array
        for call, (dist, n, B, n_spaces, blocks, seed) in enumerate((("uniform", 12000, 256, 4, 4096, 11),
                                                                     ("hotset", 9000, 64, 2, 256, 12))):
            pk, tbl, lay = TG.fault_trace(n, seed, dist, n_spaces, total_va_bytes=blocks << 21)
            gp = TG.gpu_params(B)
            r = _preprocess(ctx, pk, gp, tbl)
            layout = S.make_layout(lay)
            st0 = S.random_state(lay["blocks_per_space"] * n_spaces, seed)
            want = _service(ctx, "uvmtier_service_fault_batches_device", r, layout, st0, run_cap)
            got = _service(ctx, "uvmtier_service_fault_batches_noalloc_device", r, layout, st0, run_cap)
            assert want["n_runs"] <= run_cap and int((want["flags"] & 1).sum()) > 100
            _same(want, got, "call %d (%s)" % (call, dist))


def test_no_alloc_service_as_the_first_servicing_call_of_a_context():
    """A caller that may not allocate has, by definition, not run the allocating entry point first.  On a fresh
    context the no-alloc call must find everything it needs already in place (uvmtier_init sets up the side streams,
    their events and the fixed array's address), and give the regular entry point's results."""
    run_cap = 1 << 18
    pk, tbl, lay = TG.fault_trace(6000, 21, "uniform", 2, total_va_bytes=512 << 21)
    gp = TG.gpu_params(128)
    layout = S.make_layout(lay)
    st0 = S.random_state(lay["blocks_per_space"] * 2, 21)
    with uvmtier.Context(0) as fresh:
        r = _preprocess(fresh, pk, gp, tbl)
        got = _service(fresh, "uvmtier_service_fault_batches_noalloc_device", r, layout, st0, run_cap)
    with uvmtier.Context(0) as other:
        r2 = _preprocess(other, pk, gp, tbl)
        want = _service(other, "uvmtier_service_fault_batches_device", r2, layout, st0, run_cap)
    assert int((want["flags"] & 1).sum()) > 100
    _same(want, got, "no-alloc first")


def test_no_alloc_service_rejects_sizes_beyond_its_fixed_array():
    lib = uvmtier.lib()
    with uvmtier.Context(0) as ctx:
        pk, tbl, lay = TG.fault_trace(512, 3, "uniform", 1, total_va_bytes=64 << 21)
        gp = TG.gpu_params(256)
        r = _preprocess(ctx, pk, gp, tbl)
        big = S.make_layout(dict(va_base=lay["va_base"], blocks_per_space=65537, n_spaces=1))     # one slot too many
        dummy = BS.BlockStatePtrs()
        one = ctx.alloc(64)
        try:
            st = lib.uvmtier_service_fault_batches_noalloc_device(ctx.handle, ctypes.c_uint32(2), P(one), P(one), P(one), P(big),
                                                                  ctypes.byref(dummy), P(one), P(one), P(one), ctypes.c_uint32(1), P(one))
            assert st == 0x1F                                    # NV_ERR_INVALID_ARGUMENT, nothing was launched
            assert b"exceed the fixed scratch" in lib.uvmtier_last_error()
            ok = S.make_layout(lay)
            st = lib.uvmtier_service_fault_batches_noalloc_device(ctx.handle, ctypes.c_uint32(4097), P(one), P(one), P(one), P(ok),
                                                                  ctypes.byref(dummy), P(one), P(one), P(one), ctypes.c_uint32(1), P(one))
            assert st == 0x1F                                    # too many batches / entries
        finally:
            ctx.free(one)
