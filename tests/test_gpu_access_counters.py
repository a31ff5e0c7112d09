"""GPU tests of the access-counter notification path through the C ABI vs the CPU restatement: fetch / parse / sort /
va_block runs, then expansion and servicing -- records, copy runs and the final va_block state bit-exact."""
import numpy as np
import pytest

import aclib as A
import servicelib as S
import uvmtier
import uvmtier_testlib as T
from test_oracle_access_counters import notifications, state_with_mappings
from uvmtier import tracegen as TG

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("B,n,blocks,cpu_chunk,get", [(256, 20000, 512, 2 << 20, 0), (64, 5000, 64, 4096, 0), (128, 3000, 2000, 2 << 20, 900),
                                                      (1024, 9000, 300, 2 << 20, 0)])
def test_gpu_access_counters_match_oracle(B, n, blocks, cpu_chunk, get):
    orc = T.load_oracle()
    pk, tbl, lay = notifications(n, 5 + B, 2, blocks)
    pk[n // 3, 7] &= 0x7FFFFFFF                                  # an entry that is not ready
    ring_entries = n + 1000
    ring = np.zeros((ring_entries, 8), np.uint32)
    ring[(get + np.arange(n)) % ring_entries] = pk
    put = (get + n) % ring_entries
    gp = TG.gpu_params(B)
    layout = S.make_layout(lay)
    st0 = state_with_mappings(lay["blocks_per_space"] * 2, 31)
    with uvmtier.Context(0) as ctx:
        ctx.fault_buffer_init(gp, tbl)
        ctx.access_counters_init(cpu_chunk_bytes=cpu_chunk)
        ds = ctx.block_state_alloc(layout)
        ds.upload(st0)
        got = ctx.access_counters_service(ring, get, put, layout, ds)
        ds.free()
    want_k = A.run_preprocess(orc, "oracle_access_counters_fetch_preprocess", ring, get, put, B, tbl)
    got["entries"] = got["entries"].view(A.AC_ENTRY_DT)
    A.assert_preprocess_equal(want_k, got, B)
    assert got["get"] == put
    want = A.run_service_oracle(orc, want_k, B, tbl, layout, st0, cpu_chunk_2m=cpu_chunk == 2 << 20)
    S.assert_service_equal(want, got, B, "access counters B=%d" % B)
    assert int(want["records"]["page_fault_count"].sum()) > 100 and want["n_runs"] > 10


def test_promotions_move_va_blocks_to_the_tail_of_the_eviction_order():
    """The records of the access-counter servicing feed uvmtier_pmm_note_service_device like the fault path's: every
    va_block that received pages gets the stamp of its last promoting run."""
    import ctypes
    orc = T.load_oracle()
    B = 128
    pk, tbl, lay = notifications(3000, 77, 2, 128)
    n = len(pk)
    ring = np.zeros((n + 1, 8), np.uint32)
    ring[:n] = pk
    layout = S.make_layout(lay)
    st0 = state_with_mappings(lay["blocks_per_space"] * 2, 5)
    with uvmtier.Context(0) as ctx:
        ctx.fault_buffer_init(TG.gpu_params(B), tbl)
        ctx.access_counters_init()
        ds = ctx.block_state_alloc(layout)
        ds.upload(st0)
        pmm = ctx.pmm_create(layout, 1 << 20)
        got = ctx.access_counters_service(ring, 0, n, layout, ds)
        nb = len(got["batch"])
        d_bat, d_svc = ctx.to_device(got["batch"]), ctx.to_device(got["service"])
        uvmtier._check(uvmtier.lib().uvmtier_pmm_note_service_device(ctx.handle, ctypes.byref(pmm), ctypes.c_uint32(nb), uvmtier.P(d_bat),
                                                                     uvmtier.P(d_svc)), "note_service")
        ctx.synchronize()
        stamps = ctx.pmm_stamps(pmm)
        ctx.free(d_bat); ctx.free(d_svc)
        ctx.pmm_destroy(pmm)
        ds.free()
    want = np.zeros(st0.n, "<u8")
    orc.oracle_pmm_note_service(T.P(want), ctypes.c_uint32(st0.n), ctypes.c_uint64(0), ctypes.c_uint32(nb), ctypes.c_uint32(B),
                                T.P(got["batch"]), T.P(got["service"]))
    assert np.array_equal(want, stamps) and int((stamps != 0).sum()) > 20
    promoted = np.unique(got["service"]["slot"][(got["service"]["page_fault_count"] > 0)])
    assert set(np.nonzero(stamps)[0].tolist()) <= set(promoted.tolist())
