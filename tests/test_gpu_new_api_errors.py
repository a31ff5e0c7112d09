"""Error behaviour of the entry points of sections 1d / 1e / 3b: the reference's convention -- bad arguments and
wrong state come back as NV_STATUS codes, nothing is launched."""
import ctypes

import numpy as np
import pytest

import servicelib as S
import uvmtier
from uvmtier import tracegen as TG
from uvmtier.blockstate import MAKE_RESIDENT_REQ_DT, MR_READ_DUPLICATE

pytestmark = pytest.mark.gpu
OK, INVALID_ARGUMENT, INVALID_STATE = 0, 0x1F, 0x40


def test_access_counter_and_pmm_entry_points_reject_bad_calls():
    lib = uvmtier.lib()
    lay = dict(va_base=TG.VA_BASE, blocks_per_space=16, n_spaces=1)
    layout = S.make_layout(lay)
    with uvmtier.Context(0) as ctx:
        h = ctx.handle
        launches = uvmtier.kernel_launch_count()
        # before uvmtier_fault_buffer_init there is no table and no batch size
        assert lib.uvmtier_access_counters_init(h, None) == INVALID_STATE
        d = ctx.alloc(4096)
        out_get = ctypes.c_uint32(7)
        assert lib.uvmtier_access_counters_fetch_preprocess_device(h, uvmtier.P(d), ctypes.c_uint32(64), ctypes.c_uint32(0), ctypes.c_uint32(8),
                                                                   ctypes.c_int(1), uvmtier.P(d), uvmtier.P(d), uvmtier.P(d), None,
                                                                   ctypes.byref(out_get)) == INVALID_STATE
        ctx.fault_buffer_init(TG.gpu_params(64), TG.instance_ptr_table(1))
        assert lib.uvmtier_access_counters_init(h, None) == OK
        for gran, chunk in ((3 << 20, 2 << 20), (2 << 20, 8192)):
            cfg = (ctypes.c_uint64 * 2)(gran, chunk)
            assert lib.uvmtier_access_counters_init(h, cfg) == INVALID_ARGUMENT
        # GET / PUT outside the ring, missing outputs; an empty range is fine and reports the new GET
        f = lib.uvmtier_access_counters_fetch_preprocess_device
        assert f(h, uvmtier.P(d), ctypes.c_uint32(64), ctypes.c_uint32(64), ctypes.c_uint32(0), ctypes.c_int(1), uvmtier.P(d), uvmtier.P(d),
                 uvmtier.P(d), None, None) == INVALID_ARGUMENT
        assert f(h, uvmtier.P(d), ctypes.c_uint32(64), ctypes.c_uint32(0), ctypes.c_uint32(8), ctypes.c_int(1), None, uvmtier.P(d),
                 uvmtier.P(d), None, None) == INVALID_ARGUMENT
        assert f(h, uvmtier.P(d), ctypes.c_uint32(64), ctypes.c_uint32(5), ctypes.c_uint32(5), ctypes.c_int(1), None, None, None, None,
                 ctypes.byref(out_get)) == OK and out_get.value == 5
        # the eviction order: layout / table mismatch, a destination that is not a CPU tier
        ds = ctx.block_state_alloc(layout)
        pmm = ctx.pmm_create(layout, 4)
        cnt = ctx.alloc(64)
        assert lib.uvmtier_pmm_pick_chunks_to_evict_device(h, ctypes.byref(pmm), ctypes.c_uint32(2), ctypes.c_uint32(0), uvmtier.P(d),
                                                           uvmtier.P(cnt)) == INVALID_ARGUMENT
        assert lib.uvmtier_pmm_pick_chunks_to_evict_device(h, ctypes.byref(pmm), ctypes.c_uint32(2), ctypes.c_uint32(1), None,
                                                           uvmtier.P(cnt)) == INVALID_ARGUMENT
        other = S.make_layout(dict(va_base=TG.VA_BASE, blocks_per_space=32, n_spaces=1))
        assert lib.uvmtier_pmm_count_needed_device(h, ctypes.byref(pmm), ctypes.c_uint32(0), None, None, uvmtier.P(other),
                                                   uvmtier.P(cnt)) == INVALID_ARGUMENT
        stats = uvmtier.TieringStats()
        assert lib.uvmtier_service_fault_batches_tiered_device(h, ctypes.byref(pmm), ctypes.c_uint32(1), uvmtier.P(d), uvmtier.P(d), uvmtier.P(d),
                                                               uvmtier.P(layout), ctypes.byref(ds.ptrs), uvmtier.P(d), uvmtier.P(d),
                                                               ctypes.c_uint32(0), ctypes.c_uint64(0), ctypes.c_int(0),
                                                               ctypes.byref(stats)) == INVALID_ARGUMENT      # chunk_batches = 0
        assert lib.uvmtier_service_fault_batches_tiered_device(h, ctypes.byref(pmm), ctypes.c_uint32(1), uvmtier.P(d), uvmtier.P(d), uvmtier.P(d),
                                                               uvmtier.P(layout), ctypes.byref(ds.ptrs), uvmtier.P(d), uvmtier.P(d),
                                                               ctypes.c_uint32(1), ctypes.c_uint64(0), ctypes.c_int(1),
                                                               ctypes.byref(stats)) == INVALID_STATE         # copies without tiers
        # read duplication needs the mask
        rq = np.zeros(1, MAKE_RESIDENT_REQ_DT)
        rq["region_outer"] = 512; rq["flags"] = MR_READ_DUPLICATE
        res = ctx.make_resident(rq, None, None, layout, ds)
        assert res["results"]["slot"][0] == 0xFFFFFFFF
        assert uvmtier.kernel_launch_count() - launches <= 2                 # nothing but that make_resident call launched kernels
        ctx.pmm_destroy(pmm)
        ds.free()
        ctx.free(d); ctx.free(cnt)
