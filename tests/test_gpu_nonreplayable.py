"""GPU parity of non-replayable fault servicing through the C ABI (uvmtier_service_non_replayable_faults_device) against
the CPU restatement, whose block function is pinned by the reference's own code (tests/test_oracle_nonreplayable.py)."""
import ctypes

import numpy as np
import pytest

import nonreplaylib as N
import servicelib as S
import uvmtier
import uvmtier_testlib as T
from uvmtier import tracegen as TG

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,blocks,n_spaces,B,seed", [(20000, 512, 4, 256, 1), (6000, 6, 1, 64, 2), (3000, 2000, 2, 100, 3), (1, 4, 1, 256, 4)])
def test_gpu_matches_restatement(ctx, n, blocks, n_spaces, B, seed):
    tbl = TG.instance_ptr_table(n_spaces)
    lay = dict(va_base=TG.VA_BASE, blocks_per_space=blocks, n_spaces=n_spaces, gpu_id=1)
    gp = TG.gpu_params(B)
    st0 = S.random_state(blocks * n_spaces, seed, populated=0.6)
    pk = N.shadow_buffer(n, seed, lay, tbl)
    layout = S.make_layout(lay)
    want = N.run_checker(T.load_oracle(), "oracle", pk, gp, tbl, layout, st0, run_cap=1 << 20)
    ctx.fault_buffer_init(gp, tbl)
    ds = ctx.block_state_alloc(layout)
    ds.upload(st0)
    try:
        got = ctx.service_non_replayable_faults(pk, layout, ds, run_cap=1 << 20)
    finally:
        ds.free()
    N.assert_equal(want, got, "n %d blocks %d B %d" % (n, blocks, B))
    S.check_service_invariants(got, st0)


def test_gpu_bad_arguments(ctx):
    lib = uvmtier.lib()
    tbl = TG.instance_ptr_table(1)
    ctx.fault_buffer_init(TG.gpu_params(256), tbl)
    layout = S.make_layout(dict(va_base=TG.VA_BASE, blocks_per_space=4, n_spaces=1, gpu_id=1))
    ds = ctx.block_state_alloc(layout)
    one = ctx.alloc(4096)
    try:
        f = lib.uvmtier_service_non_replayable_faults_device
        assert f(None, uvmtier.P(one), ctypes.c_uint32(1), uvmtier.P(layout), ctypes.byref(ds.ptrs), None, uvmtier.P(one), uvmtier.P(one), uvmtier.P(one), ctypes.c_uint32(8), uvmtier.P(one)) == 0x1F
        assert f(ctx.handle, None, ctypes.c_uint32(1), uvmtier.P(layout), ctypes.byref(ds.ptrs), None, uvmtier.P(one), uvmtier.P(one), uvmtier.P(one), ctypes.c_uint32(8), uvmtier.P(one)) == 0x1F
        assert f(ctx.handle, uvmtier.P(one), ctypes.c_uint32(1), uvmtier.P(layout), ctypes.byref(ds.ptrs), None, None, uvmtier.P(one), uvmtier.P(one), ctypes.c_uint32(8), uvmtier.P(one)) == 0x1F
        assert f(ctx.handle, uvmtier.P(one), ctypes.c_uint32(0), uvmtier.P(layout), ctypes.byref(ds.ptrs), None, None, None, None, ctypes.c_uint32(0), None) == 0
    finally:
        ctx.free(one)
        ds.free()
