import sys, time, ctypes
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/open-gpu-kernel-modules_b200")
import numpy as np, uvmtier
from uvmtier import tracegen as TG
lib = uvmtier.lib()
n = 1 << 20
pk, tbl, lay = TG.fault_trace(n, 1, "uniform", 4)
for B in (256, 1024, 4096, 16384):
    with uvmtier.Context(0) as ctx:
        gp = TG.gpu_params(B)
        ctx.fault_buffer_init(gp, tbl)
        ring = np.zeros((n + 1, 8), np.uint32); ring[:n] = pk
        d_ring = ctx.to_device(ring)
        nb = (n + B - 1) // B
        d_disp, d_ord, d_grp, d_bat = ctx.alloc(n * 4), ctx.alloc(nb * B * 16), ctx.alloc(nb * B * 16), ctx.alloc(nb * 32)
        new_get = ctypes.c_uint32(0)
        ts = []
        for it in range(5):
            ctx.synchronize(); t0 = time.perf_counter()
            st = lib.uvmtier_fetch_preprocess_fault_batches_device(ctx.handle, uvmtier.P(d_ring), ctypes.c_uint32(n + 1), ctypes.c_uint32(0), ctypes.c_uint32(n), ctypes.c_int(0), uvmtier.P(d_disp), uvmtier.P(d_ord), uvmtier.P(d_grp), uvmtier.P(d_bat), None, ctypes.byref(new_get))
            assert st == 0, hex(st)
            ctx.synchronize(); ts.append(time.perf_counter() - t0)
        print("B", B, "batches", nb, "ms", min(ts) * 1e3)
