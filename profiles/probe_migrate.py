"""Bandwidth probe of the migration leg on one B200: copy-engine peak (cudaMemcpyAsync) vs the SM-driven TMA
pipeline vs plain ld/st, per tier pair, direction and page size.  Writes one JSON line per measurement.
    python profiles/probe_migrate.py [GiB]   (run under gpurun)
"""
import ctypes
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "open-gpu-kernel-modules_b200"))
import uvmtier  # noqa: E402
from uvmtier import P, TIER_CXL, TIER_DRAM, TIER_HBM  # noqa: E402

NAMES = {TIER_HBM: "HBM", TIER_DRAM: "DRAM", TIER_CXL: "CXL"}


def main():
    gib = float(sys.argv[1]) if len(sys.argv) > 1 else 1.0
    nbytes = int(gib * (1 << 30)) // (2 << 20) * (2 << 20)
    lib = uvmtier.lib()
    with uvmtier.Context(0) as c:
        info = {t: c.tier_create(t, nbytes) for t in (TIER_HBM, TIER_DRAM, TIER_CXL)}
        n, dram, cxl = uvmtier.numa_probe()
        print(json.dumps({"numa_nodes": n, "dram_node": dram, "cxl_node": cxl,
                          "cxl_emulated": bool(info[TIER_CXL].flags & 1), "bytes": nbytes}))
        for t in (TIER_DRAM, TIER_CXL):
            info[t].host_view()[:] = 0x5A
        ms = ctypes.c_float(0)
        for dst, src in ((TIER_HBM, TIER_DRAM), (TIER_HBM, TIER_CXL), (TIER_DRAM, TIER_HBM), (TIER_CXL, TIER_HBM),
                         (TIER_CXL, TIER_DRAM), (TIER_HBM, TIER_HBM)):
            peak = c.tier_memcpy_peak_gbs(dst, src, nbytes if dst != src else nbytes // 2, 5) if dst != src else None
            for page_size in (4096, 65536, 2 << 20):
                n_pages = nbytes // page_size
                if dst == src:
                    n_pages //= 2
                rng = np.random.default_rng(1)
                for order in ("sequential", "permutation"):
                    pages = np.arange(n_pages, dtype=np.uint64)
                    s_off = pages * np.uint64(page_size)
                    d_pages = rng.permutation(n_pages).astype(np.uint64) if order == "permutation" else pages
                    d_off = d_pages * np.uint64(page_size)
                    if dst == src:
                        d_off = d_off + np.uint64(n_pages * page_size)
                    d_s = c.to_device(s_off); d_d = c.to_device(d_off)
                    for method, mname in ((0, "tma"), (1, "ldst")):
                        best = 1e30
                        for rep in range(4):
                            lib.uvmtier_event_record(c.handle, 0)
                            st = lib.uvmtier_copy_page_list_device(c.handle, dst, src, ctypes.c_size_t(n_pages), P(d_d), P(d_s),
                                                                   page_size, method)
                            assert st == 0, hex(st)
                            lib.uvmtier_event_record(c.handle, 1)
                            lib.uvmtier_event_elapsed_ms(c.handle, 0, 1, ctypes.byref(ms))
                            if rep:
                                best = min(best, ms.value)
                        gbs = n_pages * page_size / (best * 1e-3) / 1e9
                        print(json.dumps({"dst": NAMES[dst], "src": NAMES[src], "page_size": page_size, "order": order,
                                          "method": mname, "GBs": round(gbs, 2), "ce_peak_GBs": round(peak, 2) if peak else None,
                                          "frac_of_ce_peak": round(gbs / peak, 3) if peak else None}), flush=True)
                    c.free(d_s); c.free(d_d)


if __name__ == "__main__":
    main()
