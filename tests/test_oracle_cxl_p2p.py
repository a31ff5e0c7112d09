"""The fork's CXL P2P DMA surface on the CPU: the restatement (oracle/oracle_cxl_p2p.c), which the GPU tests compare
the library's uvmtier_p2p_* entry points with, against the reference's own file src/nvidia/src/kernel/mem_mgr/p2p_cxl.c
compiled unmodified (oracle/_ref/libref_cxl_p2p.so): status code of every argument check, page geometry of a
registration, the window RmP2PGetCxlPages returns, the bytes one RmP2PCxlDmaRequest moves."""
import ctypes

import numpy as np
import pytest

import uvmtier_testlib as T

U32, U64 = ctypes.c_uint32, ctypes.c_uint64
MB = 1 << 20


def _libs():
    ref = T.load_ref("cxl_p2p")
    if ref is None:
        pytest.skip("reference libraries not built (no /root/reference)")
    return T.load_oracle(), ref


def test_register_checks_and_page_geometry():
    orc, ref = _libs()
    rng = np.random.default_rng(1)
    base0 = 0x7F12_3440_0000
    cases = [(base0, 0, 2), (base0, MB, 0), (base0, MB, 4), (base0, (1 << 40) + 1, 2), (base0, 1 << 40, 2), ((1 << 64) - 4096, 8192, 2),
             (base0, 4 * MB, 1), (base0 + 4096, 4 * MB, 2), (base0, 4 * MB + 4096, 2), (base0 + 100, 5000, 3), (base0 + 4095, 2, 2),
             (base0, 1, 2), (base0, 2 * MB, 3), (base0, MB, 2)]
    for _ in range(300):
        cases.append((int(rng.integers(0, 1 << 47)) & ~int(rng.choice([0, 0xFFF, 0x1FFFFF])), int(rng.integers(1, 1 << int(rng.integers(1, 36)))) &
                      ~int(rng.choice([0, 0, 0xFFF, 0x1FFFFF])) or 4096, int(rng.integers(0, 5))))
    seen = set()
    for base, size, ver in cases:
        for has_out in (1, 0):
            for n_reg in (0, 255, 256):
                out = []
                for lib, pre in ((orc, "oracle"), (ref, "ref")):
                    ps, pc, first = U32(0), U64(0), U64(0)
                    fn = getattr(lib, pre + "_cxl_register_check")
                    fn.restype = U32
                    st = fn(U64(base), U64(size), U32(ver), has_out, U32(n_reg), ctypes.byref(ps), ctypes.byref(pc), ctypes.byref(first))
                    out.append((st, ps.value, pc.value, first.value) if st == 0 else (st,))
                assert out[0] == out[1], (hex(base), size, ver, has_out, n_reg, out)
                seen.add(out[0][0])
    assert {0, 0x1F, 0x1A} <= seen


def test_get_pages_window():
    orc, ref = _libs()
    rng = np.random.default_rng(2)
    seen = set()
    for _ in range(3000):
        page_size = int(rng.choice([4096, 2 * MB]))
        pages = int(rng.integers(1, 300))
        buf = pages * page_size - int(rng.integers(0, page_size)) * int(rng.integers(0, 2))
        off = int(rng.integers(0, buf + page_size)); n = int(rng.integers(0, buf + page_size))
        reg = int(rng.random() < 0.9)
        out = []
        for lib, pre in ((orc, "oracle"), (ref, "ref")):
            sp, npg = U32(0), U32(0)
            fn = getattr(lib, pre + "_cxl_get_pages")
            fn.restype = U32
            st = fn(U64(buf), U32(page_size), U32(pages), reg, U64(off), U64(n), ctypes.byref(sp), ctypes.byref(npg))
            out.append((st, sp.value, npg.value) if st == 0 and npg.value else (st,))
        assert out[0] == out[1], (buf, page_size, pages, reg, off, n, out)
        seen.add(out[0][0])
    assert {0, 0x1F, 0x40} <= seen


def test_dma_request_checks_and_clamp():
    orc, ref = _libs()
    rng = np.random.default_rng(3)
    cases = [(1, 1, 1, 8 * MB, 0, 0), (0, 1, 1, 8 * MB, 0, 16), (1, 0, 1, 8 * MB, 0, 16), (1, 1, 0, 8 * MB, 0, 16), (1, 1, 1, 8 * MB, 8 * MB, 1),
             (1, 1, 1, 8 * MB, 8 * MB + 1, 0), (1, 1, 1, 8 * MB, 0, 8 * MB + 1), (1, 1, 1, 8 * MB, 8 * MB - 1, 1), (1, 1, 1, 8 * MB, 0, 8 * MB),
             (1, 1, 1, 1 << 33, 0, 1 << 33), (1, 1, 1, 1 << 33, 4096, (1 << 32) - 1), (1, 1, 1, 1 << 33, 0, 1 << 32), (1, 1, 1, 1 << 33, 0, 0xFFFFFFFF)]
    for _ in range(2000):
        buf = int(rng.integers(1, 1 << 34))
        cases.append((int(rng.random() < 0.95), int(rng.random() < 0.95), int(rng.random() < 0.9), buf, int(rng.integers(0, buf + 10)),
                      int(rng.integers(0, buf + 10))))
    seen = set()
    for has_gpu, has_h, reg, buf, off, n in cases:
        out = []
        for lib, pre in ((orc, "oracle"), (ref, "ref")):
            b = U64(0)
            fn = getattr(lib, pre + "_cxl_dma_check")
            fn.restype = U32
            st = fn(has_gpu, has_h, reg, U64(buf), U64(off), U64(n), ctypes.byref(b))
            out.append((st, b.value) if st == 0 else (st,))
        assert out[0] == out[1], (has_gpu, has_h, reg, buf, off, n, out)
        seen.add(out[0])
    assert (0, 0xFFFFF000) in seen and (0x1F,) in seen and (0x40,) in seen


def test_system_info_defaults_when_enumeration_fails():
    _, ref = _libs()
    d, m, up, ver = U32(9), U32(9), U32(9), U32(9)
    ref.ref_cxl_system_info_defaults(ctypes.byref(d), ctypes.byref(m), ctypes.byref(up), ctypes.byref(ver))
    assert (d.value, m.value, up.value, ver.value) == (0, 0, 0, 2)     # what uvmtier_p2p_get_cxl_system_info reports without a CXL node
