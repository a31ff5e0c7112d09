"""GPU tests of the fork's CXL P2P DMA surface (include/uvmtier.h section 3b): the register -> DMA -> verify ->
unregister flow and byte pattern of the reference's own test (tests/cxl_p2p_test.c:467-513, 532-632), the status
codes of every argument check and the page arrays against the CPU restatement (oracle/oracle_cxl_p2p.c)."""
import ctypes
import mmap

import numpy as np
import pytest

import uvmtier
import uvmtier_testlib as T
from uvmtier import TIER_HBM

pytestmark = pytest.mark.gpu

OK, INVALID_ARGUMENT, INVALID_STATE, INSUFFICIENT_RESOURCES = 0, 0x1F, 0x40, 0x1A
CXL_TO_GPU, ASYNC = 1, 2
MB = 1 << 20


def pattern(n, seed):
    """init_test_pattern (tests/cxl_p2p_test.c:467-476): byte i = (i + seed) & 0xFF"""
    return ((np.arange(n, dtype=np.uint64) + seed) & 0xFF).astype(np.uint8)


class HostBuffer:
    """Page-aligned anonymous memory standing for the CXL-backed allocation of the reference test (:410-451);
    `skew` moves the base off the 2 MB boundary."""

    def __init__(self, size, skew=0):
        self.map = mmap.mmap(-1, size + 4 * MB)
        raw = ctypes.addressof(ctypes.c_char.from_buffer(self.map))
        self.addr = ((raw + 2 * MB - 1) & ~(2 * MB - 1)) + skew
        self.size = size
        self.view = np.frombuffer(self.map, np.uint8, size, self.addr - raw)


def hbm_bytes(ctx, info, off, n):
    return ctx.from_device(info.device_ptr + off, n, np.uint8)


def test_reference_flow_and_pattern():
    """The fork's test: a 4 MB buffer filled with (i + 0x42) & 0xFF goes CXL -> GPU, the buffer is overwritten with
    the 0xAB pattern, GPU -> CXL brings the first pattern back."""
    size = 4 * MB
    buf = HostBuffer(size)
    with uvmtier.Context(0) as c:
        info = c.tier_create(TIER_HBM, 8 * MB)
        buf.view[:] = pattern(size, 0x42)
        st, h = c.p2p_register_cxl_buffer(buf.addr, size, 2)
        assert st == OK
        st, addrs, ps = c.p2p_get_cxl_pages(h, 0, size)
        assert st == OK and ps == 2 * MB and list(addrs) == [buf.addr, buf.addr + 2 * MB]
        st, moved = c.p2p_cxl_dma_request(h, 0, 0, size, CXL_TO_GPU)
        assert (st, moved) == (OK, size)
        assert np.array_equal(hbm_bytes(c, info, 0, size), pattern(size, 0x42))
        buf.view[:] = pattern(size, 0xAB)
        st, moved = c.p2p_cxl_dma_request(h, 0, 0, size, 0)
        assert (st, moved) == (OK, size)
        assert np.array_equal(buf.view, pattern(size, 0x42))            # verify_test_pattern (:478-513)
        assert c.p2p_unregister_cxl_buffer(h) == OK


@pytest.mark.parametrize("skew,size,gpu_off,cxl_off,n", [
    (0, 6 * MB, 0, 0, 6 * MB),                       # 2 MB pages, whole buffer
    (0, 6 * MB, 2 * MB + 4096, 4096 * 3, 3 * MB + 8192),     # 4 KB-aligned interior range
    (4096, 5 * MB + 4096 * 3, 64, 16, 1 * MB + 48),   # 16-byte aligned, not page aligned: pieces + tail
    (4096 + 7, 3 * MB + 5, 3, 1, 2 * MB + 11),        # byte-granular everything
    (0, 2 * MB, 0, 100, 17),                          # a few bytes
])
def test_dma_ranges_are_byte_exact_in_both_directions(skew, size, gpu_off, cxl_off, n):
    buf = HostBuffer(size, skew)
    rng = np.random.default_rng(5)
    with uvmtier.Context(0) as c:
        info = c.tier_create(TIER_HBM, 16 * MB)
        st, h = c.p2p_register_cxl_buffer(buf.addr, size, 3)
        assert st == OK
        data = rng.integers(0, 256, size, dtype=np.uint8)
        buf.view[:] = data
        guard = hbm_bytes(c, info, 0, 16 * MB).copy()
        st, moved = c.p2p_cxl_dma_request(h, gpu_off, cxl_off, n, CXL_TO_GPU)
        assert (st, moved) == (OK, n)
        after = hbm_bytes(c, info, 0, 16 * MB)
        assert np.array_equal(after[gpu_off:gpu_off + n], data[cxl_off:cxl_off + n])
        assert np.array_equal(after[:gpu_off], guard[:gpu_off]) and np.array_equal(after[gpu_off + n:], guard[gpu_off + n:])
        # back into a different place of the buffer, asynchronously
        back_off = (size - n) // 2
        buf.view[:] = 0x5A
        st, moved = c.p2p_cxl_dma_request(h, gpu_off, back_off, n, ASYNC)
        c.synchronize()
        assert (st, moved) == (OK, n)
        want = np.full(size, 0x5A, np.uint8)
        want[back_off:back_off + n] = data[cxl_off:cxl_off + n]
        assert np.array_equal(buf.view, want)
        assert c.p2p_unregister_cxl_buffer(h) == OK


def test_argument_checks_and_page_arrays_match_the_restatement():
    orc = T.load_oracle()
    lib = uvmtier.lib()
    buf = HostBuffer(8 * MB)
    cnt, tot = ctypes.c_uint32(), ctypes.c_uint64()
    with uvmtier.Context(0) as c:
        c.tier_create(TIER_HBM, 8 * MB)
        lib.uvmtier_p2p_cxl_registered(ctypes.byref(cnt), ctypes.byref(tot))
        base_count, base_bytes = cnt.value, tot.value
        ps, pc, first = ctypes.c_uint32(), ctypes.c_uint64(), ctypes.c_uint64()
        # registration: bad arguments, then geometry of accepted buffers (2 MB pages only for 2 MB multiples)
        cases = [(buf.addr, 0, 2), (buf.addr, MB, 0), (buf.addr, MB, 4), (buf.addr, (1 << 40) + 1, 2), ((1 << 64) - 4096, 8192, 2),
                 (buf.addr, 4 * MB, 1), (buf.addr + 4096, 4 * MB, 2), (buf.addr, 4 * MB + 4096, 2), (buf.addr + 100, 5000, 3),
                 (buf.addr + 4095, 2, 2), (buf.addr, 1, 2)]
        for addr, size, ver in cases:
            want = orc.oracle_cxl_register_check(ctypes.c_uint64(addr), ctypes.c_uint64(size), ctypes.c_uint32(ver), 1,
                                                 ctypes.c_uint32(0), ctypes.byref(ps), ctypes.byref(pc), ctypes.byref(first))
            st, h = c.p2p_register_cxl_buffer(addr, size, ver)
            assert st == want, (hex(addr), size, ver)
            if st == OK:
                # the whole range: ceil(size / page size) pages from the start of the array -- for a base that is not
                # page aligned that can be one fewer than the buffer has (the reference's arithmetic, :455-456)
                st2, addrs, page_size = c.p2p_get_cxl_pages(h, 0, size)
                assert st2 == OK and page_size == ps.value and len(addrs) in (pc.value, pc.value - 1)
                assert len(addrs) == (size + ps.value - 1) // ps.value
                assert np.array_equal(addrs, first.value + np.arange(len(addrs), dtype=np.uint64) * np.uint64(ps.value))
                # sub-ranges of the page array
                rng = np.random.default_rng(size & 0xFFFF)
                for _ in range(20):
                    off = int(rng.integers(0, size + 2)); n = int(rng.integers(0, size + 2))
                    sp, npg = ctypes.c_uint32(), ctypes.c_uint32()
                    w = orc.oracle_cxl_get_pages(ctypes.c_uint64(size), ps, ctypes.c_uint32(pc.value), 1, ctypes.c_uint64(off),
                                                 ctypes.c_uint64(n), ctypes.byref(sp), ctypes.byref(npg))
                    st3, sub, _ = c.p2p_get_cxl_pages(h, off, n)
                    assert st3 == w
                    if w == OK:
                        assert np.array_equal(sub, addrs[sp.value:sp.value + npg.value])
                # DMA request checks
                for cxl_off, n in [(0, 0), (size, 1), (size + 1, 0), (0, size + 1), (size - 1, 1), (size - 1, 2), (0, size)]:
                    nbytes = ctypes.c_uint64()
                    w = orc.oracle_cxl_dma_check(1, 1, 1, ctypes.c_uint64(size), ctypes.c_uint64(cxl_off), ctypes.c_uint64(n),
                                                 ctypes.byref(nbytes))
                    st4, moved = c.p2p_cxl_dma_request(h, 0, cxl_off, n, CXL_TO_GPU)
                    assert st4 == w and (w != OK or moved == nbytes.value)
                assert lib.uvmtier_p2p_put_cxl_pages(h, ctypes.c_uint64(0), ctypes.c_uint64(size)) == OK
                assert c.p2p_unregister_cxl_buffer(h) == OK
        # NULL handle, NULL out pointer
        assert c.p2p_unregister_cxl_buffer(ctypes.c_void_p()) == INVALID_ARGUMENT
        assert c.p2p_cxl_dma_request(ctypes.c_void_p(), 0, 0, 16, 0)[0] == INVALID_ARGUMENT
        assert lib.uvmtier_p2p_register_cxl_buffer(c.handle, ctypes.c_void_p(buf.addr), ctypes.c_uint64(MB), ctypes.c_uint32(2),
                                                   None) == INVALID_ARGUMENT
        # a request that reaches outside the HBM tier
        st, h = c.p2p_register_cxl_buffer(buf.addr, 8 * MB, 2)
        assert st == OK and c.p2p_cxl_dma_request(h, 4 * MB, 0, 4 * MB + 16, CXL_TO_GPU)[0] == INVALID_ARGUMENT
        assert c.p2p_unregister_cxl_buffer(h) == OK
        # the 257th buffer is refused (CXL_P2P_MAX_REGISTERED_BUFFERS), and the counters come back to where they were
        handles = []
        for k in range(256 - base_count):
            st, h = c.p2p_register_cxl_buffer(buf.addr + 4096 * k, 4096, 2)
            assert st == OK
            handles.append(h)
        assert c.p2p_register_cxl_buffer(buf.addr, 4096, 2)[0] == INSUFFICIENT_RESOURCES
        lib.uvmtier_p2p_cxl_registered(ctypes.byref(cnt), ctypes.byref(tot))
        assert cnt.value == 256 and tot.value == base_bytes + 4096 * len(handles)
        for h in handles:
            assert c.p2p_unregister_cxl_buffer(h) == OK
        lib.uvmtier_p2p_cxl_registered(ctypes.byref(cnt), ctypes.byref(tot))
        assert (cnt.value, tot.value) == (base_count, base_bytes)


def test_system_info_reports_cpu_less_nodes():
    n, m, up, ver = ctypes.c_uint32(), ctypes.c_uint32(), ctypes.c_uint32(), ctypes.c_uint32()
    assert uvmtier.lib().uvmtier_p2p_get_cxl_system_info(ctypes.byref(n), ctypes.byref(m), ctypes.byref(up), ctypes.byref(ver)) == OK
    dram = ctypes.c_int(); cxl = ctypes.c_int()
    uvmtier.lib().uvmtier_numa_probe(ctypes.byref(dram), ctypes.byref(cxl))
    assert (m.value > 0) == (cxl.value >= 0) and up.value == (1 if m.value else 0) and ver.value in (1, 2, 3)
    assert uvmtier.lib().uvmtier_p2p_get_cxl_system_info(None, None, None, None) == OK
