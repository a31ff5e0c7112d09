"""uvmtier -- thin ctypes binding over libuvmtier.so (the C ABI declared in include/uvmtier.h).

This module is plumbing for tests/ and bench.py; the product is the C library.  It mirrors the
reference-facing names one to one and never computes anything itself: if the CUDA library cannot
be loaded or no B200 is present, calls raise -- there is no CPU fallback.
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "libuvmtier.so")

NV_OK = 0
NV_WARN_MORE_PROCESSING_REQUIRED = 0x10005
ID_CPU = 0
ID_INVALID = 257


class UvmtierError(RuntimeError):
    def __init__(self, status, where, detail=""):
        super().__init__("%s failed: NV_STATUS 0x%x %s" % (where, status, detail))
        self.status = status


_lib = None


def lib():
    """Loads libuvmtier.so (once). Raises if the extension has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise UvmtierError(0xFFFF, "load", "%s not built (run __graft_entry__.build())" % LIB_PATH)
        _lib = ctypes.CDLL(LIB_PATH)
        _lib.uvmtier_version.restype = ctypes.c_char_p
        _lib.uvmtier_last_error.restype = ctypes.c_char_p
        _lib.uvmtier_kernel_launch_count.restype = ctypes.c_uint64
        _lib.uvmtier_stream.restype = ctypes.c_void_p
        _lib.uvmtier_stream.argtypes = [ctypes.c_void_p]
    return _lib


def _check(st, where):
    if st != NV_OK:
        raise UvmtierError(st, where, lib().uvmtier_last_error().decode())


def P(a):
    """numpy array / int device pointer / None -> void*"""
    if a is None:
        return ctypes.c_void_p(0)
    if isinstance(a, (int, np.integer)):
        return ctypes.c_void_p(int(a))
    assert a.flags["C_CONTIGUOUS"], "array must be C contiguous"
    return a.ctypes.data_as(ctypes.c_void_p)


class Config(ctypes.Structure):
    _fields_ = [
        ("uvm_perf_prefetch_enable", ctypes.c_uint32),
        ("uvm_perf_prefetch_threshold", ctypes.c_uint32),
        ("uvm_perf_prefetch_min_faults", ctypes.c_uint32),
        ("uvm_perf_fault_batch_count", ctypes.c_uint32),
        ("uvm_perf_fault_coalesce", ctypes.c_uint32),
        ("reserved", ctypes.c_uint32 * 11),
    ]


def default_config():
    cfg = Config()
    lib().uvmtier_config_default(ctypes.byref(cfg))
    return cfg


def kernel_launch_count():
    return int(lib().uvmtier_kernel_launch_count())


class Context:
    """uvmtier_ctx_t: one per GPU (the reference's per-parent-GPU batch_service_context)."""

    def __init__(self, device=0, **tunables):
        cfg = default_config()
        for k, v in tunables.items():
            if not hasattr(cfg, k):
                raise AttributeError(k)
            setattr(cfg, k, v)
        self._h = ctypes.c_void_p()
        _check(lib().uvmtier_init(ctypes.c_int(device), ctypes.byref(cfg), ctypes.byref(self._h)), "uvmtier_init")
        self.device = device

    def close(self):
        if self._h:
            lib().uvmtier_deinit(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    @property
    def handle(self):
        return self._h

    @property
    def stream(self):
        return lib().uvmtier_stream(self._h)

    def config(self):
        cfg = Config()
        lib().uvmtier_get_config(self._h, ctypes.byref(cfg))
        return cfg

    def synchronize(self):
        _check(lib().uvmtier_synchronize(self._h), "uvmtier_synchronize")

    # ---- raw device memory (for the *_device entry points) ----
    def alloc(self, nbytes):
        p = ctypes.c_void_p()
        _check(lib().uvmtier_mem_alloc_device(self._h, ctypes.c_size_t(nbytes), ctypes.byref(p)), "mem_alloc_device")
        return p.value

    def free(self, dptr):
        _check(lib().uvmtier_mem_free_device(self._h, ctypes.c_void_p(dptr)), "mem_free_device")

    def to_device(self, arr):
        arr = np.ascontiguousarray(arr)
        d = self.alloc(arr.nbytes)
        _check(lib().uvmtier_memcpy_h2d(self._h, ctypes.c_void_p(d), P(arr), ctypes.c_size_t(arr.nbytes)), "memcpy_h2d")
        return d

    def from_device(self, dptr, shape, dtype):
        out = np.empty(shape, dtype)
        _check(lib().uvmtier_memcpy_d2h(self._h, P(out), ctypes.c_void_p(dptr), ctypes.c_size_t(out.nbytes)), "memcpy_d2h")
        return out

    # ---- (1) fault batches ----
    def fault_buffer_init(self, gpu_params, table):
        """uvmtier_fault_buffer_init: parent-GPU parameters + instance-pointer table."""
        _check(lib().uvmtier_fault_buffer_init(self._h, P(gpu_params), P(table), ctypes.c_size_t(len(table))),
               "uvmtier_fault_buffer_init")
        self._fault_B = int(gpu_params["max_batch_size"][0])

    def fault_buffer_parse_entries(self, packets):
        from .layouts import FAULT_ENTRY_DT
        out = np.zeros(len(packets), FAULT_ENTRY_DT)
        _check(lib().uvmtier_fault_buffer_parse_entries(self._h, P(packets), ctypes.c_size_t(len(packets)), P(out)),
               "uvmtier_fault_buffer_parse_entries")
        return out

    def fetch_preprocess_fault_batches(self, ring, get, put, clear_valid=True, want_entries=True):
        """Host-buffer form. ring: uint32[ring_entries, 8] (updated in place when clear_valid).
        Returns dict(disp, ordered, groups, batch, entries, get)."""
        from .layouts import BATCH_DT, FAULT_ENTRY_DT, GROUP_DT, ORDERED_DT
        ring_entries = len(ring)
        n = put - get if put >= get else ring_entries - get + put
        B = self._fault_B
        nb = (n + B - 1) // B
        disp = np.zeros(n, "<u4")
        ordered = np.zeros(n, ORDERED_DT)
        groups = np.zeros(n, GROUP_DT)
        batch = np.zeros(nb, BATCH_DT)
        entries = np.zeros(n, FAULT_ENTRY_DT) if want_entries else None
        new_get = ctypes.c_uint32(0)
        _check(lib().uvmtier_fetch_preprocess_fault_batches(self._h, P(ring), ctypes.c_uint32(ring_entries),
                                                            ctypes.c_uint32(get), ctypes.c_uint32(put),
                                                            ctypes.c_int(1 if clear_valid else 0), P(disp), P(ordered),
                                                            P(groups), P(batch), P(entries), ctypes.byref(new_get)),
               "uvmtier_fetch_preprocess_fault_batches")
        return dict(disp=disp, ordered=ordered, groups=groups, batch=batch, entries=entries, get=new_get.value)

    # ---- (2) prefetch ----
    def perf_prefetch_compute_ats(self, faulted, faulted_region, max_prefetch_region, residency=None):
        """uvm_perf_prefetch_compute_ats for n blocks; host arrays in, uint64[n,8] mask out."""
        n = len(faulted)
        out = np.zeros((n, 8), "<u8")
        _check(lib().uvmtier_perf_prefetch_compute_ats(self._h, ctypes.c_size_t(n), P(faulted), P(faulted_region),
                                                       P(max_prefetch_region), P(residency), P(out)),
               "uvmtier_perf_prefetch_compute_ats")
        return out

    def perf_prefetch_get_hint_va_block(self, blocks, resident, faulted, thrashing, cpu_read_mapped, info):
        """uvm_perf_prefetch_get_hint_va_block for n blocks. Returns (mask[n,8], residency[n], info[n])."""
        n = len(blocks)
        out = np.zeros((n, 8), "<u8")
        res = np.zeros(n, "<u4")
        info = info.copy()
        _check(lib().uvmtier_perf_prefetch_get_hint_va_block(self._h, ctypes.c_size_t(n), P(blocks), P(resident),
                                                             P(faulted), P(thrashing), P(cpu_read_mapped), P(info),
                                                             P(out), P(res)),
               "uvmtier_perf_prefetch_get_hint_va_block")
        return out, res, info
