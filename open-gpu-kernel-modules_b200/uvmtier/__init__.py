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


TIER_HBM, TIER_DRAM, TIER_CXL, TIER_PEER0 = 0, 1, 2, 8
TIERF_CXL_EMULATED = 0x1


class PmmState(ctypes.Structure):
    """uvmtier_pmm_t"""
    _fields_ = [("d_stamp", ctypes.c_void_p), ("d_scratch", ctypes.c_void_p), ("clock", ctypes.c_uint64),
                ("n_slots", ctypes.c_uint32), ("capacity_chunks", ctypes.c_uint32), ("d_chunk_of_slot", ctypes.c_void_p),
                ("d_free_bits", ctypes.c_void_p), ("d_dram_stamp", ctypes.c_void_p), ("round", ctypes.c_uint64)]


class ThrashingState(ctypes.Structure):
    """uvmtier_thrashing_t"""
    _fields_ = [("d_blocks", ctypes.c_void_p), ("d_page_word", ctypes.c_void_p), ("d_page_processors", ctypes.c_void_p),
                ("d_thrashing_pages", ctypes.c_void_p), ("d_scratch", ctypes.c_void_p), ("lapse_ns", ctypes.c_uint64),
                ("threshold", ctypes.c_uint32), ("n_slots", ctypes.c_uint32), ("d_page_throttle", ctypes.c_void_p),
                ("d_page_hint", ctypes.c_void_p), ("d_page_throttled", ctypes.c_void_p), ("d_page_deadline", ctypes.c_void_p),
                ("d_pinned_pages", ctypes.c_void_p), ("d_hint_config", ctypes.c_void_p), ("hint_config", ctypes.c_uint8 * 560)]


class TieringStats(ctypes.Structure):
    """uvmtier_tiering_stats_t"""
    _fields_ = [("passes", ctypes.c_uint32), ("eviction_waves", ctypes.c_uint32), ("evicted_to_dram", ctypes.c_uint32),
                ("evicted_to_cxl", ctypes.c_uint32), ("peak_chunks_in_use", ctypes.c_uint32), ("chunks_in_use", ctypes.c_uint32),
                ("pages_evicted", ctypes.c_uint64), ("pages_promoted", ctypes.c_uint64), ("overcommitted_passes", ctypes.c_uint32),
                ("blocks_demoted", ctypes.c_uint32), ("reserved", ctypes.c_uint32 * 2), ("pages_demoted", ctypes.c_uint64)]


class TierInfo(ctypes.Structure):
    """uvmtier_tier_info_t"""
    _fields_ = [("bytes", ctypes.c_uint64), ("device_ptr", ctypes.c_void_p), ("host_ptr", ctypes.c_void_p),
                ("numa_node", ctypes.c_int32), ("flags", ctypes.c_uint32)]

    def host_view(self):
        """numpy uint8 view of a host tier's arena."""
        assert self.host_ptr, "not a host tier"
        return np.frombuffer((ctypes.c_uint8 * self.bytes).from_address(self.host_ptr), dtype=np.uint8)


def numa_probe():
    """(online nodes, dram node, cxl node or -1)"""
    d, c = ctypes.c_int(-1), ctypes.c_int(-1)
    n = lib().uvmtier_numa_probe(ctypes.byref(d), ctypes.byref(c))
    return n, d.value, c.value


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

    def fault_buffer_set_utlb_pending_output(self, dptr):
        """uvmtier_fault_buffer_set_utlb_pending_output: device array of uint16 [n_batches, utlb_count], or None."""
        _check(lib().uvmtier_fault_buffer_set_utlb_pending_output(self._h, ctypes.c_void_p(dptr if dptr else None)),
               "uvmtier_fault_buffer_set_utlb_pending_output")

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

    def fetch_preprocess_fault_batches_streamed(self, ring, get, put, clear_valid=True, chunk_batches=0):
        """uvmtier_fetch_preprocess_fault_batches_streamed on a copy of `ring` in pinned host memory: the ring is
        copied to the device chunk by chunk while earlier chunks are preprocessed.  Returns the same dict as
        fetch_preprocess_fault_batches plus ring_device (the device ring after the call: VALID bits cleared)."""
        from .layouts import BATCH_DT, FAULT_ENTRY_DT, GROUP_DT, ORDERED_DT
        ring = np.ascontiguousarray(ring)
        ring_entries = len(ring)
        n = put - get if put >= get else ring_entries - get + put
        B = self._fault_B
        nb = (n + B - 1) // B
        hp = ctypes.c_void_p()
        _check(lib().uvmtier_mem_alloc_host_pinned(self._h, ctypes.c_size_t(ring.nbytes), ctypes.byref(hp)), "pinned")
        ctypes.memmove(hp.value, ring.ctypes.data, ring.nbytes)
        d_ring = self.alloc(ring.nbytes)
        d_disp, d_ord, d_grp, d_bat = self.alloc(n * 4), self.alloc(n * 16), self.alloc(n * 16), self.alloc(nb * 32)
        d_ent = self.alloc(n * FAULT_ENTRY_DT.itemsize)
        try:
            for p, nbytes, v in ((d_ring, ring.nbytes, 0), (d_ord, n * 16, 0xff), (d_grp, n * 16, 0xff),
                                 (d_ent, n * FAULT_ENTRY_DT.itemsize, 0)):
                _check(lib().uvmtier_memset_device(self._h, ctypes.c_void_p(p), ctypes.c_int(v), ctypes.c_size_t(nbytes)), "memset")
            new_get = ctypes.c_uint32(0)
            _check(lib().uvmtier_fetch_preprocess_fault_batches_streamed(
                self._h, hp, P(d_ring), ctypes.c_uint32(ring_entries), ctypes.c_uint32(get), ctypes.c_uint32(put),
                ctypes.c_int(1 if clear_valid else 0), ctypes.c_uint32(chunk_batches), P(d_disp), P(d_ord), P(d_grp), P(d_bat),
                P(d_ent), ctypes.byref(new_get)), "uvmtier_fetch_preprocess_fault_batches_streamed")
            self.synchronize()
            return dict(disp=self.from_device(d_disp, n, "<u4"), ordered=self.from_device(d_ord, n, ORDERED_DT),
                        groups=self.from_device(d_grp, n, GROUP_DT), batch=self.from_device(d_bat, nb, BATCH_DT),
                        entries=self.from_device(d_ent, n, FAULT_ENTRY_DT), get=new_get.value,
                        ring_device=self.from_device(d_ring, ring.shape, ring.dtype))
        finally:
            for p in (d_ring, d_disp, d_ord, d_grp, d_bat, d_ent):
                self.free(p)
            lib().uvmtier_mem_free_host_pinned(self._h, hp)

    # ---- (1b) servicing ----
    def block_state_alloc(self, layout):
        """uvmtier_block_state_alloc -> DeviceBlockState (device pointers + upload/download helpers)."""
        return DeviceBlockState(self, layout)

    def service_fault_batches(self, k1, layout, dstate, run_cap=1 << 20):
        """Full device path on the results of fetch_preprocess_fault_batches (dict k1, host arrays): uploads the
        sorted view, runs uvmtier_service_fault_batches_device, downloads flags / records / runs / counters."""
        from .blockstate import COPY_RUN_DT, SERVICE_GROUP_DT
        from .layouts import BATCH_DT
        nb = len(k1["batch"])
        total = len(k1["ordered"])
        d_ord = self.to_device(k1["ordered"]); d_grp = self.to_device(k1["groups"]); d_bat = self.to_device(k1["batch"])
        d_flags = self.alloc(total * 2); d_svc = self.alloc(total * 96); d_runs = self.alloc(max(run_cap, 1) * 8)
        d_cnt = self.alloc(8)
        try:
            _check(lib().uvmtier_memset_device(self._h, ctypes.c_void_p(d_svc), 0, ctypes.c_size_t(total * 96)), "memset")
            _check(lib().uvmtier_service_fault_batches_device(self._h, ctypes.c_uint32(nb), P(d_ord), P(d_grp), P(d_bat),
                                                              P(layout), ctypes.byref(dstate.ptrs), P(d_flags), P(d_svc),
                                                              P(d_runs), ctypes.c_uint32(run_cap), P(d_cnt)),
                   "uvmtier_service_fault_batches_device")
            self.synchronize()
            cnt = self.from_device(d_cnt, 2, "<u4")
            return dict(flags=self.from_device(d_flags, total, "<u2"), service=self.from_device(d_svc, total, SERVICE_GROUP_DT),
                        runs=self.from_device(d_runs, max(run_cap, 1), COPY_RUN_DT), n_runs=int(cnt[0]),
                        batch=self.from_device(d_bat, nb, BATCH_DT), state=dstate.download())
        finally:
            for p in (d_ord, d_grp, d_bat, d_flags, d_svc, d_runs, d_cnt):
                self.free(p)

    def service_fault_batches_thrashing(self, k1, layout, dstate, thrashing, batch_time_ns, run_cap=1 << 20):
        """uvmtier_service_fault_batches_thrashing_device: service_fault_batches with the thrashing hint inside the fault loop.
        thrashing: a ThrashingState with hints enabled; batch_time_ns: uint64 per batch.  Adds the eight thrashing counters."""
        from .blockstate import COPY_RUN_DT, SERVICE_GROUP_DT
        from .layouts import BATCH_DT
        nb = len(k1["batch"])
        total = len(k1["ordered"])
        d_ord = self.to_device(k1["ordered"]); d_grp = self.to_device(k1["groups"]); d_bat = self.to_device(k1["batch"])
        d_time = self.to_device(np.ascontiguousarray(batch_time_ns, "<u8"))
        d_flags = self.alloc(total * 2); d_svc = self.alloc(total * 96); d_runs = self.alloc(max(run_cap, 1) * 8)
        d_cnt = self.alloc(8); d_tcnt = self.alloc(32)
        try:
            _check(lib().uvmtier_memset_device(self._h, ctypes.c_void_p(d_svc), 0, ctypes.c_size_t(total * 96)), "memset")
            _check(lib().uvmtier_service_fault_batches_thrashing_device(
                self._h, ctypes.c_uint32(nb), P(d_ord), P(d_grp), P(d_bat), P(layout), ctypes.byref(dstate.ptrs), ctypes.byref(thrashing),
                P(d_time), P(d_flags), P(d_svc), P(d_runs), ctypes.c_uint32(run_cap), P(d_cnt), P(d_tcnt)),
                "uvmtier_service_fault_batches_thrashing_device")
            self.synchronize()
            cnt = self.from_device(d_cnt, 2, "<u4")
            return dict(flags=self.from_device(d_flags, total, "<u2"), service=self.from_device(d_svc, total, SERVICE_GROUP_DT),
                        runs=self.from_device(d_runs, max(run_cap, 1), COPY_RUN_DT), n_runs=int(cnt[0]),
                        batch=self.from_device(d_bat, nb, BATCH_DT), state=dstate.download(),
                        thrashing_counters=self.from_device(d_tcnt, 8, "<u4"))
        finally:
            for p in (d_ord, d_grp, d_bat, d_time, d_flags, d_svc, d_runs, d_cnt, d_tcnt):
                self.free(p)

    def thrashing_upload(self, t, tables):
        """Copies host tables (a dict with the names thrashing_tables returns; blocks as 32-byte rows) into a ThrashingState."""
        for name, dptr in (("blocks", t.d_blocks), ("page_word", t.d_page_word), ("page_procs", t.d_page_processors),
                           ("thrashing_pages", t.d_thrashing_pages), ("page_throttle", t.d_page_throttle), ("page_hint", t.d_page_hint),
                           ("page_throttled", t.d_page_throttled), ("page_deadline", t.d_page_deadline), ("pinned_pages", t.d_pinned_pages)):
            if name in tables and dptr:
                a = np.ascontiguousarray(tables[name])
                _check(lib().uvmtier_memcpy_h2d(self._h, ctypes.c_void_p(dptr), P(a), ctypes.c_size_t(a.nbytes)), "h2d")

    def service_replayable_faults_streamed(self, ring, get, put, layout, dstate, run_cap=1 << 20, chunk_batches=0,
                                           clear_valid=True, repeat=1, before_each=None):
        """uvmtier_service_replayable_faults_streamed on a copy of `ring` in pinned host memory: fetch -> preprocess ->
        service chunk by chunk while later chunks cross the link; batch headers, entry flags and counters come back
        through the chunked copy-back (pinned buffers), everything else is downloaded afterwards.  Same dict as
        service_fault_batches plus the preprocess outputs (ordered, groups, disp).  repeat > 1 makes the same call
        again with the same buffers (before_each() restores the state table): the library then replays a CUDA graph."""
        from .blockstate import COPY_RUN_DT, SERVICE_GROUP_DT
        from .layouts import BATCH_DT, GROUP_DT, ORDERED_DT

        class Readback(ctypes.Structure):
            _fields_ = [("h_batches", ctypes.c_void_p), ("h_entry_flags", ctypes.c_void_p), ("h_counters", ctypes.c_void_p)]

        ring = np.ascontiguousarray(ring)
        ring_entries = len(ring)
        n = put - get if put >= get else ring_entries - get + put
        B = self._fault_B
        nb = (n + B - 1) // B
        total = nb * B
        pinned = []

        def pin(nbytes):
            hp = ctypes.c_void_p()
            _check(lib().uvmtier_mem_alloc_host_pinned(self._h, ctypes.c_size_t(max(nbytes, 8)), ctypes.byref(hp)), "pinned")
            pinned.append(hp)
            return hp

        hp = pin(ring.nbytes)
        ctypes.memmove(hp.value, ring.ctypes.data, ring.nbytes)
        rb = Readback(pin(nb * 32).value, pin(total * 2).value, pin(8).value)
        ctypes.memset(rb.h_entry_flags, 0, total * 2)
        d_ring = self.alloc(ring.nbytes)
        d_disp, d_ord, d_grp, d_bat = self.alloc(total * 4), self.alloc(total * 16), self.alloc(total * 16), self.alloc(nb * 32)
        d_flags = self.alloc(total * 2); d_svc = self.alloc(total * 96); d_runs = self.alloc(max(run_cap, 1) * 8)
        d_cnt = self.alloc(8)
        devs = (d_ring, d_disp, d_ord, d_grp, d_bat, d_flags, d_svc, d_runs, d_cnt)
        try:
            for p, nbytes, v in ((d_ring, ring.nbytes, 0), (d_ord, total * 16, 0xff), (d_grp, total * 16, 0xff), (d_svc, total * 96, 0),
                                 (d_flags, total * 2, 0)):
                _check(lib().uvmtier_memset_device(self._h, ctypes.c_void_p(p), ctypes.c_int(v), ctypes.c_size_t(nbytes)), "memset")
            new_get = ctypes.c_uint32(0)
            for _ in range(repeat):
                if before_each is not None:
                    before_each()
                    ctypes.memset(rb.h_entry_flags, 0xAB, total * 2)
                    ctypes.memset(rb.h_batches, 0xAB, nb * 32)
                _check(lib().uvmtier_service_replayable_faults_streamed(
                    self._h, hp, P(d_ring), ctypes.c_uint32(ring_entries), ctypes.c_uint32(get), ctypes.c_uint32(put),
                    ctypes.c_int(1 if clear_valid else 0), ctypes.c_uint32(chunk_batches), P(d_disp), P(d_ord), P(d_grp), P(d_bat),
                    P(layout), ctypes.byref(dstate.ptrs), P(d_flags), P(d_svc), P(d_runs), ctypes.c_uint32(run_cap), P(d_cnt),
                    ctypes.byref(rb), ctypes.byref(new_get)), "uvmtier_service_replayable_faults_streamed")
                self.synchronize()
            host = lambda ptr, count, dt: np.frombuffer((ctypes.c_char * (count * np.dtype(dt).itemsize)).from_address(ptr),
                                                        dtype=dt, count=count).copy()
            cnt = host(rb.h_counters, 2, "<u4")
            ctypes.memset(rb.h_entry_flags + n * 2, 0, (total - n) * 2)      # the copy-back covers the n entries of the trace
            return dict(flags=host(rb.h_entry_flags, total, "<u2"), service=self.from_device(d_svc, total, SERVICE_GROUP_DT),
                        runs=self.from_device(d_runs, max(run_cap, 1), COPY_RUN_DT), n_runs=int(cnt[0]), n_records=int(cnt[1]),
                        batch=host(rb.h_batches, nb, BATCH_DT), state=dstate.download(), get=new_get.value,
                        ordered=self.from_device(d_ord, total, ORDERED_DT), groups=self.from_device(d_grp, total, GROUP_DT),
                        disp=self.from_device(d_disp, n, "<u4"),
                        device_copies=dict(flags=self.from_device(d_flags, total, "<u2"), batch=self.from_device(d_bat, nb, BATCH_DT),
                                           counters=self.from_device(d_cnt, 2, "<u4")))
        finally:
            for p in devs:
                self.free(p)
            for h in pinned:
                lib().uvmtier_mem_free_host_pinned(self._h, h)

    # ---- (1d) eviction order + servicing under an HBM capacity ----
    def pmm_create(self, layout, capacity_chunks):
        pmm = PmmState()
        _check(lib().uvmtier_pmm_create(self._h, P(layout), ctypes.c_uint32(capacity_chunks), ctypes.byref(pmm)), "uvmtier_pmm_create")
        return pmm

    def pmm_destroy(self, pmm):
        _check(lib().uvmtier_pmm_destroy(self._h, ctypes.byref(pmm)), "uvmtier_pmm_destroy")

    def thrashing_create(self, layout, lapse_ns=0, threshold=0):
        t = ThrashingState()
        _check(lib().uvmtier_thrashing_create(self._h, P(layout), ctypes.c_uint64(lapse_ns), ctypes.c_uint32(threshold), ctypes.byref(t)),
               "uvmtier_thrashing_create")
        return t

    def thrashing_destroy(self, t):
        _check(lib().uvmtier_thrashing_destroy(self._h, ctypes.byref(t)), "uvmtier_thrashing_destroy")

    def thrashing_note_events(self, t, events):
        """uvmtier_thrashing_note_events_device on a host array of uvmtier_perf_event_t (24-byte records, grouped by
        slot).  Returns the four counters."""
        d_ev, d_cnt = self.to_device(events), self.alloc(16)
        try:
            _check(lib().uvmtier_thrashing_note_events_device(self._h, ctypes.byref(t), ctypes.c_uint32(len(events)), P(d_ev), P(d_cnt)),
                   "uvmtier_thrashing_note_events_device")
            self.synchronize()
            return self.from_device(d_cnt, 4, "<u4")
        finally:
            self.free(d_ev); self.free(d_cnt)

    def thrashing_tables(self, t):
        """The heuristic's tables back on the host: va_block records (32-byte rows), page words, page processors, masks, and
        the hint half's tables when it is on."""
        n = t.n_slots
        out = dict(blocks=self.from_device(t.d_blocks, n * 32, "u1"), page_word=self.from_device(t.d_page_word, n * 512, "<u8"),
                   page_procs=self.from_device(t.d_page_processors, n * 512, "<u4"),
                   thrashing_pages=self.from_device(t.d_thrashing_pages, n * 16, "<u4"))
        if t.d_page_throttle:
            out.update(page_throttle=self.from_device(t.d_page_throttle, n * 512, "<u8"), page_hint=self.from_device(t.d_page_hint, n * 512, "<u4"),
                       page_throttled=self.from_device(t.d_page_throttled, n * 512, "<u4"),
                       page_deadline=self.from_device(t.d_page_deadline, n * 512, "<u8"),
                       pinned_pages=self.from_device(t.d_pinned_pages, n * 16, "<u4"))
        return out

    def thrashing_enable_hints(self, t, cfg):
        """uvmtier_thrashing_enable_hints; cfg: a 560-byte uvmtier_thrashing_hint_config_t record (numpy)."""
        _check(lib().uvmtier_thrashing_enable_hints(self._h, ctypes.byref(t), P(cfg)), "uvmtier_thrashing_enable_hints")

    def thrashing_get_hints(self, t, queries):
        """uvmtier_thrashing_get_hints_device on a host array of 24-byte queries (grouped by slot).  Returns (hints as 16-byte
        records, the eight counters)."""
        d_q, d_h, d_cnt = self.to_device(queries), self.alloc(max(len(queries), 1) * 16), self.alloc(32)
        try:
            _check(lib().uvmtier_thrashing_get_hints_device(self._h, ctypes.byref(t), ctypes.c_uint32(len(queries)), P(d_q), P(d_h), P(d_cnt)),
                   "uvmtier_thrashing_get_hints_device")
            self.synchronize()
            return self.from_device(d_h, len(queries) * 16, "u1"), self.from_device(d_cnt, 8, "<u4")
        finally:
            self.free(d_q); self.free(d_h); self.free(d_cnt)

    def thrashing_unpin_expired(self, t, now_ns):
        d_cnt = self.alloc(32)
        try:
            _check(lib().uvmtier_thrashing_unpin_expired_device(self._h, ctypes.byref(t), ctypes.c_uint64(now_ns), P(d_cnt)),
                   "uvmtier_thrashing_unpin_expired_device")
            self.synchronize()
            return self.from_device(d_cnt, 8, "<u4")
        finally:
            self.free(d_cnt)

    def pmm_stamps(self, pmm):
        return self.from_device(pmm.d_stamp, pmm.n_slots, "<u8")

    def pmm_enable_chunk_table(self, pmm):
        """uvmtier_pmm_enable_chunk_table: the HBM arena holds pmm.capacity_chunks chunks, va_blocks take and return them."""
        _check(lib().uvmtier_pmm_enable_chunk_table(self._h, ctypes.byref(pmm)), "uvmtier_pmm_enable_chunk_table")

    def pmm_chunk_table(self, pmm):
        """chunk_of_slot (0xFFFFFFFF = none); None when the table is off."""
        return self.from_device(pmm.d_chunk_of_slot, pmm.n_slots, "<u4") if pmm.d_chunk_of_slot else None

    def service_fault_batches_tiered(self, k1, layout, dstate, pmm, chunk_batches, dram_capacity_pages, execute_copies=True,
                                     demote=False):
        """uvmtier_service_fault_batches_tiered_device on the results of fetch_preprocess_fault_batches (dict k1, host
        arrays, padded to whole batches).  Returns flags / records / batch headers / state / stats / stamps."""
        from .blockstate import SERVICE_GROUP_DT
        from .layouts import BATCH_DT
        nb = len(k1["batch"])
        total = len(k1["ordered"])
        d_ord = self.to_device(k1["ordered"]); d_grp = self.to_device(k1["groups"]); d_bat = self.to_device(k1["batch"])
        d_flags = self.alloc(total * 2); d_svc = self.alloc(total * 96)
        stats = TieringStats()
        try:
            _check(lib().uvmtier_memset_device(self._h, ctypes.c_void_p(d_svc), 0, ctypes.c_size_t(total * 96)), "memset")
            _check(lib().uvmtier_service_fault_batches_tiered_device(
                self._h, ctypes.byref(pmm), ctypes.c_uint32(nb), P(d_ord), P(d_grp), P(d_bat), P(layout), ctypes.byref(dstate.ptrs),
                P(d_flags), P(d_svc), ctypes.c_uint32(chunk_batches), ctypes.c_uint64(dram_capacity_pages),
                ctypes.c_int((1 if execute_copies else 0) | (2 if demote else 0)), ctypes.byref(stats)),
                "uvmtier_service_fault_batches_tiered_device")
            self.synchronize()
            return dict(flags=self.from_device(d_flags, total, "<u2"), service=self.from_device(d_svc, total, SERVICE_GROUP_DT),
                        batch=self.from_device(d_bat, nb, BATCH_DT), state=dstate.download(), stats=stats,
                        stamps=self.pmm_stamps(pmm), dram_stamps=self.from_device(pmm.d_dram_stamp, pmm.n_slots, "<u8"),
                        chunk_of_slot=self.pmm_chunk_table(pmm))
        finally:
            for p in (d_ord, d_grp, d_bat, d_flags, d_svc):
                self.free(p)

    # ---- (1e) access-counter notifications ----
    def access_counters_init(self, granularity_bytes=2 << 20, cpu_chunk_bytes=2 << 20):
        cfg = (ctypes.c_uint64 * 2)(granularity_bytes, cpu_chunk_bytes)
        _check(lib().uvmtier_access_counters_init(self._h, cfg), "uvmtier_access_counters_init")

    def access_counters_service(self, ring, get, put, layout, dstate, run_cap=1 << 20, clear_valid=True):
        """uvmtier_access_counters_fetch_preprocess_device + uvmtier_access_counters_service_device on a notification
        ring (host array, uploaded).  Returns the preprocess outputs, records, runs and the state."""
        from .blockstate import COPY_RUN_DT, SERVICE_GROUP_DT
        from .layouts import BATCH_DT, GROUP_DT
        onotif = np.dtype([("address", "<u8"), ("index", "<u2"), ("ctx", "<u2"), ("sub_granularity", "<u4")])
        ring = np.ascontiguousarray(ring)
        ring_entries = len(ring)
        n = put - get if put >= get else ring_entries - get + put
        B = self._fault_B
        nb = (n + B - 1) // B
        total = nb * B
        d_ring = self.to_device(ring)
        d_ord, d_grp, d_bat, d_ent = self.alloc(total * 16), self.alloc(total * 16), self.alloc(nb * 32), self.alloc(total * 48)
        d_svc, d_runs, d_cnt = self.alloc(total * 96), self.alloc(max(run_cap, 1) * 8), self.alloc(8)
        devs = (d_ring, d_ord, d_grp, d_bat, d_ent, d_svc, d_runs, d_cnt)
        try:
            for p_, nbytes in ((d_ord, total * 16), (d_grp, total * 16), (d_ent, total * 48), (d_svc, total * 96)):
                _check(lib().uvmtier_memset_device(self._h, ctypes.c_void_p(p_), 0, ctypes.c_size_t(nbytes)), "memset")
            new_get = ctypes.c_uint32(0)
            _check(lib().uvmtier_access_counters_fetch_preprocess_device(
                self._h, P(d_ring), ctypes.c_uint32(ring_entries), ctypes.c_uint32(get), ctypes.c_uint32(put),
                ctypes.c_int(1 if clear_valid else 0), P(d_ord), P(d_grp), P(d_bat), P(d_ent), ctypes.byref(new_get)),
                "uvmtier_access_counters_fetch_preprocess_device")
            _check(lib().uvmtier_access_counters_service_device(
                self._h, ctypes.c_uint32(nb), P(d_ord), P(d_grp), P(d_bat), P(layout), ctypes.byref(dstate.ptrs), P(d_svc),
                P(d_runs), ctypes.c_uint32(run_cap), P(d_cnt)), "uvmtier_access_counters_service_device")
            self.synchronize()
            cnt = self.from_device(d_cnt, 2, "<u4")
            return dict(ordered=self.from_device(d_ord, total, onotif), groups=self.from_device(d_grp, total, GROUP_DT),
                        batch=self.from_device(d_bat, nb, BATCH_DT), entries=self.from_device(d_ent, total, np.dtype("V48")),
                        ring=self.from_device(d_ring, ring.shape, ring.dtype), n=n, get=new_get.value,
                        service=self.from_device(d_svc, total, SERVICE_GROUP_DT),
                        runs=self.from_device(d_runs, max(run_cap, 1), COPY_RUN_DT), n_runs=int(cnt[0]),
                        flags=np.zeros(total, "<u2"), state=dstate.download())
        finally:
            for p_ in devs:
                self.free(p_)

    # ---- (1g) non-replayable faults ----
    def service_non_replayable_faults(self, packets, layout, dstate, run_cap=1 << 20):
        """uvmtier_service_non_replayable_faults_device on a host array of packets (the shadow-buffer copy).  Downloads
        entries, per-fault flags, records, copy runs, the run count and the state."""
        from .blockstate import COPY_RUN_DT, SERVICE_GROUP_DT
        from .layouts import FAULT_ENTRY_DT
        n = len(packets)
        B = self._fault_B
        total = (n + B - 1) // B * B
        d_pk = self.to_device(np.ascontiguousarray(packets))
        d_ent, d_flags, d_svc = self.alloc(max(n, 1) * 48), self.alloc(max(total, 1) * 2), self.alloc(max(total, 1) * 96)
        d_runs, d_cnt = self.alloc(max(run_cap, 1) * 8), self.alloc(8)
        try:
            _check(lib().uvmtier_service_non_replayable_faults_device(self._h, P(d_pk), ctypes.c_uint32(n), P(layout),
                                                                      ctypes.byref(dstate.ptrs), P(d_ent), P(d_flags), P(d_svc), P(d_runs),
                                                                      ctypes.c_uint32(run_cap), P(d_cnt)),
                   "uvmtier_service_non_replayable_faults_device")
            self.synchronize()
            cnt = self.from_device(d_cnt, 2, "<u4")
            return dict(entries=self.from_device(d_ent, n, FAULT_ENTRY_DT), flags=self.from_device(d_flags, total, "<u2"),
                        service=self.from_device(d_svc, total, SERVICE_GROUP_DT), runs=self.from_device(d_runs, max(run_cap, 1), COPY_RUN_DT),
                        n_runs=int(cnt[0]), state=dstate.download())
        finally:
            for p_ in (d_pk, d_ent, d_flags, d_svc, d_runs, d_cnt):
                self.free(p_)

    # ---- (1c) make_resident ----
    def make_resident(self, requests, page_masks, prefetch_masks, layout, dstate, run_cap=1 << 18, execute=False):
        """uvmtier_make_resident_device on host arrays (requests: MAKE_RESIDENT_REQ_DT; masks: uint64 [n, 8] or
        None).  Downloads results, copy runs, the run count and the state; execute=True also runs the copy runs
        (uvmtier_copy_runs_device) before returning."""
        from .blockstate import COPY_RUN_DT, MAKE_RESIDENT_RESULT_DT
        n = len(requests)
        d_req = self.to_device(requests)
        d_pm = self.to_device(page_masks) if page_masks is not None else 0
        d_pf = self.to_device(prefetch_masks) if prefetch_masks is not None else 0
        d_res = self.alloc(max(n, 1) * 160); d_runs = self.alloc(max(run_cap, 1) * 8); d_cnt = self.alloc(8)
        try:
            _check(lib().uvmtier_memset_device(self._h, ctypes.c_void_p(d_cnt), 0, ctypes.c_size_t(8)), "memset")
            _check(lib().uvmtier_make_resident_device(self._h, ctypes.c_uint32(n), P(d_req), P(d_pm), P(d_pf), P(layout),
                                                      ctypes.byref(dstate.ptrs), P(d_res), P(d_runs), ctypes.c_uint32(run_cap),
                                                      P(d_cnt)), "uvmtier_make_resident_device")
            if execute:
                _check(lib().uvmtier_copy_runs_device(self._h, P(d_runs), P(d_cnt), ctypes.c_uint32(run_cap), ctypes.c_int(0)),
                       "uvmtier_copy_runs_device")
            self.synchronize()
            cnt = self.from_device(d_cnt, 2, "<u4")
            return dict(results=self.from_device(d_res, n, MAKE_RESIDENT_RESULT_DT),
                        runs=self.from_device(d_runs, max(run_cap, 1), COPY_RUN_DT), n_runs=int(cnt[0]), state=dstate.download())
        finally:
            for p in (d_req, d_pm, d_pf, d_res, d_runs, d_cnt):
                if p:
                    self.free(p)

    # ---- (3) tiers + page-list migration ----
    def tier_create(self, tier, nbytes, numa_node=-1):
        """uvmtier_tier_create -> TierInfo (device_ptr, host view as numpy uint8 for host tiers)."""
        info = TierInfo()
        _check(lib().uvmtier_tier_create(self._h, ctypes.c_uint32(tier), ctypes.c_uint64(nbytes), ctypes.c_int(numa_node),
                                         ctypes.byref(info)), "uvmtier_tier_create")
        return info

    def tier_map_peer(self, peer_index, peer_ctx):
        _check(lib().uvmtier_tier_map_peer(self._h, ctypes.c_uint32(peer_index), peer_ctx.handle), "uvmtier_tier_map_peer")

    def tier_destroy(self, tier):
        _check(lib().uvmtier_tier_destroy(self._h, ctypes.c_uint32(tier)), "uvmtier_tier_destroy")

    def copy_page_list(self, dst_tier, src_tier, dst_off, src_off, page_size, method=0):
        """uvmtier_copy_page_list (host lists of byte offsets, uint64)."""
        dst_off = np.ascontiguousarray(dst_off, "<u8"); src_off = np.ascontiguousarray(src_off, "<u8")
        _check(lib().uvmtier_copy_page_list(self._h, ctypes.c_uint32(dst_tier), ctypes.c_uint32(src_tier),
                                            ctypes.c_size_t(len(dst_off)), P(dst_off), P(src_off), ctypes.c_uint32(page_size),
                                            ctypes.c_int(method)), "uvmtier_copy_page_list")

    # ---- (3b) the fork's CXL P2P DMA surface ----
    def p2p_register_cxl_buffer(self, base_addr, size, cxl_version=2):
        """uvmtier_p2p_register_cxl_buffer -> (status, handle).  Does not raise: the status is what callers test."""
        h = ctypes.c_void_p()
        st = lib().uvmtier_p2p_register_cxl_buffer(self._h, ctypes.c_void_p(base_addr), ctypes.c_uint64(size),
                                                   ctypes.c_uint32(cxl_version), ctypes.byref(h))
        return int(st), h

    def p2p_unregister_cxl_buffer(self, handle):
        return int(lib().uvmtier_p2p_unregister_cxl_buffer(self._h, handle))

    def p2p_get_cxl_pages(self, handle, offset, size, cap=1 << 20):
        """-> (status, page addresses, page size)"""
        addrs = np.zeros(cap, "<u8")
        n, ps = ctypes.c_uint32(0), ctypes.c_uint32(0)
        st = lib().uvmtier_p2p_get_cxl_pages(handle, ctypes.c_uint64(offset), ctypes.c_uint64(size), P(addrs), ctypes.byref(n),
                                             ctypes.byref(ps))
        return int(st), addrs[:n.value].copy(), ps.value

    def p2p_cxl_dma_request(self, handle, gpu_offset, cxl_offset, size, flags=0):
        """-> (status, bytes moved)"""
        moved = ctypes.c_uint64(0)
        st = lib().uvmtier_p2p_cxl_dma_request(self._h, handle, ctypes.c_uint64(gpu_offset), ctypes.c_uint64(cxl_offset),
                                               ctypes.c_uint64(size), ctypes.c_uint32(flags), ctypes.byref(moved))
        return int(st), moved.value

    def checksum_pages(self, tier, off, page_size):
        off = np.ascontiguousarray(off, "<u8")
        d_off = self.to_device(off); d_out = self.alloc(len(off) * 8)
        try:
            _check(lib().uvmtier_checksum_pages_device(self._h, ctypes.c_uint32(tier), ctypes.c_size_t(len(off)), P(d_off),
                                                       ctypes.c_uint32(page_size), P(d_out)), "uvmtier_checksum_pages_device")
            self.synchronize()
            return self.from_device(d_out, len(off), "<u8")
        finally:
            self.free(d_off); self.free(d_out)

    def tier_memcpy_peak_gbs(self, dst_tier, src_tier, nbytes, reps=5):
        ms = ctypes.c_float(0)
        _check(lib().uvmtier_tier_memcpy_peak(self._h, ctypes.c_uint32(dst_tier), ctypes.c_uint32(src_tier),
                                              ctypes.c_uint64(nbytes), ctypes.c_int(reps), ctypes.byref(ms)), "tier_memcpy_peak")
        return nbytes / (ms.value * 1e-3) / 1e9

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


class DeviceBlockState:
    """uvmtier_block_state_t in device memory (one table per context / GPU)."""

    def __init__(self, ctx, layout):
        from .blockstate import BlockStatePtrs
        self.ctx = ctx
        self.layout = layout
        self.n = int(layout["n_spaces"][0]) * int(layout["blocks_per_space"][0])
        self.ptrs = BlockStatePtrs()
        _check(lib().uvmtier_block_state_alloc(ctx.handle, P(layout), ctypes.byref(self.ptrs)), "uvmtier_block_state_alloc")

    def enable_read_duplication(self):
        """uvmtier_block_state_enable_read_duplication: adds the (zeroed) read_duplicated mask array."""
        if not self.ptrs.read_duplicated:
            _check(lib().uvmtier_block_state_enable_read_duplication(self.ctx.handle, P(self.layout), ctypes.byref(self.ptrs)),
                   "uvmtier_block_state_enable_read_duplication")
        return self

    def add_peer(self, gpu_id, direct=True):
        """uvmtier_block_state_add_peer: another GPU that holds pages of these va_blocks (ascending gpu id)."""
        _check(lib().uvmtier_block_state_add_peer(self.ctx.handle, P(self.layout), ctypes.byref(self.ptrs), ctypes.c_uint32(gpu_id),
                                                  ctypes.c_int(1 if direct else 0)), "uvmtier_block_state_add_peer")
        return self

    def _fields(self):
        p = self.ptrs
        f = (("gpu_resident", p.gpu_resident), ("cpu_resident0", p.cpu_resident[0]), ("cpu_resident1", p.cpu_resident[1]),
             ("gpu_pte0", p.gpu_pte[0]), ("gpu_pte1", p.gpu_pte[1]), ("gpu_pte2", p.gpu_pte[2]),
             ("prefetch_info", p.prefetch_info), ("logical_prot", p.logical_prot), ("preferred_location", p.preferred_location),
             ("gpu_evicted", p.gpu_evicted))
        return f + ((("read_duplicated", p.read_duplicated),) if p.read_duplicated else ())

    def upload(self, host_state):
        assert host_state.n == self.n
        if host_state.read_duplicated is not None:
            self.enable_read_duplication()
        for name, dptr in self._fields():
            a = np.ascontiguousarray(getattr(host_state, name))
            _check(lib().uvmtier_memcpy_h2d(self.ctx.handle, ctypes.c_void_p(dptr), P(a), ctypes.c_size_t(a.nbytes)), "h2d")
        for k, (gid, direct, mask) in enumerate(host_state.peers):
            if k >= self.ptrs.n_peers:
                self.add_peer(gid, direct)
            assert self.ptrs.peer_gpu_id[k] == gid and bool((self.ptrs.peer_direct >> k) & 1) == bool(direct), "peer set differs"
            a = np.ascontiguousarray(mask)
            _check(lib().uvmtier_memcpy_h2d(self.ctx.handle, ctypes.c_void_p(self.ptrs.peer_resident[k]), P(a),
                                            ctypes.c_size_t(a.nbytes)), "h2d")

    def download(self):
        from .blockstate import HostBlockState
        h = HostBlockState(self.n)
        if self.ptrs.read_duplicated:
            h.enable_read_duplication()
        for name, dptr in self._fields():
            a = getattr(h, name)
            _check(lib().uvmtier_memcpy_d2h(self.ctx.handle, P(a), ctypes.c_void_p(dptr), ctypes.c_size_t(a.nbytes)), "d2h")
        for k in range(self.ptrs.n_peers):
            h.add_peer(self.ptrs.peer_gpu_id[k], bool((self.ptrs.peer_direct >> k) & 1))
            a = h.peers[k][2]
            _check(lib().uvmtier_memcpy_d2h(self.ctx.handle, P(a), ctypes.c_void_p(self.ptrs.peer_resident[k]),
                                            ctypes.c_size_t(a.nbytes)), "d2h")
        return h

    def free(self):
        if self.ptrs.gpu_resident:
            lib().uvmtier_block_state_free(self.ctx.handle, ctypes.byref(self.ptrs))
