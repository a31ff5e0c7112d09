"""Host-side containers for the per-slot va_block state table and the servicing-pass records
(numpy mirrors of uvmtier_va_layout_t / uvmtier_block_state_t / uvmtier_service_group_t / uvmtier_copy_run_t)."""
import ctypes

import numpy as np

VA_LAYOUT_DT = np.dtype([("va_base", "<u8"), ("blocks_per_space", "<u4"), ("n_spaces", "<u4"), ("gpu_id", "<u4"),
                         ("reserved", "<u4")])
SERVICE_GROUP_DT = np.dtype([("slot", "<u4"), ("region_first", "<u2"), ("region_outer", "<u2"), ("page_fault_count", "<u2"),
                             ("prefetch_count", "<u2"), ("migrated", "<u2", 2), ("first_touch", "<u2"), ("hint_residency", "<u2"),
                             ("run_offset", "<u4"), ("run_count", "<u2"), ("batch", "<u2"), ("group", "<u4"),
                             ("new_residency", "<u8", 8)])
COPY_RUN_DT = np.dtype([("slot", "<u4"), ("first_page", "<u2"), ("bits", "<u2")])   # bits: num_pages:10 src:3 prefetch:1
PREFETCH_INFO_DT = np.dtype([("last_migration_proc_id", "<u2"), ("fault_migrations_to_last_proc", "<u2")])
assert (VA_LAYOUT_DT.itemsize, SERVICE_GROUP_DT.itemsize, COPY_RUN_DT.itemsize) == (24, 96, 8)
# uvmtier_make_resident_req_t / _result_t (include/uvmtier.h section 1c)
MAKE_RESIDENT_REQ_DT = np.dtype([("slot", "<u4"), ("region_first", "<u2"), ("region_outer", "<u2"), ("dest_tier", "u1"),
                                 ("cause", "u1"), ("flags", "<u2"), ("mask_row", "<u4")])
MAKE_RESIDENT_RESULT_DT = np.dtype([("slot", "<u4"), ("migrated", "<u2", (3,)), ("first_touch", "<u2"), ("run_offset", "<u4"),
                                    ("run_count", "<u2"), ("unmapped", "<u2"), ("migrated_peer", "<u2", (4,)), ("reserved", "<u4"),
                                    ("pages_migrated", "<u8", (8,)), ("pages_changed_residency", "<u8", (8,))])
assert (MAKE_RESIDENT_REQ_DT.itemsize, MAKE_RESIDENT_RESULT_DT.itemsize) == (16, 160)
MR_HAS_PAGE_MASK, MR_HAS_PREFETCH_MASK, MR_EVICT_GPU_PAGES, MR_READ_DUPLICATE = 1, 2, 4, 8
CAUSE_REPLAYABLE_FAULT, CAUSE_EVICTION, CAUSE_API_MIGRATE = 0, 4, 6
TIER_HBM, TIER_DRAM, TIER_CXL = 0, 1, 2

STATE_FIELDS = ("gpu_resident", "cpu_resident0", "cpu_resident1", "gpu_pte0", "gpu_pte1", "gpu_pte2", "prefetch_info",
                "logical_prot", "preferred_location", "gpu_evicted", "read_duplicated")


class BlockStatePtrs(ctypes.Structure):
    _fields_ = [("gpu_resident", ctypes.c_void_p), ("cpu_resident", ctypes.c_void_p * 2), ("gpu_pte", ctypes.c_void_p * 3),
                ("prefetch_info", ctypes.c_void_p), ("logical_prot", ctypes.c_void_p), ("preferred_location", ctypes.c_void_p),
                ("gpu_evicted", ctypes.c_void_p), ("read_duplicated", ctypes.c_void_p),
                ("peer_resident", ctypes.c_void_p * 4), ("peer_gpu_id", ctypes.c_uint16 * 4), ("n_peers", ctypes.c_uint32),
                ("peer_direct", ctypes.c_uint32)]

MAX_PEERS = 4


class HostBlockState:
    """Host copy of the per-slot va_block state (numpy)."""

    def __init__(self, n_slots):
        z = lambda: np.zeros((n_slots, 8), "<u8")
        self.n = n_slots
        self.gpu_resident, self.cpu_resident0, self.cpu_resident1 = z(), z(), z()
        self.gpu_pte0, self.gpu_pte1, self.gpu_pte2 = z(), z(), z()
        self.gpu_evicted = z()
        self.read_duplicated = None                      # read duplication is off until enable_read_duplication()
        self.prefetch_info = np.zeros(n_slots, PREFETCH_INFO_DT)
        self.prefetch_info["last_migration_proc_id"] = 257
        self.logical_prot = np.full(n_slots, 3, "u1")
        self.preferred_location = np.full(n_slots, 257, "<u2")
        # other GPUs holding pages of these va_blocks: (gpu id, direct copy access, resident masks), ascending gpu id
        self.peers = []

    def add_peer(self, gpu_id, direct=True, resident=None):
        assert len(self.peers) < 4 and all(p[0] < gpu_id for p in self.peers), "peers in ascending gpu id, at most 4"
        self.peers.append([int(gpu_id), bool(direct), np.zeros((self.n, 8), "<u8") if resident is None else resident.copy()])
        return self

    def enable_read_duplication(self):
        if self.read_duplicated is None:
            self.read_duplicated = np.zeros((self.n, 8), "<u8")
        return self

    def copy(self):
        c = HostBlockState(self.n)
        for f in STATE_FIELDS:
            v = getattr(self, f)
            setattr(c, f, None if v is None else v.copy())
        c.peers = [[g, d, m.copy()] for g, d, m in self.peers]
        return c

    def ptrs(self):
        s = BlockStatePtrs()
        s.gpu_resident = self.gpu_resident.ctypes.data
        s.cpu_resident[0] = self.cpu_resident0.ctypes.data
        s.cpu_resident[1] = self.cpu_resident1.ctypes.data
        s.gpu_pte[0] = self.gpu_pte0.ctypes.data
        s.gpu_pte[1] = self.gpu_pte1.ctypes.data
        s.gpu_pte[2] = self.gpu_pte2.ctypes.data
        s.prefetch_info = self.prefetch_info.ctypes.data
        s.logical_prot = self.logical_prot.ctypes.data
        s.preferred_location = self.preferred_location.ctypes.data
        s.gpu_evicted = self.gpu_evicted.ctypes.data
        s.read_duplicated = None if self.read_duplicated is None else self.read_duplicated.ctypes.data
        s.n_peers = len(self.peers)
        s.peer_direct = 0
        for k, (g, d, m) in enumerate(self.peers):
            s.peer_resident[k] = m.ctypes.data
            s.peer_gpu_id[k] = g
            s.peer_direct |= (1 << k) if d else 0
        return s

    def assert_equal(self, other, what=""):
        for f in STATE_FIELDS:
            a, b = getattr(self, f), getattr(other, f)
            if a is None or b is None:
                assert a is None and b is None, "%s state.%s present on one side only" % (what, f)
                continue
            if not np.array_equal(a, b):
                bad = np.nonzero((a != b).reshape(self.n, -1).any(axis=1))[0] if a.dtype.names is None else np.nonzero(a != b)[0]
                raise AssertionError("%s state.%s differs at slots %s" % (what, f, bad[:8]))
        assert len(self.peers) == len(other.peers), what + " number of peers"
        for k, ((g, d, a), (g2, d2, b)) in enumerate(zip(self.peers, other.peers)):
            assert (g, d) == (g2, d2), what + " peer ids"
            if not np.array_equal(a, b):
                raise AssertionError("%s state.peer_resident[%d] differs at slots %s" % (what, k, np.nonzero((a != b).any(axis=1))[0][:8]))


def make_layout(lay, gpu_id=1):
    v = np.zeros(1, VA_LAYOUT_DT)
    v["va_base"] = lay["va_base"]
    v["blocks_per_space"] = lay["blocks_per_space"]
    v["n_spaces"] = lay["n_spaces"]
    v["gpu_id"] = gpu_id
    return v
