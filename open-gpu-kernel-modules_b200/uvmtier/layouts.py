"""numpy dtypes mirroring the structs of include/uvmtier.h (byte-identical layouts)."""
import numpy as np

PREFETCH_BLOCK_DT = np.dtype([("start", "<u8"), ("big_page_size", "<u4"), ("num_pages", "<u2"), ("faulted_first", "<u2"),
                              ("faulted_outer", "<u2"), ("new_residency", "<u2"), ("preferred_location", "<u2"),
                              ("flags", "<u2")])
PREFETCH_INFO_DT = np.dtype([("last_migration_proc_id", "<u2"), ("fault_migrations_to_last_proc", "<u2")])
FAULT_ENTRY_DT = np.dtype([
    ("fault_address", "<u8"), ("timestamp", "<u8"), ("instance_ptr_address", "<u8"),
    ("client_id", "<u2"), ("mmu_engine_id", "<u2"), ("utlb_id", "<u2"), ("num_instances", "<u2"),
    ("instance_ptr_aperture", "u1"), ("fault_type", "u1"), ("fault_access_type", "u1"), ("client_type", "u1"),
    ("mmu_engine_type", "u1"), ("gpc_id", "u1"), ("ve_id", "u1"), ("flags", "u1"),
    ("access_type_mask", "u1"), ("pad", "u1", 7)])
BATCH_DT = np.dtype([("num_cached_faults", "<u4"), ("num_coalesced_faults", "<u4"), ("is_single_instance_ptr", "<u4"),
                     ("max_utlb_id", "<u4"), ("status", "<u4"), ("num_duplicate_faults", "<u4"),
                     ("num_invalid_prefetch_faults", "<u4"), ("num_block_groups", "<u4")])
ORDERED_DT = np.dtype([("fault_address", "<u8"), ("fault_index", "<u2"), ("ctx", "<u2"), ("num_instances", "<u2"),
                       ("access_type_mask", "u1"), ("fault_access_type", "u1")])
GROUP_DT = np.dtype([("block_start", "<u8"), ("ctx", "<u2"), ("first_fault_index", "<u2"), ("block_faults", "<u2"),
                     ("flags", "<u2")])
assert (PREFETCH_BLOCK_DT.itemsize, PREFETCH_INFO_DT.itemsize, FAULT_ENTRY_DT.itemsize, BATCH_DT.itemsize,
        ORDERED_DT.itemsize, GROUP_DT.itemsize) == (24, 4, 48, 32, 16, 16)
