"""The C-ABI library loads and exports every symbol include/*.h declares (no compute calls)."""
import ctypes
import glob
import os
import re

import uvmtier

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    names = set()
    for h in glob.glob(os.path.join(ROOT, "include", "*.h")):
        text = open(h).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        for m in re.finditer(r"\b(uvmtier_[a-z0-9_]+)\s*\(", text):
            names.add(m.group(1))
    return sorted(names)


def test_library_exports_every_declared_symbol():
    lib = uvmtier.lib()
    syms = _declared_symbols()
    assert len(syms) >= 10
    missing = [s for s in syms if not hasattr(lib, s)]
    assert not missing, "declared in include/ but not exported: %s" % missing


def test_ops_table_points_at_the_entry_points():
    """uvmtier_gpu_ops (the table a reference-side harness dispatches through) holds the exported functions, in the
    order include/uvmtier.h declares its slots."""
    lib = uvmtier.lib()
    slots = ("uvmtier_fetch_preprocess_fault_batches", "uvmtier_fetch_preprocess_fault_batches_device",
             "uvmtier_service_fault_batches_device", "uvmtier_perf_prefetch_get_hint_va_block", "uvmtier_make_resident_device",
             "uvmtier_copy_runs_device", "uvmtier_copy_page_list", "uvmtier_service_replayable_faults_streamed",
             "uvmtier_access_counters_fetch_preprocess_device", "uvmtier_access_counters_service_device")
    table = (ctypes.c_void_p * len(slots)).in_dll(lib, "uvmtier_gpu_ops")
    for k, name in enumerate(slots):
        assert table[k] == ctypes.cast(getattr(lib, name), ctypes.c_void_p).value, name


def test_version_and_defaults():
    assert b"sm_100a" in uvmtier.lib().uvmtier_version()
    cfg = uvmtier.default_config()
    assert (cfg.uvm_perf_prefetch_enable, cfg.uvm_perf_prefetch_threshold, cfg.uvm_perf_prefetch_min_faults) == (1, 51, 1)
    assert cfg.uvm_perf_fault_batch_count == 256 and cfg.uvm_perf_fault_coalesce == 1


def test_struct_layouts_match_header():
    import uvmtier_testlib as T
    assert T.PREFETCH_BLOCK_DT.itemsize == 24 and T.PREFETCH_INFO_DT.itemsize == 4
    assert ctypes.sizeof(uvmtier.Config) == 64


def test_no_gpu_fails_loudly():
    """Without a CUDA device the product must refuse, not fall back to the CPU."""
    import pytest
    h = ctypes.c_void_p()
    st = uvmtier.lib().uvmtier_init(0, None, ctypes.byref(h))
    if st == 0:
        uvmtier.lib().uvmtier_deinit(h)
        pytest.skip("a GPU is present")
    assert st != 0 and not h.value
