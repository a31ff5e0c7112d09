"""Shared helpers for the test-suite: library loaders, numpy struct layouts, input generators.

Three implementations are driven with the same numpy arrays:
  * product   : open-gpu-kernel-modules_b200/libuvmtier.so through its C ABI (GPU only)
  * oracle    : oracle/liboracle.so        -- CPU restatement (test infrastructure)
  * reference : oracle/_ref/libref_*.so    -- the reference's own C compiled against shims
  * simt      : tests/simt/libsimt_kernels.so -- the product's device source under a CPU SIMT
                emulator (logic check without a GPU; never a fallback)
"""
import ctypes
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "open-gpu-kernel-modules_b200")
if PKG not in sys.path:
    sys.path.insert(0, PKG)

ID_INVALID = 257
PF_HAS_RESIDENT_MASK = 0x01
PF_BLOCK_RESIDENT_EMPTY = 0x02
PF_HAS_THRASHING = 0x04
PF_GPU_VA_SPACE_REGISTERED = 0x08

PREFETCH_BLOCK_DT = np.dtype(
    [
        ("start", "<u8"),
        ("big_page_size", "<u4"),
        ("num_pages", "<u2"),
        ("faulted_first", "<u2"),
        ("faulted_outer", "<u2"),
        ("new_residency", "<u2"),
        ("preferred_location", "<u2"),
        ("flags", "<u2"),
    ],
    align=False,
)
assert PREFETCH_BLOCK_DT.itemsize == 24
PREFETCH_INFO_DT = np.dtype([("last_migration_proc_id", "<u2"), ("fault_migrations_to_last_proc", "<u2")])


def P(a):
    """numpy array (or None) -> void*"""
    if a is None:
        return ctypes.c_void_p(0)
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(ctypes.c_void_p)


def _load(path):
    if not os.path.exists(path):
        return None
    return ctypes.CDLL(path)


def load_oracle():
    lib = _load(os.path.join(ROOT, "oracle", "liboracle.so"))
    if lib is None:
        raise RuntimeError("oracle/liboracle.so missing: run `python -c 'import __graft_entry__ as g; g.build()'`")
    return lib


def load_ref(leg):
    """oracle/_ref/libref_<leg>.so or None when it has not been built (it is built from /root/reference)."""
    return _load(os.path.join(ROOT, "oracle", "_ref", "libref_%s.so" % leg))


def load_simt():
    lib = _load(os.path.join(ROOT, "tests", "simt", "libsimt_kernels.so"))
    if lib is None:
        raise RuntimeError("tests/simt/libsimt_kernels.so missing: run build()")
    return lib


# ----------------------------------------------------------------------------- mask helpers
def masks_from_bool(b):
    """bool array [n,512] -> uint64 [n,8] little-endian page masks"""
    b = np.ascontiguousarray(b, dtype=np.uint8)
    packed = np.packbits(b, axis=1, bitorder="little")  # [n,64] bytes
    return np.ascontiguousarray(packed).view("<u8").reshape(b.shape[0], 8)


def bool_from_masks(m):
    m = np.ascontiguousarray(m, dtype="<u8")
    return np.unpackbits(m.view(np.uint8).reshape(m.shape[0], 64), axis=1, bitorder="little").astype(bool)


# ----------------------------------------------------------------------------- prefetch inputs
def gen_prefetch_blocks(n, seed, *, aligned=True, thrashing=False, dst="gpu", first_touch_frac=0.05,
                        density=None, full_blocks=True):
    """Random but valid inputs of uvm_perf_prefetch_get_hint_va_block for n independent blocks."""
    rng = np.random.default_rng(seed)
    blk = np.zeros(n, PREFETCH_BLOCK_DT)
    if full_blocks:
        num_pages = np.full(n, 512)
        start_page = np.zeros(n, np.int64)
    else:
        num_pages = rng.integers(1, 513, n)
        if aligned:
            start_page = np.zeros(n, np.int64)
        else:
            start_page = np.array([rng.integers(0, 512 - p + 1) for p in num_pages], np.int64)
    base = (0x7F0000000000 >> 21) + np.arange(n, dtype=np.int64)
    blk["start"] = (base << 21) + (start_page << 12)
    blk["num_pages"] = num_pages
    blk["big_page_size"] = 65536
    idx = np.arange(512)[None, :]
    inblk = idx < num_pages[:, None]
    p = rng.choice([0.0, 0.02, 0.25, 0.6, 0.9], n) if density is None else np.full(n, density)
    resident = (rng.random((n, 512)) < p[:, None]) & inblk
    # run-structured residency for a third of the blocks
    runs = rng.random(n) < 0.33
    for i in np.nonzero(runs)[0]:
        a = rng.integers(0, num_pages[i])
        b = rng.integers(a, num_pages[i]) + 1
        resident[i] = False
        resident[i, a:b] = True
    nf = rng.integers(1, 9, n)
    faulted = np.zeros((n, 512), bool)
    for i in range(n):
        pages = rng.integers(0, num_pages[i], nf[i])
        faulted[i, pages] = True
    faulted &= ~resident if dst == "gpu" else True
    empty = ~faulted.any(axis=1)
    for i in np.nonzero(empty)[0]:
        faulted[i, rng.integers(0, num_pages[i])] = True
        resident[i] &= ~faulted[i]
    first = faulted.argmax(axis=1)
    last = 511 - faulted[:, ::-1].argmax(axis=1)
    blk["faulted_first"] = first
    blk["faulted_outer"] = last + 1
    flags = np.full(n, PF_HAS_RESIDENT_MASK | PF_GPU_VA_SPACE_REGISTERED, np.uint16)
    ft = rng.random(n) < first_touch_frac
    resident[ft] = False
    flags[ft] |= PF_BLOCK_RESIDENT_EMPTY
    nores = (rng.random(n) < 0.05) & ~ft
    if dst == "gpu":
        flags[nores] &= ~np.uint16(PF_HAS_RESIDENT_MASK)
    blk["new_residency"] = 0 if dst == "cpu" else rng.integers(1, 9, n)
    pref = np.full(n, ID_INVALID)
    pick = rng.random(n) < 0.5
    pref[pick] = blk["new_residency"][pick]
    blk["preferred_location"] = pref
    thr = None
    if thrashing:
        thr_b = (rng.random((n, 512)) < 0.01) & inblk
        has = rng.random(n) < 0.7
        thr_b[~has] = False
        flags[has] |= PF_HAS_THRASHING
        thr = masks_from_bool(thr_b)
    blk["flags"] = flags
    info = np.zeros(n, PREFETCH_INFO_DT)
    info["last_migration_proc_id"] = np.where(rng.random(n) < 0.5, blk["new_residency"], ID_INVALID)
    info["fault_migrations_to_last_proc"] = rng.integers(0, 4, n)
    cpu_read = masks_from_bool((rng.random((n, 512)) < 0.5) & inblk) if dst == "cpu" else None
    return dict(blocks=blk, resident=masks_from_bool(resident), faulted=masks_from_bool(faulted), thrashing=thr,
                cpu_read=cpu_read, info=info)


def run_prefetch_get_hint(lib, prefix, inp, init=(1, 51, 1)):
    """Run <prefix>_prefetch_get_hint_va_block of a CPU checker library. Returns (mask, residency, info)."""
    n = len(inp["blocks"])
    getattr(lib, prefix + "_prefetch_init")(*[ctypes.c_uint(v) for v in init])
    out = np.zeros((n, 8), "<u8")
    res = np.zeros(n, "<u4")
    info = inp["info"].copy()
    getattr(lib, prefix + "_prefetch_get_hint_va_block")(
        ctypes.c_size_t(n), P(inp["blocks"]), P(inp["resident"]), P(inp["faulted"]), P(inp["thrashing"]),
        P(inp["cpu_read"]), P(info), P(out), P(res))
    return out, res, info


def run_prefetch_get_hint_simt(lib, inp, init=(1, 51, 1), grid=2):
    n = len(inp["blocks"])
    out = np.zeros((n, 8), "<u8")
    res = np.zeros(n, "<u4")
    info = inp["info"].copy()
    # same tunable validation as uvm_perf_prefetch_init
    en, thr, mf = init
    thr = thr if thr <= 100 else 51
    mf = mf if 1 <= mf <= 20 else 1
    lib.simt_prefetch_get_hint(ctypes.c_size_t(n), P(inp["blocks"]), P(inp["resident"]), P(inp["faulted"]),
                               P(inp["thrashing"]), P(inp["cpu_read"]), P(info), P(out), P(res),
                               ctypes.c_uint(1 if en else 0), ctypes.c_uint(thr), ctypes.c_uint(mf), ctypes.c_uint(grid))
    return out, res, info


def gen_compute_ats(n, seed):
    rng = np.random.default_rng(seed)
    mf = rng.integers(0, 400, n)
    mo = np.array([rng.integers(a + 1, 513) for a in mf])
    idx = np.arange(512)[None, :]
    inreg = (idx >= mf[:, None]) & (idx < mo[:, None])
    p = rng.choice([0.0, 0.25, 0.6, 0.9], n)
    resid = rng.random((n, 512)) < p[:, None]          # may have bits outside the window, like hmm_range_fault output
    faulted = np.zeros((n, 512), bool)
    for i in range(n):
        k = rng.integers(1, 6)
        faulted[i, rng.integers(mf[i], mo[i], k)] = True
    first = faulted.argmax(axis=1)
    last = 511 - faulted[:, ::-1].argmax(axis=1)
    fr = np.stack([first, last + 1], 1).astype("<u2")
    mr = np.stack([mf, mo], 1).astype("<u2")
    return dict(faulted=masks_from_bool(faulted), residency=masks_from_bool(resid), fregion=np.ascontiguousarray(fr),
                mregion=np.ascontiguousarray(mr))


def run_compute_ats(lib, prefix, inp, with_residency=True, init=(1, 51, 1)):
    n = len(inp["faulted"])
    getattr(lib, prefix + "_prefetch_init")(*[ctypes.c_uint(v) for v in init])
    out = np.zeros((n, 8), "<u8")
    getattr(lib, prefix + "_prefetch_compute_ats")(ctypes.c_size_t(n), P(inp["faulted"]), P(inp["fregion"]),
                                                   P(inp["mregion"]), P(inp["residency"] if with_residency else None),
                                                   P(out))
    return out


def run_compute_ats_simt(lib, inp, with_residency=True, init=(1, 51, 1), grid=2):
    n = len(inp["faulted"])
    out = np.zeros((n, 8), "<u8")
    en, thr, mf = init
    thr = thr if thr <= 100 else 51
    lib.simt_prefetch_compute_ats(ctypes.c_size_t(n), P(inp["faulted"]), P(inp["fregion"]), P(inp["mregion"]),
                                  P(inp["residency"] if with_residency else None), P(out), ctypes.c_uint(1 if en else 0),
                                  ctypes.c_uint(thr), ctypes.c_uint(mf), ctypes.c_uint(grid))
    return out
