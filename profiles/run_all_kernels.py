"""Runs every hot kernel of the library once or twice at a meaningful size, for one ncu capture:
    ncu --set full --clock-control none --import-source on -k regex:'^k_' -o gpurun_out/all python profiles/run_all_kernels.py
(run it without ncu first).  Sizes: the 1M-entry / 64 GB step of config B for the fault path, 32768 and 1M va_blocks
for the prefetch tree, 256 MB page lists for the copy kernel (host DRAM <-> HBM and HBM -> HBM)."""
import ctypes
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "open-gpu-kernel-modules_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402
import uvmtier  # noqa: E402
from uvmtier import P, TIER_DRAM, TIER_HBM, TIER_PEER0, blockstate as BSL, layouts as LY  # noqa: E402

lib = uvmtier.lib()
ctx = uvmtier.Context(0)

# ---- fault path: one step of config B ----
print(bench.other_trace_numbers(ctx, lib, P, "uniform", steps=1, warmup=1)["ms_per_step"])

# ---- prefetch tree: 32768 and 1M va_blocks ----
import residentlib as R  # noqa: E402
import uvmtier_testlib as T  # noqa: E402
for n in (32768, 1 << 20):
    rng = np.random.default_rng(5)
    res = T.masks_from_bool(rng.random((n, 512)) < 0.4)
    flt = T.masks_from_bool(rng.random((n, 512)) < 0.02)
    blk = np.zeros(n, LY.PREFETCH_BLOCK_DT)
    blk["start"] = 0x7F0000000000 + np.arange(n, dtype=np.uint64) * np.uint64(2 << 20)
    blk["big_page_size"] = 65536; blk["num_pages"] = 512; blk["faulted_first"] = 0; blk["faulted_outer"] = 512
    blk["new_residency"] = 1; blk["preferred_location"] = 257; blk["flags"] = 0x01 | 0x08
    info = np.zeros(n, LY.PREFETCH_INFO_DT)
    info["last_migration_proc_id"] = 257
    out = ctx.perf_prefetch_get_hint_va_block(blk, res, flt, None, None, info)
    print("prefetch", n, int(np.unpackbits(out[0].view(np.uint8)).sum()))

# ---- make_resident: 32768 requests ----
import servicelib as S  # noqa: E402
n = 32768
lay = dict(va_base=0x7F0000000000, blocks_per_space=n, n_spaces=1, gpu_id=1)
layout = S.make_layout(lay)
st0 = S.random_state(n, 3)
req = np.zeros(n, BSL.MAKE_RESIDENT_REQ_DT)
if True:
    req["slot"] = np.arange(n); req["region_first"] = 0; req["region_outer"] = 512; req["dest_tier"] = 0; req["cause"] = 6
    ds = ctx.block_state_alloc(layout); ds.upload(st0)
    r = ctx.make_resident(req, None, None, layout, ds, run_cap=1 << 23)
    print("make_resident done", type(r).__name__)
    ds.free()

# ---- copy kernel: host DRAM <-> HBM, HBM -> HBM (second arena registered as a peer tier) ----
nbytes = 256 << 20
info = {t: ctx.tier_create(t, nbytes) for t in (TIER_HBM, TIER_DRAM)}
d_peer = ctx.alloc(nbytes)
uvmtier._check(lib.uvmtier_tier_register(ctx.handle, ctypes.c_uint32(TIER_PEER0), ctypes.c_void_p(d_peer), None, ctypes.c_uint64(nbytes)), "reg")
npg = nbytes // 4096
lin = np.arange(npg, dtype=np.uint64) * np.uint64(4096)
perm = np.random.default_rng(7).permutation(npg).astype(np.uint64) * np.uint64(4096)
for dst, src in ((TIER_HBM, TIER_DRAM), (TIER_DRAM, TIER_HBM), (TIER_PEER0, TIER_HBM)):
    ctx.copy_page_list(dst, src, perm, lin, 4096, 0)
print("copies done")

# ---- thrashing detector and access counters at the bench's sizes ----
print(bench.thrashing_numbers(ctx, lib, P, 0)["ms"])
print(bench.access_counter_numbers(ctx, lib, P, 0)["service_ms"])
print(bench.nonreplayable_numbers(ctx, lib, P, 0))

# ---- round 2: thrashing hints (queries over 4096 va_blocks), the tiered loop on a small table, the big-batch K1 ----
import thrashlib as TH  # noqa: E402
from uvmtier import tracegen as TG  # noqa: E402
n_slots = 4096
cfg = TH.make_cfg("pcie", gpus=(1,), pin_threshold=2)
layout = BSL.make_layout(dict(va_base=TG.VA_BASE, blocks_per_space=n_slots, n_spaces=1))
t = ctx.thrashing_create(layout)
ctx.thrashing_enable_hints(t, cfg)
rng = np.random.default_rng(3)
now = 10000
for rnd in range(6):
    ev = np.zeros(n_slots * 4, TH.EVENT_DT)
    ev["slot"] = np.repeat(np.arange(n_slots), 4)
    ev["first_page"] = np.tile(np.arange(4) * 16, n_slots); ev["num_pages"] = 16
    ev["processor"] = rnd & 1; ev["time_stamp"] = now
    ctx.thrashing_note_events(t, ev)
    now += 50000
q = np.zeros(n_slots * 8, TH.QUERY_DT)
q["slot"] = np.repeat(np.arange(n_slots), 8); q["page"] = np.tile(np.arange(8) * 2, n_slots)
q["requester"] = 1; q["resident_processors"] = 1; q["preferred_location"] = TH.NO_PREF
for rnd in range(4):
    q["time_stamp"] = now
    q["requester"] = rnd & 1
    h, c = ctx.thrashing_get_hints(t, q)
    now += 20000
    print("hints round", rnd, c[3:6])
print("unpinned", ctx.thrashing_unpin_expired(t, now + 10 ** 9)[1])
ctx.thrashing_destroy(t)
print(bench.oversubscription_numbers(ctx, lib, P, 0, blocks=1024, hbm_chunks=4096 // 8, dram_blocks=384)["uniform"]["entries_per_s"])
pk, tbl, lay = TG.fault_trace(1 << 18, 1, "uniform", 4)
gp = TG.gpu_params(4096)
ctx.fault_buffer_init(gp, tbl)
ring = np.zeros(((1 << 18) + 1, 8), np.uint32); ring[:1 << 18] = pk
r = ctx.fetch_preprocess_fault_batches(ring, 0, 1 << 18)
print("big batches", len(r["batch"]))
ctx.close()
