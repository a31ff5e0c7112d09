"""One launch (or a handful) of each kernel for an ncu --set full capture that stays short:
    ncu --set full --clock-control none --import-source on -k regex:'^k_' -c 110 -o /tmp/new python profiles/run_new_kernels.py
(run it without ncu first).  The first nine launches are one device-resident step of config B (K1 + the servicing pass:
their dram bytes are the `traffic` of bench.py's roofline, profiles/ncu_traffic.json); then this round's new kernels:
thrashing events / hints / expiry, one round of the tiered loop, the big-batch K1, non-replayable faults."""
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
from uvmtier import P, blockstate as BSL, tracegen as TG  # noqa: E402
import thrashlib as TH  # noqa: E402

lib = uvmtier.lib()
ctx = uvmtier.Context(0)
print(bench.other_trace_numbers(ctx, lib, P, "uniform", steps=1, warmup=0)["ms_per_step"])

n_slots = 4096
cfg = TH.make_cfg("pcie", gpus=(1,), pin_threshold=2)
layout = BSL.make_layout(dict(va_base=TG.VA_BASE, blocks_per_space=n_slots, n_spaces=1))
t = ctx.thrashing_create(layout)
ctx.thrashing_enable_hints(t, cfg)
now = 10000
for rnd in range(5):
    ev = np.zeros(n_slots * 4, TH.EVENT_DT)
    ev["slot"] = np.repeat(np.arange(n_slots), 4)
    ev["first_page"] = np.tile(np.arange(4) * 16, n_slots); ev["num_pages"] = 16
    ev["processor"] = rnd & 1; ev["time_stamp"] = now
    ctx.thrashing_note_events(t, ev)
    now += 50000
q = np.zeros(n_slots * 8, TH.QUERY_DT)
q["slot"] = np.repeat(np.arange(n_slots), 8); q["page"] = np.tile(np.arange(8) * 2, n_slots)
q["resident_processors"] = 1; q["preferred_location"] = TH.NO_PREF
for rnd in range(2):
    q["time_stamp"] = now; q["requester"] = 1 - (rnd & 1)
    h, c = ctx.thrashing_get_hints(t, q)
    now += 20000
    print("hints", c[3:6])
print("unpinned", ctx.thrashing_unpin_expired(t, now + 10 ** 9)[1])
ctx.thrashing_destroy(t)

pk, tbl, lay = TG.fault_trace(1 << 18, 1, "uniform", 4)
ctx.fault_buffer_init(TG.gpu_params(4096), tbl)
ring = np.zeros(((1 << 18) + 1, 8), np.uint32); ring[:1 << 18] = pk
print("big batches", len(ctx.fetch_preprocess_fault_batches(ring, 0, 1 << 18)["batch"]))
print(bench.nonreplayable_numbers(ctx, lib, P, 0)["ms"])

# one round of the tiered loop: 2048 va_blocks over 1024 chunks, 1024 entries (4 batches of 256 = one round), HBM full before it
from uvmtier import TIER_CXL, TIER_DRAM, TIER_HBM  # noqa: E402
from uvmtier.blockstate import HostBlockState  # noqa: E402
blocks, chunks = 2048, 1200
gp = TG.gpu_params(256)
info = {TIER_HBM: ctx.tier_create(TIER_HBM, chunks << 21), TIER_DRAM: ctx.tier_create(TIER_DRAM, blocks << 21),
        TIER_CXL: ctx.tier_create(TIER_CXL, blocks << 21)}
pk, tbl, lay = TG.fault_trace(1024 * 2, 9, "uniform", 1, blocks << 21)
ctx.fault_buffer_init(gp, tbl)
layout = BSL.make_layout(lay)
st = HostBlockState(blocks)
st.cpu_resident0[:] = np.uint64(0xFFFFFFFFFFFFFFFF)
ds = ctx.block_state_alloc(layout); ds.upload(st)
pmm = ctx.pmm_create(layout, chunks)
ctx.pmm_enable_chunk_table(pmm)
ring = np.zeros((len(pk) + 1, 8), np.uint32); ring[:len(pk)] = pk
k1 = ctx.fetch_preprocess_fault_batches(ring, 0, len(pk))
got = ctx.service_fault_batches_tiered(k1, layout, ds, pmm, 4, 600 * 512, execute_copies=True, demote=True)
s = got["stats"]
print("tiered", s.passes, s.eviction_waves, s.blocks_demoted, s.pages_evicted, s.pages_promoted)
ctx.close()
