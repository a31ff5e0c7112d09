#!/usr/bin/env python
"""bench.py -- headline benchmark of the nvidia-uvm tiering hot path on B200 (BASELINE.json).

A step = one pass of the hot path over one synthetic batch of work per GPU (config B):
  1,048,576 replayable-fault packets (uniform over 64 GB of VA, 4 VA spaces, 256-entry batches)
  -> K0/K1  decode + fetch-time coalescing + sort + va_block grouping        (uvmtier_fetch_preprocess_fault_batches)
  -> service per va_block: duplicates / permissions / authorised skip, K2 prefetch bitmap trees, residency algebra,
             copy lists, mapping state over the 32,768 va_blocks of that VA  (uvmtier_service_fault_batches)
Metric: fault entries/s preprocessed (whole job = all ranks).  `value` is measured with inputs resident in HBM
(CUDA events on the library's stream, L2 flushed between iterations); `e2e` goes through the host-buffer C-ABI
call with pinned host buffers, H2D and D2H inside the timed region.  The second half of the metric, page-migration
GB/s, is reported in the `migration` object (SM-driven copies vs the copy-engine peak measured in the same run).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--trace uniform|stream|hotset]

Multi-GPU (torchrun): the VA space is sharded, every rank runs its own trace on its own GPU, no data-path
collective; torch.distributed is used for the barrier, the max-over-ranks and to pass CUDA IPC
handles for the NVLink peer-copy measurement.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "open-gpu-kernel-modules_b200"))

N_ENTRIES = 1 << 20
BATCH = 256
N_SPACES = 4
VA_BYTES = 64 << 30
N_BLOCKS = VA_BYTES >> 21
L2_FLUSH_BYTES = 256 << 20
RUN_CAP = 1 << 23
METRIC = "fault entries/s preprocessed (1M-entry trace: decode+coalesce+sort+group by va_block, serviced with prefetch trees over 64 GB VA)"
UNIT = "entries/s"


def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def bind_near_gpu(index):
    """Pins this process (the thread that services GPU `index`, and the pinned buffers it allocates next: first touch) to
    the CPUs of the GPU's NUMA node, as UVM/uvm_gpu_isr.c:598-625 does for the fault-servicing bottom half.  Returns what was
    found and done, for the bench line; the caller restores the old affinity afterwards."""
    info = {"gpu": index, "affinity_before": len(os.sched_getaffinity(0))}
    try:
        bdf = subprocess.run(["nvidia-smi", "-i", str(index), "--query-gpu=pci.bus_id", "--format=csv,noheader"],
                             capture_output=True, text=True, timeout=20).stdout.strip().lower()
        info["pci_bus_id"] = bdf
        short = bdf[4:] if len(bdf) > 12 else bdf                      # nvidia-smi prints an 8-digit domain, sysfs a 4-digit one
        node = int(open("/sys/bus/pci/devices/%s/numa_node" % short).read().strip())
        info["numa_node"] = node
        nodes = [d for d in os.listdir("/sys/devices/system/node") if d.startswith("node")]
        info["numa_nodes_on_host"] = len(nodes)
        if node >= 0:
            cpus = set()
            for part in open("/sys/devices/system/node/node%d/cpulist" % node).read().strip().split(","):
                a, _, b = part.partition("-")
                cpus.update(range(int(a), int(b or a) + 1))
            cpus &= os.sched_getaffinity(0)
            if cpus:
                os.sched_setaffinity(0, cpus)
                info["bound_to_cpus"] = len(cpus)
    except Exception as e:                                              # no sysfs entry, no nvidia-smi: run unbound and say so
        info["error"] = repr(e)[:120]
    return info


class ClockSampler(threading.Thread):
    """Samples nvidia-smi clocks / throttle reasons while the timed regions run."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    PAD = 0.06

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.rows = []
        self.proc = None

    def run(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "50"], stdout=subprocess.PIPE, text=True)
            for line in self.proc.stdout:
                self.rows.append((time.perf_counter(), [c.strip() for c in line.split(",")]))
        except Exception:
            pass

    def stop(self, windows=()):
        if self.proc:
            self.proc.terminate()
        self.join(timeout=2)
        sm, mx, reasons, n_in = [], 0, set(), 0
        for t, r in self.rows:
            try:
                # the timed regions last a few milliseconds and nvidia-smi answers every 50 ms: a sample counts when it
                # falls within one sampling period of a timed region (the GPU runs the warm-up steps right before it
                # and the next phase right after it, so it is under the same load)
                inside = any(a - self.PAD <= t <= b + self.PAD for a, b in windows) if windows else True
                mx = max(mx, float(r[2]))
                if not inside:
                    continue
                n_in += 1
                sm.append(float(r[1]))
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                continue
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples_under_load": n_in, "samples": len(self.rows)}


# ----------------------------------------------------------------------------------------------- workload
def build_workload(seed, trace, rank=0, world=1):
    """Trace + instance-pointer table + initial va_block state (everything resident on the CPU tiers: 70 % host
    DRAM, 30 % CXL node, nothing on the GPU yet, read-write VMAs).
    world > 1: ONE VA of world x 64 GB and ONE trace of world x 1M entries over it; the va_blocks of
    every VA space are cut into `world` contiguous ranges (block b -> GPU b * world / n_blocks) and every GPU gets
    the faults on the blocks it owns, in ring order, plus the state table of its range."""
    from uvmtier import sharding as SH, tracegen as TG
    from uvmtier.blockstate import HostBlockState, make_layout
    pk, tbl, lay = TG.fault_trace(N_ENTRIES * world, seed, trace, N_SPACES, VA_BYTES * world)
    if world > 1:
        bps = lay["blocks_per_space"]
        pk = np.ascontiguousarray(SH.split_trace_by_owner(pk, lay["va_base"], bps, world)[rank])
        first, outer = SH.block_range(bps, world, rank)
        lay = dict(lay, va_base=lay["va_base"] + (first << 21), blocks_per_space=outer - first)
    gp = TG.gpu_params(BATCH)
    n_slots = lay["blocks_per_space"] * N_SPACES
    st = HostBlockState(n_slots)
    # ~25 % of every va_block's 64 KB chunks live on the CXL node, the rest on host DRAM (tiering works on
    # multi-page extents, not single pages)
    rs = TG.Stream((seed + 1000 * rank) ^ 0x5EED, n_slots * 2)
    chunk_bits = (rs.take(n_slots) & rs.take(n_slots)).astype("<u4")                  # 32 chunks of 16 pages per block
    chunks = np.unpackbits(chunk_bits.view(np.uint8).reshape(n_slots, 4), axis=1, bitorder="little")
    pages = np.repeat(chunks, 16, axis=1)
    st.cpu_resident1 = np.ascontiguousarray(np.packbits(pages, axis=1, bitorder="little")).view("<u8").reshape(n_slots, 8).copy()
    st.cpu_resident0 = (~st.cpu_resident1).copy()
    return dict(packets=pk, table=tbl, gpu_params=gp, layout=make_layout(lay), lay=lay, state=st, n_slots=n_slots)


# ----------------------------------------------------------------------------------------------- CPU arm
def _ref_worker(args):
    """One process: the CPU implementation on a shard of whole batches.  Stage A/B run the reference's own C
    (oracle/_ref/libref_faults.so) when it was built; stage C/D/E the reference's own per-va_block entry loop
    (oracle/_ref/libref_service.so) with the CPU port of the uvm_va_block.c back half behind it, else the port."""
    pk, tbl, gp, layout, state, reps = args
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import faultlib as F
    import servicelib as S
    import uvmtier_testlib as T
    orc = T.load_oracle()
    rf = T.load_ref("faults")
    rs = T.load_ref("service")
    # the back half of the reference's entry loop: the reference's own prefetch file and make_resident when they were
    # built too (then nothing on the CPU arm's data path is a restatement except the mapping of the serviced pages)
    back = rs is not None and S.use_reference_back_half(rs, T.load_ref("prefetch"), T.load_ref("resident"))
    kind = "reference" if rf is not None and rs is not None and back else "port"
    sim_like = None
    best = 1e30
    B = int(gp["max_batch_size"][0])
    for _ in range(reps):
        t0 = time.perf_counter()
        r = F.run_fetch_preprocess(rf if rf is not None else orc, "ref" if rf is not None else "oracle", pk, gp, tbl)
        t_ab = time.perf_counter() - t0
        # stage C/D/E needs the sorted view in the product layout: built by the port (not timed twice: the
        # reference already produced the same order above; only the conversion below is untimed)
        if sim_like is None:
            sim_like = _port_sorted_view(orc, pk, gp, tbl, r)
        t1 = time.perf_counter()
        if rs is not None:
            S.run_service_ref(rs, sim_like, gp, tbl, layout, state, run_cap=1 << 22)
        else:
            S.run_service_oracle(orc, sim_like, gp, tbl, layout, state, run_cap=1 << 22)
        t_c = time.perf_counter() - t1
        best = min(best, t_ab + t_c)
    return best, kind


def _port_sorted_view(orc, pk, gp, tbl, r):
    """Product-layout sorted view (ordered records + va_block runs) from a CPU checker's fetch result."""
    import faultlib as F
    B = int(gp["max_batch_size"][0])
    n = len(pk)
    nb = len(r["batch"])
    e = r["entries"]
    ordered = np.zeros(nb * B, F.ORDERED_DT)
    groups = np.zeros(nb * B, F.GROUP_DT)
    batch = r["batch"].copy()
    inst_to_ctx = {int(v): k for k, v in enumerate(tbl["instance_ptr"])}
    ctx_of = np.array([inst_to_ctx.get(int(v), 0) for v in np.unique(e["instance_ptr_address"])])
    uniq = np.unique(e["instance_ptr_address"])
    ctx_all = ctx_of[np.searchsorted(uniq, e["instance_ptr_address"])]
    for b in range(nb):
        if batch["status"][b] != 0:
            continue
        base = b * B
        m = int(batch["num_coalesced_faults"][b])
        idx = r["ordered"][base:base + m].astype(np.int64) + base
        rec = ordered[base:base + m]
        rec["fault_address"] = e["fault_address"][idx]
        rec["fault_index"] = idx - base
        rec["ctx"] = ctx_all[idx]
        rec["num_instances"] = e["num_instances"][idx]
        rec["access_type_mask"] = e["access_type_mask"][idx]
        fatal = (e["flags"][idx] & 1) != 0
        rec["fault_access_type"] = e["fault_access_type"][idx] | np.where(fatal, 0x80, 0).astype("u1")
        key = (tbl["sort_rank"][rec["ctx"]].astype(np.uint64) << np.uint64(52)) | (rec["fault_address"] >> np.uint64(21))
        bnd = np.ones(m, bool)
        if m > 1:
            bnd[1:] = (key[1:] != key[:-1]) | fatal[1:] | fatal[:-1]
        starts = np.nonzero(bnd)[0]
        g = groups[base:base + len(starts)]
        g["block_start"] = rec["fault_address"][starts] & ~np.uint64((2 << 20) - 1)
        g["ctx"] = rec["ctx"][starts]
        g["first_fault_index"] = starts
        g["block_faults"] = np.diff(np.append(starts, m))
        g["flags"] = fatal[starts]
        batch["num_block_groups"][b] = len(starts)
    return dict(ordered_rec=ordered, groups=groups, batch=batch)


def cpu_arm(wl, threads, reps=3):
    """Times the CPU implementation on `threads` processes, each with a contiguous shard of whole batches and its
    own copy of the va_block state (like one service thread per GPU in the reference)."""
    import multiprocessing as mp
    n = N_ENTRIES // (BATCH * threads) * (BATCH * threads)
    jobs = []
    for t in range(threads):
        a, b = t * n // threads, (t + 1) * n // threads
        jobs.append((wl["packets"][a:b].copy(), wl["table"], wl["gpu_params"], wl["layout"], wl["state"], reps))
    if threads == 1:
        res = [_ref_worker(jobs[0])]
    else:
        with mp.get_context("fork").Pool(threads) as pool:
            res = pool.map(_ref_worker, jobs)
    t = max(r[0] for r in res)
    return n / t, res[0][1], n


def workload_name(trace):
    return ("1xB200 config B: 1M-entry %s fault trace (4 VA spaces, batch 256) sort/dedupe/coalesce + service with prefetch "
            "trees over 64 GB VA (32768 va_blocks)" % trace)


def config_dict(trace, world):
    """The `config` object of the JSON line: the same keys and values in both arms."""
    return {"workload": workload_name(trace), "batch": BATCH, "va_spaces": N_SPACES,
            "l2": "flushed between iterations (256 MiB memset)", "rng": "uvm_test_rng KISS stream, seed 1",
            "initial_residency": "all pages on the CPU tiers (75% DRAM / 25% CXL node), none on the GPU",
            "sharding": ("one GPU: the whole 64 GB VA and the whole 1M-entry trace" if world == 1 else
                         "ONE VA of %d x 64 GB (4 VA spaces) cut into %d contiguous va_block ranges, ONE trace of %d x 1M "
                         "entries split by the owner of the faulting va_block (uvmtier/sharding.py); every GPU services "
                         "its own range, no data-path collective; weak scaling" % (world, world, world))}


def run_reference_arm(args, rank):
    if rank != 0:
        return
    wl = build_workload(1, args.trace)
    cores = os.cpu_count() or 1
    for _ in range(args.warmup):
        cpu_arm(wl, cores, reps=1)
    vals, kind, n = [], "port", N_ENTRIES
    t0 = time.perf_counter()
    for _ in range(args.steps):
        v, kind, n = cpu_arm(wl, cores, reps=1)
        vals.append(v)
    dt = (time.perf_counter() - t0) / max(1, args.steps)
    value = float(np.mean(vals))
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u64", "data": "synthetic",
            "config": config_dict(args.trace, 1),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind,
                             "sample": "%d entries per step sharded over %d processes, each with its own va_block state; "
                                       "fetch/merge/sort, the per-va_block entry loop, the prefetch hint and make_resident are the "
                                       "reference's own C (oracle/_ref), the mapping of the serviced pages is restated" % (n, cores)},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


# ----------------------------------------------------------------------------------------------- migration leg
def migration_numbers(ctx, lib, P, rank, world, dist, arena=8 << 30):
    """Page-migration GB/s (BASELINE config C as a compact sweep, and config A's CPU copy beside it):
      * 4 KB / 64 KB / 2 MB pages x {1 GB, 8 GB} x {DRAM->HBM, CXL->HBM, HBM->DRAM, HBM->CXL}, pages placed by a random
        permutation, SM-driven TMA copies (uvmtier_copy_page_list_device).  Every point is measured right after the
        copy-engine peak of the same direction, size and buffers (cudaMemcpyAsync, best of 2), all ranks doing the same
        pair at the same time, so the fraction compares like with like also when eight GPUs share the host;
      * DRAM->CXL for 64 MB and 1 GB: the reference's CPU page loop (uvm_va_block.c:3946-3991, kmap + memcpy per page,
        compiled verbatim: oracle/_ref/libref_copy.so) on 1 thread and on all cores, and the GPU moving the same pages;
      * the fork's CXL P2P DMA request (one contiguous range, 2 MB pages)."""
    import uvmtier
    from uvmtier import TIER_CXL, TIER_DRAM, TIER_HBM
    info = {t: ctx.tier_create(t, arena) for t in (TIER_HBM, TIER_DRAM, TIER_CXL)}
    for t in (TIER_DRAM, TIER_CXL):
        info[t].host_view()[:] = 0x5A
    ms = ctypes.c_float(0)
    out = {"order": "random permutation", "unit": "GB/s",
           "cxl_node": "emulated on a DRAM node (no CPU-less NUMA node on this machine)" if info[TIER_CXL].flags & 1 else
           "NUMA node %d" % info[TIER_CXL].numa_node}

    def sync_all():
        ctx.synchronize()
        if dist is not None:
            dist.barrier()

    def timed(dst, src, d_d, d_s, n, page_size, reps=2):
        best = 1e30
        for r in range(reps + 1):
            lib.uvmtier_event_record(ctx.handle, 900)
            st = lib.uvmtier_copy_page_list_device(ctx.handle, dst, src, ctypes.c_size_t(n), P(d_d), P(d_s), page_size, 0)
            assert st == 0, hex(st)
            lib.uvmtier_event_record(ctx.handle, 901)
            lib.uvmtier_event_elapsed_ms(ctx.handle, 900, 901, ctypes.byref(ms))
            if r:
                best = min(best, ms.value)
        return n * page_size / (best * 1e-3) / 1e9

    names = {TIER_HBM: "hbm", TIER_DRAM: "dram", TIER_CXL: "cxl"}
    sweep = []
    rng = np.random.default_rng(7)
    for nbytes in (1 << 30, arena):
        for page_size in (4096, 65536, 2 << 20):
            n = nbytes // page_size
            s_off = np.arange(n, dtype=np.uint64) * np.uint64(page_size)
            d_off = rng.permutation(n).astype(np.uint64) * np.uint64(page_size)
            d_s, d_d = ctx.to_device(s_off), ctx.to_device(d_off)
            for dst, src in ((TIER_HBM, TIER_DRAM), (TIER_HBM, TIER_CXL), (TIER_DRAM, TIER_HBM), (TIER_CXL, TIER_HBM)):
                sync_all()
                peak = ctx.tier_memcpy_peak_gbs(dst, src, nbytes, 2)
                gbs = timed(dst, src, d_d, d_s, n, page_size)
                row = {"direction": "%s_to_%s" % (names[src], names[dst]), "page_size": page_size, "bytes": nbytes,
                       "GBs": round(gbs, 2), "copy_engine_peak_GBs": round(peak, 2), "frac_of_link_peak": round(gbs / peak, 3)}
                sweep.append(row)
                if page_size == 4096 and nbytes == 1 << 30:
                    out[row["direction"]] = {k: row[k] for k in ("GBs", "copy_engine_peak_GBs", "frac_of_link_peak")}
            ctx.free(d_s); ctx.free(d_d)
    out["sweep"] = sweep
    out["min_frac_of_link_peak"] = min(r["frac_of_link_peak"] for r in sweep)

    # ---- DRAM -> CXL: the reference's CPU page loop beside the GPU (rank 0 only: the host cores are shared) ----
    if rank == 0:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import uvmtier_testlib as T
        rc = T.load_ref("copy")
        cpu = {"what": "uvm_va_block.c:3946-3991 compiled verbatim (kmap + memcpy(PAGE_SIZE) per page), DRAM tier -> CXL tier",
               "available": rc is not None}
        src_p, dst_p = info[TIER_DRAM].host_ptr, info[TIER_CXL].host_ptr
        cores = os.cpu_count() or 1
        for nbytes, key in ((64 << 20, "64MB"), (1 << 30, "1GB")):
            blocks = nbytes >> 21
            entry = {}
            if rc is not None:
                rc.ref_copy_pages_cpu.restype = ctypes.c_uint64
                rc.ref_copy_pages_cpu.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint32,
                                                  ctypes.c_uint32]
                for threads in (1, cores):
                    best = 1e30
                    for _ in range(3):
                        th = [threading.Thread(target=rc.ref_copy_pages_cpu,
                                               args=(dst_p, src_p, t * blocks // threads, (t + 1) * blocks // threads, 0, 512))
                              for t in range(threads)]
                        t0 = time.perf_counter()
                        for x in th: x.start()
                        for x in th: x.join()
                        best = min(best, time.perf_counter() - t0)
                    entry["cpu_%d_thread%s_GBs" % (threads, "" if threads == 1 else "s")] = round(nbytes / best / 1e9, 2)
            n = nbytes // 4096
            lin = np.arange(n, dtype=np.uint64) * np.uint64(4096)
            d_l = ctx.to_device(lin)
            entry["gpu_GBs"] = round(timed(TIER_CXL, TIER_DRAM, d_l, d_l, n, 4096), 2)
            ctx.free(d_l)
            cpu[key] = entry
        cpu["cores"] = cores
        out["dram_to_cxl"] = cpu

    # the fork's CXL P2P DMA request (one contiguous range between the CXL buffer and HBM, 2 MB pages)
    nbytes = 1 << 30
    hnd = ctypes.c_void_p()
    st = lib.uvmtier_p2p_register_cxl_buffer(ctx.handle, ctypes.c_void_p(info[TIER_CXL].host_ptr), ctypes.c_uint64(nbytes),
                                             ctypes.c_uint32(2), ctypes.byref(hnd))
    assert st == 0, hex(st)
    for flag, key, dst, src in ((1, "cxl_p2p_dma_cxl_to_gpu", TIER_HBM, TIER_CXL), (0, "cxl_p2p_dma_gpu_to_cxl", TIER_CXL, TIER_HBM)):
        sync_all()
        peak = ctx.tier_memcpy_peak_gbs(dst, src, nbytes, 2)
        best = 1e30
        for r in range(3):
            lib.uvmtier_event_record(ctx.handle, 900)
            st = lib.uvmtier_p2p_cxl_dma_request(ctx.handle, hnd, ctypes.c_uint64(0), ctypes.c_uint64(0), ctypes.c_uint64(nbytes),
                                                 ctypes.c_uint32(flag | 2), None)
            assert st == 0, hex(st)
            lib.uvmtier_event_record(ctx.handle, 901)
            lib.uvmtier_event_elapsed_ms(ctx.handle, 900, 901, ctypes.byref(ms))
            if r:
                best = min(best, ms.value)
        out[key] = {"GBs": round(nbytes / (best * 1e-3) / 1e9, 2), "copy_engine_peak_GBs": round(peak, 2),
                    "frac_of_link_peak": round(nbytes / (best * 1e-3) / 1e9 / peak, 3)}
    assert lib.uvmtier_p2p_unregister_cxl_buffer(ctx.handle, hnd) == 0
    for t in (TIER_HBM, TIER_DRAM, TIER_CXL):
        ctx.tier_destroy(t)
    return out


def oversubscription_numbers(ctx, lib, P, rank, blocks=8192, hbm_chunks=4096, dram_blocks=3072):
    """BASELINE config D at a bounded size: the fault path with HBM oversubscribed 2x.  16 GB of VA (8192 va_blocks) over an
    HBM arena of 4096 root chunks (8 GB, slot -> chunk table), host DRAM budget 6 GB of this table's pages, the rest on the
    CXL node.  Every round of 16 batches evicts least recently used va_blocks, demotes the coldest DRAM va_blocks to CXL to
    keep the budget, services, and copies: real pages move over the host link in both directions.  Two traces on the same
    table: uniform (every fault is a 64 KB migration: link-bound) and hotset (most faults land on pages that are already
    resident and mapped: fault-filtering bound).  Timed end to end on the host clock around one call (the loop decides
    everything on the device and synchronises once)."""
    import uvmtier
    from uvmtier import TIER_CXL, TIER_DRAM, TIER_HBM, tracegen as TG
    from uvmtier.blockstate import HostBlockState, make_layout
    B = 256
    chunk_batches = max(1, min(16, hbm_chunks // B))      # a round may not fault on more va_blocks than HBM has chunks
    t_setup = time.perf_counter()
    gp = TG.gpu_params(B)
    info = {TIER_HBM: ctx.tier_create(TIER_HBM, hbm_chunks << 21), TIER_DRAM: ctx.tier_create(TIER_DRAM, blocks << 21),
            TIER_CXL: ctx.tier_create(TIER_CXL, blocks << 21)}
    setup_s = time.perf_counter() - t_setup
    # the copy-engine peaks of this box, both directions at once is what the loop asks of the link
    peak = {}
    for name, (dst, src) in (("h2d", (TIER_HBM, TIER_DRAM)), ("d2h", (TIER_DRAM, TIER_HBM))):
        ms = ctypes.c_float(0)
        uvmtier._check(lib.uvmtier_tier_memcpy_peak(ctx.handle, ctypes.c_uint32(dst), ctypes.c_uint32(src), ctypes.c_uint64(1 << 30), 3,
                                                    ctypes.byref(ms)), "peak")
        peak[name] = (1 << 30) / (ms.value * 1e-3) / 1e9
    out = {"workload": "%d va_blocks (%d GB of VA) over an HBM arena of %d root chunks (chunk table), DRAM budget %d va_blocks of pages, rest on the CXL node; rounds of %d batches of %d"
                       % (blocks, blocks >> 9, hbm_chunks, dram_blocks, chunk_batches, B),
           "tier_setup_s": setup_s, "ce_peak_h2d_GBs": peak["h2d"], "ce_peak_d2h_GBs": peak["d2h"],
           "policy": "victims in the reference's order (pick_root_chunk_to_evict) until enough of them are va_blocks the round does not "
                     "fault on; evicted pages go to host DRAM, the coldest DRAM va_blocks are demoted to the CXL node to keep the budget"}
    for trace, n_entries in (("uniform", 262144), ("hotset", 1048576)):
        pk, tbl, lay = TG.fault_trace(n_entries, 77 + rank, trace, 1, blocks << 21)
        ctx.fault_buffer_init(gp, tbl)
        layout = make_layout(lay)
        st = HostBlockState(blocks)
        st.cpu_resident0[:dram_blocks] = np.uint64(0xFFFFFFFFFFFFFFFF)      # the table starts on the CPU tiers, DRAM at its budget
        st.cpu_resident1[dram_blocks:] = np.uint64(0xFFFFFFFFFFFFFFFF)
        ring = np.zeros((n_entries + 1, 8), np.uint32)
        ring[:n_entries] = pk
        k1 = ctx.fetch_preprocess_fault_batches(ring, 0, n_entries)
        nb = len(k1["batch"])
        ds = ctx.block_state_alloc(layout)
        ds.upload(st)
        pmm = ctx.pmm_create(layout, hbm_chunks)
        ctx.pmm_enable_chunk_table(pmm)
        d_ord, d_grp, d_bat = ctx.to_device(k1["ordered"]), ctx.to_device(k1["groups"]), ctx.to_device(k1["batch"])
        d_flags, d_svc = ctx.alloc(nb * B * 2), ctx.alloc(nb * B * 96)
        stats = uvmtier.TieringStats()
        ctx.synchronize()
        launches0 = lib.uvmtier_kernel_launch_count()
        t0 = time.perf_counter()
        st_code = lib.uvmtier_service_fault_batches_tiered_device(
            ctx.handle, ctypes.byref(pmm), ctypes.c_uint32(nb), P(d_ord), P(d_grp), P(d_bat), P(layout), ctypes.byref(ds.ptrs), P(d_flags),
            P(d_svc), ctypes.c_uint32(chunk_batches), ctypes.c_uint64(dram_blocks * 512), ctypes.c_int(3), ctypes.byref(stats))
        ctx.synchronize()
        dt = time.perf_counter() - t0
        assert st_code == 0, hex(st_code)
        out_b = stats.pages_evicted * 4096 / 1e9
        in_b = stats.pages_promoted * 4096 / 1e9
        # demoted pages cross the link twice (DRAM -> GPU -> CXL node), once in each direction
        dem_b = stats.pages_demoted * 4096 / 1e9
        out[trace] = {"entries": n_entries, "entries_per_s": n_entries / dt, "ms": dt * 1e3, "rounds": stats.passes,
                      "eviction_waves": stats.eviction_waves, "root_chunks_evicted_to_dram": stats.evicted_to_dram,
                      "root_chunks_evicted_to_cxl": stats.evicted_to_cxl, "va_blocks_demoted": stats.blocks_demoted,
                      "peak_chunks_in_use": stats.peak_chunks_in_use, "overcommitted_rounds": stats.overcommitted_passes,
                      "GB_evicted": out_b, "GB_promoted": in_b, "GB_demoted": dem_b,
                      "h2d_GBs": (in_b + dem_b) / dt, "d2h_GBs": (out_b + dem_b) / dt,
                      "h2d_frac_of_ce_peak": (in_b + dem_b) / dt / peak["h2d"], "d2h_frac_of_ce_peak": (out_b + dem_b) / dt / peak["d2h"],
                      # the loop's copies run one after the other on one stream: the share of the wall time the link spends
                      # moving this call's bytes at the copy-engine rate of their direction
                      "link_busy_frac": ((in_b + dem_b) / peak["h2d"] + (out_b + dem_b) / peak["d2h"]) / dt,
                      "gpu_launches": int(lib.uvmtier_kernel_launch_count() - launches0)}
        for p in (d_ord, d_grp, d_bat, d_flags, d_svc):
            ctx.free(p)
        ctx.pmm_destroy(pmm)
        ds.free()
    for t in (TIER_HBM, TIER_DRAM, TIER_CXL):
        ctx.tier_destroy(t)
    return out


def access_counter_numbers(ctx, lib, P, rank):
    """The access-counter path on the same 64 GB table shape, bounded: 262144 notifications (2 MB tracking, random
    sub-granularity masks) over 4 VA spaces x 8192 va_blocks whose pages live on the CPU tiers and are mapped remotely by
    the GPU; fetch + sort, then expansion + servicing.  Device-resident inputs, CUDA events on the library's stream."""
    import uvmtier
    from uvmtier import tracegen as TG
    from uvmtier.blockstate import HostBlockState, make_layout
    n, B, n_spaces, blocks = 262144, 256, 4, 8192
    rng = np.random.default_rng(5 + rank)
    tbl = TG.instance_ptr_table(n_spaces)
    gp = TG.gpu_params(B)
    ctx.fault_buffer_init(gp, tbl)
    lay = dict(va_base=TG.VA_BASE, blocks_per_space=blocks, n_spaces=n_spaces)
    layout = make_layout(lay)
    sp = rng.integers(0, n_spaces, n)
    addr = (TG.VA_BASE + rng.integers(0, blocks, n).astype(np.uint64) * np.uint64(2 << 20)).astype(np.uint64)
    sub = (rng.integers(0, 1 << 32, n, dtype=np.uint64) & rng.integers(0, 1 << 32, n, dtype=np.uint64)).astype(np.uint32)
    w = np.zeros((n + 1, 8), np.uint32)
    ip = tbl["instance_ptr"][sp] >> np.uint64(12)
    w[:n, 0] = 1 | ((ip & np.uint64(0xFFFFF)) << np.uint64(12)).astype(np.uint32)
    w[:n, 1] = (ip >> np.uint64(20)).astype(np.uint32)
    w[:n, 2] = (addr & np.uint64(0xFFFFFFFF)).astype(np.uint32)
    w[:n, 3] = ((addr >> np.uint64(32)) & np.uint64((1 << 25) - 1)).astype(np.uint32)
    w[:n, 4] = sub
    w[:n, 6] = 256
    w[:n, 7] = (384 << 20) | (1 << 31)
    st = HostBlockState(n_spaces * blocks)
    st.cpu_resident0[:] = np.uint64(0xFFFFFFFFFFFFFFFF)
    st.gpu_pte0[:] = np.uint64(0xFFFFFFFFFFFFFFFF)                       # remote mappings: what makes the counters tick
    ds, ds0 = ctx.block_state_alloc(layout), ctx.block_state_alloc(layout)
    ds0.upload(st)
    nb = n // B
    d_ring0, d_ring = ctx.to_device(w), ctx.alloc(w.nbytes)
    d_ord, d_grp, d_bat = ctx.alloc(n * 16), ctx.alloc(n * 16), ctx.alloc(nb * 32)
    run_cap = 1 << 22
    d_svc, d_runs, d_cnt = ctx.alloc(n * 96), ctx.alloc(run_cap * 8), ctx.alloc(8)
    state_bytes = n_spaces * blocks * (6 * 64 + 7)
    new_get, ms = ctypes.c_uint32(0), ctypes.c_float(0)
    t_pre, t_svc = [], []
    for it in range(6):
        lib.uvmtier_memcpy_d2d_async(ctx.handle, ctypes.c_void_p(ds.ptrs.gpu_resident), ctypes.c_void_p(ds0.ptrs.gpu_resident),
                                     ctypes.c_size_t(state_bytes))
        lib.uvmtier_memcpy_d2d_async(ctx.handle, P(d_ring), P(d_ring0), ctypes.c_size_t(w.nbytes))
        lib.uvmtier_event_record(ctx.handle, 910)
        st1 = lib.uvmtier_access_counters_fetch_preprocess_device(ctx.handle, P(d_ring), ctypes.c_uint32(n + 1), ctypes.c_uint32(0),
                                                                  ctypes.c_uint32(n), ctypes.c_int(1), P(d_ord), P(d_grp), P(d_bat), None,
                                                                  ctypes.byref(new_get))
        lib.uvmtier_event_record(ctx.handle, 911)
        st2 = lib.uvmtier_access_counters_service_device(ctx.handle, ctypes.c_uint32(nb), P(d_ord), P(d_grp), P(d_bat), P(layout),
                                                         ctypes.byref(ds.ptrs), P(d_svc), P(d_runs), ctypes.c_uint32(run_cap), P(d_cnt))
        lib.uvmtier_event_record(ctx.handle, 912)
        assert st1 == 0 and st2 == 0, (hex(st1), hex(st2))
        lib.uvmtier_event_elapsed_ms(ctx.handle, 910, 911, ctypes.byref(ms)); a = ms.value
        lib.uvmtier_event_elapsed_ms(ctx.handle, 911, 912, ctypes.byref(ms)); b = ms.value
        if it >= 2:
            t_pre.append(a); t_svc.append(b)
    cnt = ctx.from_device(d_cnt, 2, "<u4")
    pre, svc = float(np.mean(t_pre)), float(np.mean(t_svc))
    out = {"notifications": n, "va_blocks": n_spaces * blocks, "fetch_preprocess_ms": pre, "service_ms": svc,
           "notifications_per_s": n / ((pre + svc) * 1e-3), "copy_runs": int(cnt[0]),
           "what": "2 MB tracking granularity, 2 MB CPU chunks, every page on host DRAM and remotely mapped"}
    for p in (d_ring0, d_ring, d_ord, d_grp, d_bat, d_svc, d_runs, d_cnt):
        ctx.free(p)
    ds.free(); ds0.free()
    return out


def nonreplayable_numbers(ctx, lib, P, rank):
    """Non-replayable faults on the 64 GB table shape, bounded: 262144 copy-engine faults (uniform pages, every page on
    the CPU tiers) parsed, translated and serviced one by one in buffer order per va_block.  Device-resident shadow
    buffer, CUDA events on the library's stream."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import nonreplaylib as NR
    from uvmtier import tracegen as TG
    from uvmtier.blockstate import HostBlockState, make_layout
    n, B, n_spaces, blocks = 262144, 256, 4, 8192
    tbl = TG.instance_ptr_table(n_spaces)
    ctx.fault_buffer_init(TG.gpu_params(B), tbl)
    lay = dict(va_base=TG.VA_BASE, blocks_per_space=blocks, n_spaces=n_spaces, gpu_id=1)
    layout = make_layout(lay)
    pk = NR.shadow_buffer(n, 13 + rank, lay, tbl, fatal_frac=0.0, stale_frac=0.0, outside_frac=0.0)
    st = HostBlockState(n_spaces * blocks)
    st.cpu_resident0[:] = np.uint64(0xFFFFFFFFFFFFFFFF)
    ds, ds0 = ctx.block_state_alloc(layout), ctx.block_state_alloc(layout)
    ds0.upload(st)
    d_pk = ctx.to_device(pk)
    run_cap = 1 << 22
    d_flags, d_svc, d_runs, d_cnt = ctx.alloc(n * 2), ctx.alloc(n * 96), ctx.alloc(run_cap * 8), ctx.alloc(8)
    state_bytes = n_spaces * blocks * (6 * 64 + 7)
    ms, times = ctypes.c_float(0), []
    for it in range(6):
        lib.uvmtier_memcpy_d2d_async(ctx.handle, ctypes.c_void_p(ds.ptrs.gpu_resident), ctypes.c_void_p(ds0.ptrs.gpu_resident),
                                     ctypes.c_size_t(state_bytes))
        lib.uvmtier_event_record(ctx.handle, 940)
        st1 = lib.uvmtier_service_non_replayable_faults_device(ctx.handle, P(d_pk), ctypes.c_uint32(n), P(layout), ctypes.byref(ds.ptrs), None,
                                                               P(d_flags), P(d_svc), P(d_runs), ctypes.c_uint32(run_cap), P(d_cnt))
        lib.uvmtier_event_record(ctx.handle, 941)
        assert st1 == 0, hex(st1)
        lib.uvmtier_event_elapsed_ms(ctx.handle, 940, 941, ctypes.byref(ms))
        if it >= 2:
            times.append(ms.value)
    cnt = ctx.from_device(d_cnt, 2, "<u4")
    flags = ctx.from_device(d_flags, n, "<u2")
    m = float(np.mean(times))
    out = {"faults": n, "va_blocks": n_spaces * blocks, "ms": m, "faults_per_s": n / (m * 1e-3), "serviced": int((flags & 1).sum()),
           "copy_runs": int(cnt[0]), "what": "copy-engine faults, every page on host DRAM, serviced in buffer order per va_block"}
    for p_ in (d_pk, d_flags, d_svc, d_runs, d_cnt):
        ctx.free(p_)
    ds.free(); ds0.free()
    return out


def thrashing_numbers(ctx, lib, P, rank):
    """The thrashing detector on the 64 GB table shape: 32768 va_blocks x 32 events of 16 pages each per call, two
    processors taking turns inside the window (every page of a region is read, stamped and written back).  Device
    resident events, grouped by va_block; CUDA events on the library's stream."""
    from uvmtier.blockstate import make_layout
    from uvmtier import tracegen as TG
    n_slots, per, pages, calls = 32768, 32, 16, 6
    n = n_slots * per
    rng = np.random.default_rng(11 + rank)
    layout = make_layout(dict(va_base=TG.VA_BASE, blocks_per_space=n_slots // 4, n_spaces=4))
    ev_dt = np.dtype([("time_stamp", "<u8"), ("slot", "<u4"), ("first_page", "<u2"), ("num_pages", "<u2"), ("processor", "<u2"),
                      ("kind", "u1"), ("flags", "u1"), ("reserved", "<u4")])
    t = ctx.thrashing_create(layout)
    d_cnt = ctx.alloc(16)
    ms, times, det = ctypes.c_float(0), [], 0
    for it in range(calls):
        ev = np.zeros(n, ev_dt)
        k = np.tile(np.arange(per, dtype=np.uint64), n_slots)
        ev["slot"] = np.repeat(np.arange(n_slots, dtype=np.uint32), per)
        ev["time_stamp"] = np.uint64(1000) + (np.uint64(it * per) + k) * np.uint64(20000) + rng.integers(0, 5000, n).astype(np.uint64)
        ev["first_page"] = rng.integers(0, 4, n).astype(np.uint16) * 128 + rng.integers(0, 8, n).astype(np.uint16)
        ev["num_pages"] = pages
        ev["processor"] = (k & np.uint64(1)).astype(np.uint16)
        ev["kind"] = (rng.random(n) < 0.15).astype(np.uint8)
        d_ev = ctx.to_device(ev)
        lib.uvmtier_event_record(ctx.handle, 920)
        st = lib.uvmtier_thrashing_note_events_device(ctx.handle, ctypes.byref(t), ctypes.c_uint32(n), P(d_ev), P(d_cnt))
        lib.uvmtier_event_record(ctx.handle, 921)
        assert st == 0, hex(st)
        lib.uvmtier_event_elapsed_ms(ctx.handle, 920, 921, ctypes.byref(ms))
        cnt = ctx.from_device(d_cnt, 4, "<u4")
        assert cnt[2] == 0
        det += int(cnt[0])
        if it >= 2:
            times.append(ms.value)
        ctx.free(d_ev)
    masks = ctx.from_device(t.d_thrashing_pages, n_slots * 16, "<u4")
    thrashing_pages = int(np.unpackbits(masks.view(np.uint8)).sum())
    ctx.thrashing_destroy(t)
    ctx.free(d_cnt)
    m = float(np.mean(times))
    alg = n * pages * 24 + n * 24 + n_slots * 2 * 96       # page word + processors read and written, events, va_block rows
    return {"events": n, "va_blocks": n_slots, "pages_per_event": pages, "ms": m, "events_per_s": n / (m * 1e-3),
            "algorithmic_bytes": alg, "GBps": alg / (m * 1e-3) / 1e9, "detections": det, "thrashing_pages": thrashing_pages,
            "what": "two processors alternate on 16-page regions every 20 us (window 500 us, threshold 3)"}


def peer_numbers(ctx, lib, P, rank, world, dist, region_bytes=1 << 30):
    """BASELINE config E, the NVLink leg: every GPU moves a 1 GiB page list (4 KB pages, random placement) to the
    next GPU of the ring (push: the source GPU's SMs store into the peer's HBM, uvm_va_block.c:3427-3457) and from
    it (pull), all ranks at once.  Each rank's HBM arena has a source half, filled with the fork's test pattern
    (tests/cxl_p2p_test.c:467-476: byte i = (i + seed) & 0xFF, plus the page number in the first 8 bytes so that a
    misplaced page cannot pass), and a landing half, filled with 0xEE before every leg.  After each leg the receiver
    checksums the pages that landed and compares them with the sender's checksums of the pages it sent; a mismatch
    fails the run.  The peak the copies are compared with is measured in the same run on the same arenas:
    cudaMemcpyAsync between the two GPUs (the copy engine, what ce_hal->memcopy does), 1 GiB, best of 3."""
    import torch
    from uvmtier import TIER_HBM, TIER_PEER0
    info = ctx.tier_create(TIER_HBM, 2 * region_bytes)
    hnd = (ctypes.c_uint8 * 64)()
    st = lib.uvmtier_tier_ipc_export(ctx.handle, TIER_HBM, hnd)
    assert st == 0, hex(st)
    mine = torch.tensor(list(bytes(hnd)), dtype=torch.uint8, device="cuda")
    allh = [torch.zeros_like(mine) for _ in range(world)]
    dist.all_gather(allh, mine)
    nxt, prv = (rank + 1) % world, (rank - 1) % world
    peer = (ctypes.c_uint8 * 64)(*allh[nxt].cpu().tolist())
    st = lib.uvmtier_tier_ipc_import(ctx.handle, TIER_PEER0, peer, ctypes.c_uint64(2 * region_bytes))
    assert st == 0, hex(st)

    def barrier():
        ctx.synchronize(); dist.barrier(); ctx.synchronize()

    # the copy-engine peak, every rank at once like the legs below (contents do not matter yet)
    barrier()
    peak_push = ctx.tier_memcpy_peak_gbs(TIER_PEER0, TIER_HBM, region_bytes, 3)
    barrier()
    peak_pull = ctx.tier_memcpy_peak_gbs(TIER_HBM, TIER_PEER0, region_bytes, 3)
    barrier()

    n_pages = region_bytes // 4096
    perm = np.random.default_rng(7).permutation(n_pages).astype(np.uint64)          # the same on every rank
    s_off = np.arange(n_pages, dtype=np.uint64) * np.uint64(4096)
    d_off = np.uint64(region_bytes) + perm * np.uint64(4096)
    d_s, d_d = ctx.to_device(s_off), ctx.to_device(d_off)
    seed = 17 + rank
    pat = np.tile(np.roll(np.arange(256, dtype=np.uint8), -seed), region_bytes // 256).reshape(n_pages, 4096)
    pat[:, :8] = (np.arange(n_pages, dtype=np.uint64) | (np.uint64(rank + 1) << np.uint64(48))).view(np.uint8).reshape(n_pages, 8)
    base = info.device_ptr

    def chk(st_):
        if st_ != 0:
            raise RuntimeError("peer leg: 0x%x %s" % (st_, lib.uvmtier_last_error().decode()))

    chk(lib.uvmtier_memcpy_h2d(ctx.handle, ctypes.c_void_p(base), P(pat), ctypes.c_size_t(region_bytes)))
    del pat
    cs_src = ctx.checksum_pages(TIER_HBM, s_off, 4096)
    t = torch.from_numpy(cs_src.view(np.int64).copy()).cuda()
    allcs = [torch.zeros_like(t) for _ in range(world)]
    dist.all_gather(allcs, t)
    ms = ctypes.c_float(0)

    def leg(dst, src, expect_from):
        chk(lib.uvmtier_memset_device(ctx.handle, ctypes.c_void_p(base + region_bytes), 0xEE, ctypes.c_size_t(region_bytes)))
        barrier()
        best = 1e30
        for r in range(4):
            lib.uvmtier_event_record(ctx.handle, 900)
            chk(lib.uvmtier_copy_page_list_device(ctx.handle, dst, src, ctypes.c_size_t(n_pages), P(d_d), P(d_s), 4096, 0))
            lib.uvmtier_event_record(ctx.handle, 901)
            lib.uvmtier_event_elapsed_ms(ctx.handle, 900, 901, ctypes.byref(ms))
            if r:
                best = min(best, ms.value)
        barrier()
        got = ctx.checksum_pages(TIER_HBM, d_off, 4096)
        want = allcs[expect_from].cpu().numpy().view(np.uint64)
        return region_bytes / (best * 1e-3) / 1e9, bool(np.array_equal(got, want))

    push, ok_push = leg(TIER_PEER0, TIER_HBM, prv)       # my pages land in the next GPU; what landed here came from prv
    pull, ok_pull = leg(TIER_HBM, TIER_PEER0, nxt)       # I fetch the next GPU's pages
    tt = torch.tensor([push, pull, peak_push, peak_pull, float(ok_push), float(ok_pull)], dtype=torch.float64, device="cuda")
    dist.all_reduce(tt, op=dist.ReduceOp.MIN)
    push, pull, peak_push, peak_pull, ok_push, ok_pull = [float(x) for x in tt.tolist()]
    out = {"page_size": 4096, "order": "random permutation", "bytes_per_gpu": region_bytes, "pattern": "fork pattern (i + seed) & 0xFF + page number",
           "nvlink_peer_push_ring": {"GBs_per_gpu_min": round(push, 2), "GBs_aggregate": round(push * world, 2),
                                     "copy_engine_peak_GBs": round(peak_push, 2), "frac_of_measured_peak": round(push / peak_push, 3),
                                     "nominal_per_direction_GBs": 900, "verified": ok_push == 1.0},
           "nvlink_peer_pull_ring": {"GBs_per_gpu_min": round(pull, 2), "copy_engine_peak_GBs": round(peak_pull, 2),
                                     "frac_of_measured_peak": round(pull / peak_pull, 3), "verified": ok_pull == 1.0},
           "peak": "cudaMemcpyAsync GPU to GPU, 1 GiB, best of 3, all ranks at once (minimum over ranks)"}
    lib.uvmtier_tier_destroy(ctx.handle, TIER_PEER0)
    barrier()
    ctx.free(d_s); ctx.free(d_d)
    ctx.tier_destroy(TIER_HBM)
    if not (ok_push == 1.0 and ok_pull == 1.0):
        raise RuntimeError("NVLink peer copy moved wrong bytes: %s" % json.dumps(out))
    return out


def other_trace_numbers(ctx, lib, P, trace, steps=5, warmup=3):
    """The same device-resident step (K1 + service pass, L2 flushed, CUDA events on the library's stream) on another
    address distribution of the 1M-entry trace: `stream` (sequential pages, 30-70 % duplicates, a few hundred va_blocks
    with long runs) or `hotset` (90 % of the faults on 1 % of the va_blocks)."""
    from uvmtier import layouts as L
    wl = build_workload(1, trace)
    n = len(wl["packets"]); nb = (n + BATCH - 1) // BATCH; ncap = nb * BATCH
    ctx.fault_buffer_init(wl["gpu_params"], wl["table"])
    ring_host = np.zeros((n + 1, 8), np.uint32); ring_host[:n] = wl["packets"]
    d_ring0, d_ring = ctx.to_device(ring_host), ctx.alloc(ring_host.nbytes)
    d_disp, d_ord, d_grp, d_bat = ctx.alloc(ncap * 4), ctx.alloc(ncap * 16), ctx.alloc(ncap * 16), ctx.alloc(nb * 32)
    d_flags, d_svc, d_runs, d_cnt = ctx.alloc(ncap * 2), ctx.alloc(ncap * 96), ctx.alloc(RUN_CAP * 8), ctx.alloc(8)
    ds0 = ctx.block_state_alloc(wl["layout"]); ds0.upload(wl["state"])
    ds = ctx.block_state_alloc(wl["layout"])
    state_bytes = wl["n_slots"] * (6 * 64 + 4 + 2 + 1)
    d_flush = ctx.alloc(L2_FLUSH_BYTES)
    new_get, ms = ctypes.c_uint32(0), ctypes.c_float(0)
    t_k1, t_svc = [], []
    for it in range(warmup + steps):
        lib.uvmtier_memcpy_d2d_async(ctx.handle, P(d_ring), P(d_ring0), ctypes.c_size_t(ring_host.nbytes))
        lib.uvmtier_memcpy_d2d_async(ctx.handle, ctypes.c_void_p(ds.ptrs.gpu_resident), ctypes.c_void_p(ds0.ptrs.gpu_resident),
                                     ctypes.c_size_t(state_bytes))
        lib.uvmtier_memset_device(ctx.handle, P(d_flush), 0, ctypes.c_size_t(L2_FLUSH_BYTES))
        lib.uvmtier_event_record(ctx.handle, 930)
        st1 = lib.uvmtier_fetch_preprocess_fault_batches_device(ctx.handle, P(d_ring), ctypes.c_uint32(n + 1), ctypes.c_uint32(0),
                                                                ctypes.c_uint32(n), ctypes.c_int(1), P(d_disp), P(d_ord), P(d_grp),
                                                                P(d_bat), None, ctypes.byref(new_get))
        lib.uvmtier_event_record(ctx.handle, 931)
        st2 = lib.uvmtier_service_fault_batches_device(ctx.handle, ctypes.c_uint32(nb), P(d_ord), P(d_grp), P(d_bat), P(wl["layout"]),
                                                       ctypes.byref(ds.ptrs), P(d_flags), P(d_svc), P(d_runs), ctypes.c_uint32(RUN_CAP),
                                                       P(d_cnt))
        lib.uvmtier_event_record(ctx.handle, 932)
        assert st1 == 0 and st2 == 0, (hex(st1), hex(st2))
        lib.uvmtier_event_elapsed_ms(ctx.handle, 930, 931, ctypes.byref(ms)); a = ms.value
        lib.uvmtier_event_elapsed_ms(ctx.handle, 931, 932, ctypes.byref(ms)); b = ms.value
        if it >= warmup:
            t_k1.append(a); t_svc.append(b)
    bat = ctx.from_device(d_bat, nb, L.BATCH_DT)
    cnt = ctx.from_device(d_cnt, 2, "<u4")
    k1, svc = float(np.mean(t_k1)), float(np.mean(t_svc))
    out = {"ms_per_step": k1 + svc, "k_fault_fetch_preprocess_ms": k1, "service_pass_ms": svc, "entries_per_s": n / ((k1 + svc) * 1e-3),
           "coalesced_entries": int(bat["num_coalesced_faults"].sum()), "va_block_runs": int(bat["num_block_groups"].sum()),
           "copy_runs": int(cnt[0]), "steps": steps}
    for p_ in (d_ring0, d_ring, d_disp, d_ord, d_grp, d_bat, d_flags, d_svc, d_runs, d_cnt, d_flush):
        ctx.free(p_)
    ds.free(); ds0.free()
    return out


# ----------------------------------------------------------------------------------------------- GPU arm
_REAL_STDOUT = None


def emit(line):
    sys.stdout.flush()
    data = (json.dumps(line) + "\n").encode()
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, data)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--trace", default="uniform", choices=["uniform", "stream", "hotset"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-migration", action="store_true")
    args = ap.parse_args()

    # stdout carries exactly one JSON line: libraries that print banners on fd 1 (NCCL's version line) are sent to
    # stderr, and the line is written to the real stdout at the end
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)

    rank, world, local_rank = env_int("RANK", 0), env_int("WORLD_SIZE", 1), env_int("LOCAL_RANK", 0)
    if args.impl == "reference":
        run_reference_arm(args, rank)
        return
    args.warmup = max(args.warmup, 3)

    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    import uvmtier
    from uvmtier import P, blockstate as BS, layouts as L
    lib = uvmtier.lib()
    ctx = uvmtier.Context(local_rank)
    h = ctx.handle
    wl = build_workload(1, args.trace, rank, world)
    n = len(wl["packets"])                  # this rank's share of the world x 1M-entry trace
    nb = (n + BATCH - 1) // BATCH
    ncap = nb * BATCH                       # per-entry outputs are sliced by batch: batch b at b * BATCH
    ctx.fault_buffer_init(wl["gpu_params"], wl["table"])
    layout = wl["layout"]

    def chk(st, what):
        if st != 0:
            raise RuntimeError("%s failed: 0x%x %s" % (what, st, lib.uvmtier_last_error().decode()))

    # ---- device-resident buffers ----
    ring_host = np.zeros((n + 1, 8), np.uint32)
    ring_host[:n] = wl["packets"]
    d_ring0 = ctx.to_device(ring_host)           # pristine copy
    d_ring = ctx.alloc(ring_host.nbytes)
    d_disp, d_ord, d_grp, d_bat = ctx.alloc(ncap * 4), ctx.alloc(ncap * 16), ctx.alloc(ncap * 16), ctx.alloc(nb * 32)
    d_flags, d_svc, d_runs, d_cnt = ctx.alloc(ncap * 2), ctx.alloc(ncap * 96), ctx.alloc(RUN_CAP * 8), ctx.alloc(8)
    ds0 = ctx.block_state_alloc(layout); ds0.upload(wl["state"])       # pristine state
    ds = ctx.block_state_alloc(layout)
    state_bytes = wl["n_slots"] * (6 * 64 + 4 + 2 + 1)
    d_flush = ctx.alloc(L2_FLUSH_BYTES)
    new_get = ctypes.c_uint32(0)

    def prepare():
        """Untimed: restore the ring's VALID bits and the va_block state, then push everything out of L2."""
        chk(lib.uvmtier_memcpy_d2d_async(h, P(d_ring), P(d_ring0), ctypes.c_size_t(ring_host.nbytes)), "d2d")
        chk(lib.uvmtier_memcpy_d2d_async(h, ctypes.c_void_p(ds.ptrs.gpu_resident), ctypes.c_void_p(ds0.ptrs.gpu_resident),
                                         ctypes.c_size_t(state_bytes)), "d2d")
        chk(lib.uvmtier_memset_device(h, P(d_flush), 0, ctypes.c_size_t(L2_FLUSH_BYTES)), "flush")

    def step(ev):
        chk(lib.uvmtier_event_record(h, ev), "event")
        chk(lib.uvmtier_fetch_preprocess_fault_batches_device(h, P(d_ring), ctypes.c_uint32(n + 1), ctypes.c_uint32(0),
                                                              ctypes.c_uint32(n), ctypes.c_int(1), P(d_disp), P(d_ord),
                                                              P(d_grp), P(d_bat), None, ctypes.byref(new_get)), "fetch_preprocess")
        chk(lib.uvmtier_event_record(h, ev + 1), "event")
        chk(lib.uvmtier_service_fault_batches_device(h, ctypes.c_uint32(nb), P(d_ord), P(d_grp), P(d_bat), P(layout),
                                                     ctypes.byref(ds.ptrs), P(d_flags), P(d_svc), P(d_runs),
                                                     ctypes.c_uint32(RUN_CAP), P(d_cnt)), "service")
        chk(lib.uvmtier_event_record(h, ev + 2), "event")

    def barrier():
        ctx.synchronize()
        if dist is not None:
            dist.barrier()
        ctx.synchronize()

    sampler = ClockSampler(local_rank)
    sampler.start()
    for _ in range(args.warmup):
        prepare(); step(0)
    barrier()
    launches0 = uvmtier.kernel_launch_count()
    wall0 = time.perf_counter()
    steps = args.steps
    assert 3 * steps <= 890
    for i in range(steps):
        prepare(); step(3 * i)
    barrier()
    wall1 = time.perf_counter()
    launches = uvmtier.kernel_launch_count() - launches0
    ms = ctypes.c_float(0)
    t_step, t_k1, t_svc = [], [], []
    for i in range(steps):
        chk(lib.uvmtier_event_elapsed_ms(h, 3 * i, 3 * i + 2, ctypes.byref(ms)), "elapsed"); t_step.append(ms.value)
        chk(lib.uvmtier_event_elapsed_ms(h, 3 * i, 3 * i + 1, ctypes.byref(ms)), "elapsed"); t_k1.append(ms.value)
        chk(lib.uvmtier_event_elapsed_ms(h, 3 * i + 1, 3 * i + 2, ctypes.byref(ms)), "elapsed"); t_svc.append(ms.value)
    ms_step = float(np.mean(t_step))

    # ---- counts for the algorithmic bytes ----
    bat = ctx.from_device(d_bat, nb, L.BATCH_DT)
    cnt = ctx.from_device(d_cnt, 2, "<u4")
    n_coalesced = int(bat["num_coalesced_faults"].sum()); n_groups = int(bat["num_block_groups"].sum()); n_runs = int(cnt[0])
    assert n_runs <= RUN_CAP, "copy-run buffer too small"
    k1_bytes = n * 32 + n * 4 + n_coalesced * 16 + n_groups * 16 + nb * 32
    svc_rec = ctx.from_device(d_svc, ncap, BS.SERVICE_GROUP_DT)
    written = (np.arange(ncap) % BATCH) < np.repeat(bat["num_block_groups"], BATCH)     # records the pass wrote
    live = written & (svc_rec["page_fault_count"] > 0)
    touched = len(np.unique(svc_rec["slot"][live]))
    svc_bytes = n_groups * (16 + 96 + 8) + n_coalesced * (16 + 2) + touched * 2 * (6 * 64 + 8) + n_runs * 8
    k1_ms, svc_ms = float(np.mean(t_k1)), float(np.mean(t_svc))

    # ---- e2e: host-buffer entry point for stage A/B with pinned memory (H2D + D2H in the timed region), then the
    #      service pass on the device-resident sorted view, then the D2H of what a caller consumes (records, runs) ----
    old_affinity = os.sched_getaffinity(0)
    binding = bind_near_gpu(local_rank)

    def pinned(arr):
        p = ctypes.c_void_p()
        chk(lib.uvmtier_mem_alloc_host_pinned(h, ctypes.c_size_t(max(arr.nbytes, 1)), ctypes.byref(p)), "pinned")
        v = np.frombuffer((ctypes.c_char * arr.nbytes).from_address(p.value), dtype=arr.dtype).reshape(arr.shape)
        v[...] = arr
        return v
    hp_ring = pinned(ring_host)
    hp_bat = pinned(np.zeros(nb, L.BATCH_DT)); hp_flags = pinned(np.zeros(ncap, "<u2")); hp_cnt = pinned(np.zeros(2, "<u4"))
    hp_svc = pinned(np.zeros(ncap, BS.SERVICE_GROUP_DT)); hp_runs = pinned(np.zeros(max(n_runs, 1), BS.COPY_RUN_DT))
    # What comes back to the host every step is what the reference's caller of service_fault_batch looks at
    # afterwards: the per-batch headers (status, duplicate / invalid-prefetch / fatal counts -> cancel + replay
    # policy), the per-fault flags (which entries to cancel and how) and the run counter.  Service records and copy
    # runs are consumed on the GPU (uvmtier_copy_runs_device), the ring's VALID bits are the host's own to clear
    # for [get, new_get).  "full" additionally reads ring, records and runs back (the r1_a..e definition).
    h2d = ring_host.nbytes
    d2h_result = nb * 32 + n * 2 + 8
    d2h_full = d2h_result + ring_host.nbytes + n * 96 + n_runs * 8

    class Readback(ctypes.Structure):
        _fields_ = [("h_batches", ctypes.c_void_p), ("h_entry_flags", ctypes.c_void_p), ("h_counters", ctypes.c_void_p)]
    readback = Readback(hp_bat.ctypes.data, hp_flags.ctypes.data, hp_cnt.ctypes.data)
    # 8 chunks on a quiet host; with 8 rings sharing the host's memory the copy is the long pole (23 GB/s per GPU instead of
    # 54) and finer chunks hide more of the rest behind it: measured 1.295 ms with 8 chunks, 1.206 ms with 16 at N = 8
    E2E_CHUNKS = max(1, int(os.environ.get("UVMTIER_E2E_CHUNKS", "16" if world >= 8 else "8")))
    E2E_MODE = os.environ.get("UVMTIER_E2E_MODE", "")        # diagnostics only: "copy", "fetch", "norb" (no copy-back)

    def e2e_step(full):
        hp_ring[:n] = wl["packets"]                                   # untimed: restore VALID bits / state
        chk(lib.uvmtier_memcpy_d2d_async(h, ctypes.c_void_p(ds.ptrs.gpu_resident), ctypes.c_void_p(ds0.ptrs.gpu_resident),
                                         ctypes.c_size_t(state_bytes)), "d2d")
        chk(lib.uvmtier_memset_device(h, P(d_flush), 0, ctypes.c_size_t(L2_FLUSH_BYTES)), "flush")
        ctx.synchronize()
        t0 = time.perf_counter()
        if E2E_MODE == "copy":          # diagnostic: the ring's H2D copy alone
            chk(lib.uvmtier_memcpy_h2d_async(h, P(d_ring), P(hp_ring), ctypes.c_size_t(ring_host.nbytes)), "h2d")
        elif E2E_MODE == "fetch":       # diagnostic: chunked copy + preprocess, no servicing
            chk(lib.uvmtier_fetch_preprocess_fault_batches_streamed(
                h, P(hp_ring), P(d_ring), ctypes.c_uint32(n + 1), ctypes.c_uint32(0), ctypes.c_uint32(n), ctypes.c_int(1),
                ctypes.c_uint32((nb + E2E_CHUNKS - 1) // E2E_CHUNKS), P(d_disp), P(d_ord), P(d_grp), P(d_bat), None,
                ctypes.byref(new_get)), "fetch_preprocess_streamed")
        else:
            # the ring crosses PCIe from pinned host memory in chunks; each chunk is preprocessed and serviced as soon as
            # it is there, and its batch headers + flags go back on a third stream while the next chunk is worked on
            chk(lib.uvmtier_service_replayable_faults_streamed(
                h, P(hp_ring), P(d_ring), ctypes.c_uint32(n + 1), ctypes.c_uint32(0), ctypes.c_uint32(n), ctypes.c_int(1),
                ctypes.c_uint32((nb + E2E_CHUNKS - 1) // E2E_CHUNKS), P(d_disp), P(d_ord), P(d_grp), P(d_bat), P(layout),
                ctypes.byref(ds.ptrs), P(d_flags), P(d_svc), P(d_runs), ctypes.c_uint32(RUN_CAP), P(d_cnt), ctypes.byref(readback) if E2E_MODE != "norb" else None,
                ctypes.byref(new_get)), "service_replayable_faults_streamed")
        if full:
            chk(lib.uvmtier_memcpy_d2h_async(h, P(hp_ring), P(d_ring), ctypes.c_size_t(ring_host.nbytes)), "d2h ring")
            chk(lib.uvmtier_memcpy_d2h_async(h, P(hp_svc), P(d_svc), ctypes.c_size_t(ncap * 96)), "d2h")
            chk(lib.uvmtier_memcpy_d2h_async(h, P(hp_runs), P(d_runs), ctypes.c_size_t(max(n_runs, 1) * 8)), "d2h")
        ctx.synchronize()
        return time.perf_counter() - t0

    for _ in range(args.warmup):
        e2e_step(False)
    barrier()
    wall2 = time.perf_counter()
    e2e_t = [e2e_step(False) for _ in range(max(3, min(steps, 10)))]
    barrier()
    wall3 = time.perf_counter()
    e2e_ms = float(np.mean(e2e_t)) * 1e3
    if not E2E_MODE:
        assert np.array_equal(hp_bat["num_coalesced_faults"], bat["num_coalesced_faults"]), "e2e and device paths disagree"
        assert int(hp_cnt[0]) == n_runs
    e2e_full_ms = float(np.mean([e2e_step(True) for _ in range(3)])) * 1e3
    # the floor of the end-to-end number: this step's ring alone crossing the host link, all ranks at once (the N rings
    # share the host's memory controllers and root ports)
    def h2d_alone():
        ctx.synchronize()
        t0 = time.perf_counter()
        chk(lib.uvmtier_memcpy_h2d_async(h, P(d_ring), P(hp_ring), ctypes.c_size_t(ring_host.nbytes)), "h2d")
        ctx.synchronize()
        return time.perf_counter() - t0
    h2d_alone()
    barrier()
    h2d_floor_ms = float(np.mean([h2d_alone() for _ in range(5)])) * 1e3
    barrier()
    os.sched_setaffinity(0, old_affinity)

    migration = None
    if not args.no_migration:
        migration = migration_numbers(ctx, lib, P, rank, world, dist)
        if dist is not None and world > 1:
            migration["nvlink_peer"] = peer_numbers(ctx, lib, P, rank, world, dist)
    oversub = None
    if not args.no_migration and rank == 0:
        oversub = oversubscription_numbers(ctx, lib, P, rank)
    access = None
    if not args.no_migration and rank == 0:
        access = access_counter_numbers(ctx, lib, P, rank)
    thrash = None
    if not args.no_migration and rank == 0:
        thrash = thrashing_numbers(ctx, lib, P, rank)
    nonrep = None
    if not args.no_migration and rank == 0:
        nonrep = nonreplayable_numbers(ctx, lib, P, rank)
    other = None
    if not args.no_migration and world == 1 and args.trace == "uniform":
        other = {t: other_trace_numbers(ctx, lib, P, t) for t in ("stream", "hotset")}
    clocks = sampler.stop(windows=((wall0, wall1), (wall2, wall3)))

    # ---- max over ranks ----
    if dist is not None:
        import torch
        t = torch.tensor([ms_step, e2e_ms, k1_ms, svc_ms, h2d_floor_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_step, e2e_ms, k1_ms, svc_ms, h2d_floor_ms = [float(x) for x in t.tolist()]
        gathered = [None] * world
        dist.all_gather_object(gathered, binding)
        binding = gathered

    peak, peak_src = peaks()
    dom_is_k1 = k1_ms >= svc_ms
    dom_bytes, dom_ms = (k1_bytes, k1_ms) if dom_is_k1 else (svc_bytes, svc_ms)
    achieved = dom_bytes / (dom_ms * 1e-3) / 1e9
    # DRAM bytes per launch from the committed ncu capture (dram__bytes_read.sum + dram__bytes_write.sum summed over
    # the pass's kernels; uniform trace only -- other traces were not captured)
    traffic, traffic_src = None, None
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        if args.trace == "uniform":
            traffic = tr["k_fault_fetch_preprocess_dram_bytes_per_launch" if dom_is_k1 else "service_pass_dram_bytes_per_launch"]
            traffic_src = tr["source"]
    except Exception:
        pass
    line = {
        "metric": METRIC, "value": world * N_ENTRIES / (ms_step * 1e-3), "unit": UNIT, "n_gpus": world, "steps": steps,
        "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u64", "data": "synthetic",
        "config": config_dict(args.trace, world),
        "e2e": {"value": world * N_ENTRIES / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h_result),
                "ms_per_step": e2e_ms,
                "h2d_floor_ms": h2d_floor_ms, "h2d_floor_GBs_per_gpu": h2d / (h2d_floor_ms * 1e-3) / 1e9,
                "h2d_floor_what": "the step's ring alone, one cudaMemcpyAsync from pinned host memory per rank, all ranks at once, max over ranks",
                "chunks": E2E_CHUNKS, "binding": binding,
                "result": "batch headers + per-fault flags + run counter (what the caller of service_fault_batch consumes)",
                "full_readback": {"value": n / (e2e_full_ms * 1e-3), "ms_per_step": e2e_full_ms, "d2h_bytes_per_step": int(d2h_full),
                                  "what": "also ring, service records and copy runs (rank-local, not max over ranks)"}},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": {"bound": "hbm", "kernel": "k_fault_fetch_preprocess" if dom_is_k1 else "service pass (k_service_count, k_service_alloc, k_service_fill, k_service_sort_lists/_big, k_service_blocks_grp, k_service_blocks, k_service_unmapped)",
                     "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                     "traffic_unit": "bytes of DRAM traffic per launch (ncu)", "traffic_source": traffic_src,
                     "peak_source": peak_src, "algorithmic_bytes_per_launch": int(dom_bytes), "ms_per_launch": dom_ms,
                     "frac_of_nominal_8TBs": achieved / 8000.0},
        "kernels": {"k_fault_fetch_preprocess_ms": k1_ms, "k_fault_fetch_preprocess_GBs": k1_bytes / (k1_ms * 1e-3) / 1e9,
                    "service_pass_ms": svc_ms, "service_pass_GBs": svc_bytes / (svc_ms * 1e-3) / 1e9,
                    "coalesced_entries": n_coalesced, "va_block_runs": n_groups, "copy_runs": n_runs, "va_blocks_touched": touched},
        "migration": migration,
        "oversubscription": oversub,
        "access_counters": access,
        "thrashing": thrash,
        "non_replayable_faults": nonrep,
        "other_traces": other,
        "wall_s_timed_region": wall1 - wall0,
    }
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        v1, kind, ns = cpu_arm(wl, 1, reps=2)
        line["cpu_baseline"] = {"value": v1, "unit": UNIT, "cores": 1, "kind": kind,
                                "sample": "full step (%d entries), best of 2, 1 thread (the reference services one GPU on one thread); "
                                          "fetch/merge/sort, entry loop, prefetch hint and make_resident = the reference's own C "
                                          "when kind is reference, the CPU restatement when it is port" % ns}
    if rank == 0:
        emit(line)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    ctx.close()


if __name__ == "__main__":
    main()
