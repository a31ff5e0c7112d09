#!/usr/bin/env python
"""bench.py -- headline benchmark of the nvidia-uvm tiering hot path on B200 (BASELINE.json).

A step = one pass of the hot path over one synthetic batch of work per GPU:
  1,048,576 replayable-fault packets (uniform over 64 GB of VA, 4 VA spaces, 256-entry batches)
  -> K0/K1 decode + fetch-time coalescing + sort + va_block grouping          (fault entries)
  -> K2 prefetch bitmap-tree hints for the 32,768 va_blocks of that 64 GB VA.
Metric: fault entries/s preprocessed (whole job = all ranks).  `value` is measured with inputs resident
in HBM (CUDA events on the library's stream, L2 flushed between iterations); `e2e` goes through the
host-buffer C-ABI calls with pinned host buffers, H2D and D2H inside the timed region.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--trace uniform|stream|hotset]

Multi-GPU (torchrun): the VA space is sharded, every rank runs its own trace on its own GPU, no data-path
collective; torch.distributed is used for the barrier and the max-over-ranks only.
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
METRIC = "fault entries/s preprocessed (1M-entry trace: decode+coalesce+sort+group by va_block, + prefetch trees over 64 GB VA)"
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


class ClockSampler(threading.Thread):
    """Samples nvidia-smi clocks / throttle reasons while the timed region runs."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.rows = []
        self.proc = None

    def run(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE, text=True)
            for line in self.proc.stdout:
                self.rows.append([c.strip() for c in line.split(",")])
        except Exception:
            pass

    def stop(self):
        if self.proc:
            self.proc.terminate()
        self.join(timeout=2)
        sm, mx, reasons = [], 0, set()
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx = max(mx, float(r[2]))
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                continue
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


# ----------------------------------------------------------------------------------------------- workload
def build_workload(seed, trace):
    from uvmtier import tracegen as TG
    from uvmtier.layouts import PREFETCH_BLOCK_DT, PREFETCH_INFO_DT
    pk, tbl, lay = TG.fault_trace(N_ENTRIES, seed, trace, N_SPACES, VA_BYTES)
    gp = TG.gpu_params(BATCH)
    # block state for the prefetch leg: faulted pages of every va_block = the pages the trace touches,
    # resident = seeded Bernoulli(0.25) minus faulted (pages a GPU faults on are not resident on it)
    w = pk.astype(np.uint64)
    page = ((w[:, 3] << np.uint64(20)) | (w[:, 2] >> np.uint64(12))) - np.uint64(TG.VA_BASE >> 12)
    inst = (w[:, 1] << np.uint64(20)) | (w[:, 0] >> np.uint64(12))
    space = np.searchsorted(tbl["instance_ptr"] >> np.uint64(12), inst)
    gpage = space.astype(np.int64) * (lay["blocks_per_space"] * 512) + page.astype(np.int64)
    faulted = np.zeros(N_BLOCKS * 512, bool)
    faulted[gpage] = True
    faulted = faulted.reshape(N_BLOCKS, 512)
    rs = TG.Stream(seed ^ 0x5EED, N_BLOCKS * 16)
    resident = np.unpackbits((rs.take(N_BLOCKS * 16) & rs.buf[:N_BLOCKS * 16][::-1]).view(np.uint8).reshape(N_BLOCKS, 64),
                             axis=1, bitorder="little").astype(bool) & ~faulted
    blk = np.zeros(N_BLOCKS, PREFETCH_BLOCK_DT)
    blk["start"] = np.uint64(TG.VA_BASE) + (np.arange(N_BLOCKS, dtype=np.uint64) % np.uint64(lay["blocks_per_space"]) << np.uint64(21))
    blk["big_page_size"] = 65536
    blk["num_pages"] = 512
    anyf = faulted.any(axis=1)
    blk["faulted_first"] = np.where(anyf, faulted.argmax(axis=1), 0)
    blk["faulted_outer"] = np.where(anyf, 512 - faulted[:, ::-1].argmax(axis=1), 0)
    blk["new_residency"] = 1
    blk["preferred_location"] = 257
    blk["flags"] = 0x01 | 0x08
    info = np.zeros(N_BLOCKS, PREFETCH_INFO_DT)
    info["last_migration_proc_id"] = 257
    pack = lambda b: np.ascontiguousarray(np.packbits(b, axis=1, bitorder="little")).view("<u8").reshape(N_BLOCKS, 8)
    return dict(packets=pk, table=tbl, gpu_params=gp, blocks=blk, resident=pack(resident), faulted=pack(faulted), info=info)


# ----------------------------------------------------------------------------------------------- CPU arm
def _ref_worker(args):
    """One process: the reference's own C (oracle/_ref) on a shard of whole batches."""
    pk, tbl, gp, blocks, resident, faulted, info, reps = args
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import faultlib as F
    import uvmtier_testlib as T
    rf, rp = T.load_ref("faults"), T.load_ref("prefetch")
    kind = "reference"
    if rf is None or rp is None:
        rf = rp = T.load_oracle()
        kind = "port"
    pfx = "ref" if kind == "reference" else "oracle"
    best = 1e30
    for _ in range(reps):
        t0 = time.perf_counter()
        F.run_fetch_preprocess(rf, pfx, pk, gp, tbl)
        if len(blocks):
            T.run_prefetch_get_hint(rp, pfx, dict(blocks=blocks, resident=resident, faulted=faulted, thrashing=None,
                                                  cpu_read=None, info=info))
        best = min(best, time.perf_counter() - t0)
    return best, kind


def cpu_arm(wl, threads, reps=3, frac=1.0):
    """Times the CPU implementation on `threads` processes, each with a contiguous shard of the sample."""
    import multiprocessing as mp
    n = int(N_ENTRIES * frac) // (BATCH * threads) * (BATCH * threads)
    nblk = int(N_BLOCKS * frac) // threads * threads
    jobs = []
    for t in range(threads):
        a, b = t * n // threads, (t + 1) * n // threads
        c, d = t * nblk // threads, (t + 1) * nblk // threads
        jobs.append((wl["packets"][a:b].copy(), wl["table"], wl["gpu_params"], wl["blocks"][c:d].copy(),
                     wl["resident"][c:d].copy(), wl["faulted"][c:d].copy(), wl["info"][c:d].copy(), reps))
    if threads == 1:
        res = [_ref_worker(jobs[0])]
    else:
        with mp.get_context("fork").Pool(threads) as pool:
            res = pool.map(_ref_worker, jobs)
    t = max(r[0] for r in res)
    return n / t, res[0][1], n, nblk


def run_reference_arm(args, rank, world):
    if rank != 0:
        return
    from uvmtier import tracegen  # noqa: F401  (host-only helpers; no GPU needed)
    wl = build_workload(1, args.trace)
    cores = os.cpu_count() or 1
    for _ in range(args.warmup):
        cpu_arm(wl, cores, reps=1, frac=0.25)
    vals = []
    t0 = time.perf_counter()
    kind = "port"
    for _ in range(args.steps):
        v, kind, n, nblk = cpu_arm(wl, cores, reps=1)
        vals.append(v)
    dt = (time.perf_counter() - t0) / max(1, args.steps)
    value = float(np.mean(vals))
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u64", "data": "synthetic",
            "config": {"workload": workload_name(args.trace), "cpu": "reference C (oracle/_ref) on host cores" if kind == "reference" else "CPU port (oracle)"},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind,
                             "sample": "%d entries + %d va_blocks per step, sharded over %d processes" % (n, nblk, cores)},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


def workload_name(trace):
    return ("1xB200 config B: 1M-entry %s fault trace (4 VA spaces, batch 256) sort/dedupe/coalesce + prefetch trees over "
            "64 GB VA (32768 va_blocks)" % trace)


# ----------------------------------------------------------------------------------------------- GPU arm
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--trace", default="uniform", choices=["uniform", "stream", "hotset"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank, world, local_rank = env_int("RANK", 0), env_int("WORLD_SIZE", 1), env_int("LOCAL_RANK", 0)
    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return

    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    import uvmtier
    from uvmtier import P, layouts as L
    lib = uvmtier.lib()
    ctx = uvmtier.Context(local_rank)
    h = ctx.handle
    wl = build_workload(1 + rank, args.trace)
    n, nb = N_ENTRIES, N_ENTRIES // BATCH
    ctx.fault_buffer_init(wl["gpu_params"], wl["table"])

    def chk(st, what):
        if st != 0:
            raise RuntimeError("%s failed: 0x%x %s" % (what, st, lib.uvmtier_last_error().decode()))

    # ---- device-resident buffers ----
    ring_host = np.zeros((n + 1, 8), np.uint32)
    ring_host[:n] = wl["packets"]
    d_ring0 = ctx.to_device(ring_host)           # pristine copy
    d_ring = ctx.alloc(ring_host.nbytes)
    d_disp = ctx.alloc(n * 4)
    d_ord = ctx.alloc(n * 16)
    d_grp = ctx.alloc(n * 16)
    d_bat = ctx.alloc(nb * 32)
    d_blocks = ctx.to_device(wl["blocks"]); d_res = ctx.to_device(wl["resident"]); d_f = ctx.to_device(wl["faulted"])
    d_info0 = ctx.to_device(wl["info"]); d_info = ctx.alloc(N_BLOCKS * 4)
    d_pmask = ctx.alloc(N_BLOCKS * 64); d_phint = ctx.alloc(N_BLOCKS * 4)
    d_flush = ctx.alloc(L2_FLUSH_BYTES)
    new_get = ctypes.c_uint32(0)

    def prepare():
        """Untimed: restore the ring's VALID bits and the prefetch counters, then push everything out of L2."""
        chk(lib.uvmtier_memcpy_d2d_async(h, P(d_ring), P(d_ring0), ctypes.c_size_t(ring_host.nbytes)), "d2d")
        chk(lib.uvmtier_memcpy_d2d_async(h, P(d_info), P(d_info0), ctypes.c_size_t(N_BLOCKS * 4)), "d2d")
        chk(lib.uvmtier_memset_device(h, P(d_flush), 0, ctypes.c_size_t(L2_FLUSH_BYTES)), "flush")

    def step(ev):
        chk(lib.uvmtier_event_record(h, ev), "event")
        chk(lib.uvmtier_fetch_preprocess_fault_batches_device(h, P(d_ring), ctypes.c_uint32(n + 1), ctypes.c_uint32(0),
                                                              ctypes.c_uint32(n), ctypes.c_int(1), P(d_disp), P(d_ord),
                                                              P(d_grp), P(d_bat), None, ctypes.byref(new_get)), "fetch_preprocess")
        chk(lib.uvmtier_event_record(h, ev + 1), "event")
        chk(lib.uvmtier_perf_prefetch_get_hint_va_block_device(h, ctypes.c_size_t(N_BLOCKS), P(d_blocks), P(d_res), P(d_f),
                                                               None, None, P(d_info), P(d_pmask), P(d_phint)), "get_hint")
        chk(lib.uvmtier_event_record(h, ev + 2), "event")

    def barrier():
        ctx.synchronize()
        if dist is not None:
            dist.barrier()
        ctx.synchronize()

    for _ in range(args.warmup):
        prepare(); step(0)
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    launches0 = uvmtier.kernel_launch_count()
    wall0 = time.perf_counter()
    steps = args.steps
    assert 3 * steps <= 1000
    for i in range(steps):
        prepare(); step(3 * i)
    barrier()
    wall = time.perf_counter() - wall0
    launches = uvmtier.kernel_launch_count() - launches0
    ms = ctypes.c_float(0)
    t_step, t_k1, t_k2 = [], [], []
    for i in range(steps):
        chk(lib.uvmtier_event_elapsed_ms(h, 3 * i, 3 * i + 2, ctypes.byref(ms)), "elapsed"); t_step.append(ms.value)
        chk(lib.uvmtier_event_elapsed_ms(h, 3 * i, 3 * i + 1, ctypes.byref(ms)), "elapsed"); t_k1.append(ms.value)
        chk(lib.uvmtier_event_elapsed_ms(h, 3 * i + 1, 3 * i + 2, ctypes.byref(ms)), "elapsed"); t_k2.append(ms.value)
    clocks = sampler.stop()
    ms_step = float(np.mean(t_step))

    # ---- counts for the algorithmic bytes of the dominant kernel (K1) ----
    bat = ctx.from_device(d_bat, nb, L.BATCH_DT)
    n_coalesced = int(bat["num_coalesced_faults"].sum()); n_groups = int(bat["num_block_groups"].sum())
    k1_bytes = n * 32 + n * 4 + n_coalesced * 16 + n_groups * 16 + nb * 32
    k2_bytes = N_BLOCKS * (64 + 64 + 24 + 4 + 64 + 4 + 4)
    k1_ms, k2_ms = float(np.mean(t_k1)), float(np.mean(t_k2))

    # ---- e2e: the host-buffer C-ABI calls, pinned host memory, H2D + D2H inside the timed region ----
    def pinned(arr):
        p = ctypes.c_void_p()
        chk(lib.uvmtier_mem_alloc_host_pinned(h, ctypes.c_size_t(max(arr.nbytes, 1)), ctypes.byref(p)), "pinned")
        v = np.frombuffer((ctypes.c_char * arr.nbytes).from_address(p.value), dtype=arr.dtype).reshape(arr.shape)
        v[...] = arr
        return v
    hp_ring = pinned(ring_host); hp_disp = pinned(np.zeros(n, "<u4")); hp_ord = pinned(np.zeros(n, L.ORDERED_DT))
    hp_grp = pinned(np.zeros(n, L.GROUP_DT)); hp_bat = pinned(np.zeros(nb, L.BATCH_DT))
    hp_blocks = pinned(wl["blocks"]); hp_res = pinned(wl["resident"]); hp_f = pinned(wl["faulted"]); hp_info = pinned(wl["info"])
    hp_pmask = pinned(np.zeros((N_BLOCKS, 8), "<u8")); hp_phint = pinned(np.zeros(N_BLOCKS, "<u4"))
    h2d = ring_host.nbytes + wl["blocks"].nbytes + wl["resident"].nbytes + wl["faulted"].nbytes + wl["info"].nbytes
    d2h = ring_host.nbytes + n * 4 + n * 16 + n * 16 + nb * 32 + N_BLOCKS * (64 + 4 + 4)

    def e2e_step():
        hp_ring[:n] = wl["packets"]; hp_info[...] = wl["info"]        # untimed: restore VALID bits / counters
        chk(lib.uvmtier_memset_device(h, P(d_flush), 0, ctypes.c_size_t(L2_FLUSH_BYTES)), "flush")
        ctx.synchronize()
        t0 = time.perf_counter()
        chk(lib.uvmtier_fetch_preprocess_fault_batches(h, P(hp_ring), ctypes.c_uint32(n + 1), ctypes.c_uint32(0),
                                                       ctypes.c_uint32(n), ctypes.c_int(1), P(hp_disp), P(hp_ord), P(hp_grp),
                                                       P(hp_bat), None, ctypes.byref(new_get)), "e2e fetch_preprocess")
        chk(lib.uvmtier_perf_prefetch_get_hint_va_block(h, ctypes.c_size_t(N_BLOCKS), P(hp_blocks), P(hp_res), P(hp_f), None,
                                                        None, P(hp_info), P(hp_pmask), P(hp_phint)), "e2e get_hint")
        return time.perf_counter() - t0

    for _ in range(args.warmup):
        e2e_step()
    barrier()
    e2e_t = [e2e_step() for _ in range(max(3, min(steps, 10)))]
    barrier()
    e2e_ms = float(np.mean(e2e_t)) * 1e3
    assert np.array_equal(hp_bat["num_coalesced_faults"], bat["num_coalesced_faults"]), "e2e and device paths disagree"

    # ---- max over ranks ----
    if dist is not None:
        import torch
        t = torch.tensor([ms_step, e2e_ms, k1_ms, k2_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_step, e2e_ms, k1_ms, k2_ms = [float(x) for x in t.tolist()]

    peak, peak_src = peaks()
    achieved = k1_bytes / (k1_ms * 1e-3) / 1e9
    line = {
        "metric": METRIC, "value": world * n / (ms_step * 1e-3), "unit": UNIT, "n_gpus": world, "steps": steps,
        "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u64", "data": "synthetic",
        "config": {"workload": workload_name(args.trace), "batch": BATCH, "va_spaces": N_SPACES, "l2": "flushed between iterations (256 MiB memset)",
                   "rng": "uvm_test_rng KISS stream, seed 1+rank", "sharding": "one trace + 64 GB VA per GPU, no data-path collective"},
        "e2e": {"value": world * n / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": e2e_ms},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": {"bound": "hbm", "kernel": "k_fault_fetch_preprocess", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": None, "peak_source": peak_src, "algorithmic_bytes_per_launch": int(k1_bytes),
                     "ms_per_launch": k1_ms, "frac_of_nominal_8TBs": achieved / 8000.0},
        "kernels": {"k_fault_fetch_preprocess_ms": k1_ms, "k_prefetch_get_hint_ms": k2_ms,
                    "k_prefetch_get_hint_GBs": k2_bytes / (k2_ms * 1e-3) / 1e9,
                    "coalesced_entries": n_coalesced, "va_block_groups": n_groups},
        "wall_s_timed_region": wall,
    }
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        v1, kind, ns, nblk = cpu_arm(wl, 1, reps=3)
        line["cpu_baseline"] = {"value": v1, "unit": UNIT, "cores": 1, "kind": kind,
                                "sample": "full step (%d entries + %d va_blocks), best of 3, 1 thread (the reference services one GPU single-threaded)" % (ns, nblk)}
    if rank == 0:
        print(json.dumps(line))
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    ctx.close()


if __name__ == "__main__":
    main()
