// service_tpb.cuh -- the servicing pass with ONE THREAD per va_block (replaces the warp-per-block kernel on
// the hot path).  Same semantics as k_service_blocks in service_kernels.cuh, same reference citations.
//
// Why: a va_block's runs must be applied one after the other (each one sees the residency, mapping and prefetch
// state the previous one left), so the work per block is a sequential chain whatever the formulation.  With a
// warp per block 31/32 of the scalar bookkeeping was redundant and the pass was instruction-bound (ncu r1_c).
// Here a thread keeps its block's six 512-bit masks in registers (16 words each, all loops unrolled, no shuffles),
// a warp advances 32 independent blocks at once, and the prefetch tree is evaluated on the thread's own words:
// SWAR per word for node sizes <= 32, plain adds of the 16 per-word popcounts above that.  Service-pass blocks are
// whole, 2 MB-aligned va_blocks, so the tree has offset 0, 512 leaves and no clamped node.
//
// The per-slot run lists must be in ascending (batch, run) order: k_service_sort_lists sorts them first
// (registers for <= 32 runs, shared memory up to 4096, in-place odd-even transposition beyond that).
#pragma once

#include "service_kernels.cuh"

namespace uvmtier_k {

constexpr int kTpbThreads = 64;
constexpr uint32_t kSortSmemKeys = 4096;

// ---- levels of <= 32 leaves inside one word (same arithmetic as pf_tree_select, offset 0) ----
__device__ __forceinline__ uint32_t pf_word_levels(uint32_t w, uint32_t fw, uint32_t m0, uint32_t m1, uint32_t m2, uint32_t m3,
                                                   uint32_t m4, uint32_t m5)
{
    const uint32_t c2 = w - ((w >> 1) & 0x55555555u);
    const uint32_t c2e = c2 & 0x33333333u, c2o = (c2 >> 2) & 0x33333333u;
    const uint32_t c4 = c2e + c2o;
    const uint32_t c8 = (c4 + (c4 >> 4)) & 0x0f0f0f0fu;
    const uint32_t c16 = (c8 + (c8 >> 8)) & 0x00ff00ffu;
    const uint32_t c32 = (c16 + (c16 >> 16)) & 0x3fu;
    const uint32_t a2 = (fw | (fw >> 1)) & 0x55555555u;
    const uint32_t a4 = (a2 | (a2 >> 2)) & 0x11111111u;
    const uint32_t a8 = (a4 | (a4 >> 4)) & 0x01010101u;
    const uint32_t a16 = (a8 | (a8 >> 8)) & 0x00010001u;
    const uint32_t s1 = m0 <= 1u ? (w & fw) : 0u;
    const uint32_t d2e = ((c2e + (8u - m1) * 0x11111111u) >> 3) & 0x11111111u;
    const uint32_t d2o = ((c2o + (8u - m1) * 0x11111111u) >> 3) & 0x11111111u;
    const uint32_t s2 = (d2e & a2) | ((d2o & (a2 >> 2)) << 2);
    const uint32_t s4 = (((c4 + (8u - m2) * 0x11111111u) >> 3) & 0x11111111u) & a4;
    const uint32_t s8 = (((c8 + (128u - m3) * 0x01010101u) >> 7) & 0x01010101u) & a8;
    const uint32_t s16 = (((c16 + (0x8000u - m4) * 0x00010001u) >> 15) & 0x00010001u) & a16;
    const uint32_t s32 = (c32 >= m5 && fw != 0u) ? 0xffffffffu : 0u;
    return s1 | ((s2 << 2) - s2) | ((s4 << 4) - s4) | ((s8 << 8) - s8) | ((s16 << 16) - s16) | s32;
}

// Prefetch hint of a whole aligned va_block on one thread (uvm_perf_prefetch_prenotify_fault_migrations :327-413
// + :477-500, GPU destination, no thrashing).  f = faulted (new_residency) words, gres = resident on the GPU.
// Writes the prefetch mask to out[], updates *pi, returns the hint residency.
__device__ __forceinline__ uint32_t pf_get_hint_thread(const uint32_t (&f)[16], const uint32_t (&gres)[16], bool block_empty,
                                                       uint32_t gpu_id, uint32_t pref, uint32_t thr, uint32_t min_faults,
                                                       uvmtier_prefetch_info_t *pi, uint32_t (&out)[16])
{
    if (pi->last_migration_proc_id != gpu_id) {                            // :340-343
        pi->last_migration_proc_id = (uint16_t)gpu_id;
        pi->fault_migrations_to_last_proc = 0;
    }
    const bool first_touch = block_empty && gpu_id == pref;                 // :363-366
    const uint32_t m0 = thr / 100u + 1u, m1 = (2u * thr) / 100u + 1u, m2 = (4u * thr) / 100u + 1u,
                   m3 = (8u * thr) / 100u + 1u, m4 = (16u * thr) / 100u + 1u, m5 = (32u * thr) / 100u + 1u;
    uint32_t cnt[16];
    uint32_t anyw = 0, weight = 0;
#pragma unroll
    for (int w = 0; w < 16; w++) {
        // tree = resident | faulted, every 64 KB big page holding a fault is filled (:164-220, 64 KB = 16 pages)
        uint32_t pages = gres[w] | f[w];
        pages |= (f[w] & 0xffffu) ? 0xffffu : 0u;
        pages |= (f[w] >> 16) ? 0xffff0000u : 0u;
        out[w] = pf_word_levels(pages, f[w], m0, m1, m2, m3, m4, m5);
        cnt[w] = __popc(pages);
        anyw |= (f[w] != 0u) ? (1u << w) : 0u;
        weight += __popc(f[w]);
    }
    // nodes of 64, 128, 256, 512 leaves: sums of the per-word populations
#pragma unroll
    for (int span = 2; span <= 16; span <<= 1) {
#pragma unroll
        for (int j = 0; j < 16; j += span) {
            uint32_t c = 0;
#pragma unroll
            for (int k = 0; k < span; k++) c += cnt[j + k];
            const uint32_t wm = ((1u << span) - 1u) << j;
            if ((anyw & wm) && c * 100u > (uint32_t)(32 * span) * thr) {
#pragma unroll
                for (int k = 0; k < span; k++) out[j + k] = 0xffffffffu;
            }
        }
    }
    uint32_t pending = 0;
#pragma unroll
    for (int w = 0; w < 16; w++) {
        if (first_touch) out[w] = 0xffffffffu;
        out[w] &= ~f[w];                                                    // :392
        pending += __popc(out[w]);
    }
    pi->fault_migrations_to_last_proc = (uint16_t)(pi->fault_migrations_to_last_proc + weight);   // :410
    return (pi->fault_migrations_to_last_proc >= min_faults && pending > 0) ? pi->last_migration_proc_id
                                                                            : (uint32_t)UVMTIER_ID_INVALID;
}

// Copy runs of mask m (maximal same-cause sequences of consecutive pages): counts them, and with runs != NULL
// writes them starting at runs[at].  UVM/uvm_va_block.c:4242-4335.
__device__ __forceinline__ uint32_t tpb_runs(const uint32_t (&m)[16], const uint32_t (&pf)[16], uint32_t slot, uint32_t src,
                                             uvmtier_copy_run_t *runs, uint32_t at, uint32_t run_cap)
{
    uint32_t start[16], stop[16];
    uint32_t total = 0, cm = 0, cp = 0;
#pragma unroll
    for (int w = 0; w < 16; w++) {
        const uint32_t mp = (m[w] << 1) | cm, pp = (pf[w] << 1) | cp;
        start[w] = m[w] & (~mp | (pf[w] ^ pp));
        stop[w] = ~m[w] | start[w];
        total += __popc(start[w]);
        cm = m[w] >> 31;
        cp = pf[w] >> 31;
    }
    if (!runs || !total)
        return total;
    // first stop position at or after the start of each following word
    uint32_t next_abs[16];
    uint32_t nxt = 512u;
#pragma unroll
    for (int w = 15; w >= 0; w--) {
        next_abs[w] = nxt;
        if (stop[w]) nxt = (uint32_t)(32 * w + __ffs((int)stop[w]) - 1);
    }
    uint32_t k = at;
#pragma unroll
    for (int w = 0; w < 16; w++) {
        uint32_t s_bits = start[w];
        while (s_bits) {
            const int s = __ffs((int)s_bits) - 1;
            s_bits &= s_bits - 1;
            const uint32_t rest = s == 31 ? 0u : (stop[w] >> (s + 1));
            const uint32_t first = (uint32_t)(32 * w + s);
            const uint32_t end = rest ? first + (uint32_t)__ffs((int)rest) : next_abs[w];
            if (k < run_cap) {
                uvmtier_copy_run_t r;
                r.slot = slot;
                r.first_page = (uint16_t)first;
                r.num_pages = (uint16_t)(end - first);
                r.src = (uint16_t)src;
                r.prefetch = (uint16_t)((pf[w] >> s) & 1u);
                r.pad = 0;
                runs[k] = r;
            }
            k++;
        }
    }
    return total;
}

// ------------------------------------------------------------------------------------------------
// Sorts every slot's run list ascending.  One warp per slot.
__global__ void __launch_bounds__(32)
k_service_sort_lists(uint32_t n_slots, const uint32_t *__restrict__ count, const uint32_t *__restrict__ offset,
                     uint32_t *__restrict__ list)
{
    UVMTIER_DYN_SMEM(raw);
    uint32_t *s_keys = (uint32_t *)raw;
    const int lane = threadIdx.x;
    for (uint32_t slot = blockIdx.x; slot < n_slots; slot += gridDim.x) {
        const uint32_t cnt = count[slot];
        if (cnt < 2)
            continue;
        uint32_t *l = list + offset[slot];
        if (cnt <= 32) {
            uint32_t key = (uint32_t)lane < cnt ? l[lane] : 0xffffffffu;
            for (int k = 2; k <= 32; k <<= 1)
                for (int j = k >> 1; j > 0; j >>= 1) {
                    const uint32_t other = __shfl_xor_sync(0xffffffffu, key, j);
                    const bool up = (lane & k) == 0, keep_min = ((lane & j) == 0) == up;
                    key = (keep_min == (key < other)) ? key : other;
                }
            if ((uint32_t)lane < cnt)
                l[lane] = key;
        }
        else if (cnt <= kSortSmemKeys) {
            uint32_t P = 64;
            while (P < cnt) P <<= 1;
            for (uint32_t i = lane; i < P; i += 32) s_keys[i] = i < cnt ? l[i] : 0xffffffffu;
            __syncwarp();
            for (uint32_t k = 2; k <= P; k <<= 1)
                for (uint32_t j = k >> 1; j > 0; j >>= 1) {
                    for (uint32_t i = lane; i < P / 2; i += 32) {
                        const uint32_t a = 2 * i - (i & (j - 1)), b = a + j;     // pair (a, a + j), bit j of a clear
                        const uint32_t x = s_keys[a], y = s_keys[b];
                        const bool up = (a & k) == 0;
                        if ((x > y) == up) { s_keys[a] = y; s_keys[b] = x; }
                    }
                    __syncwarp();
                }
            for (uint32_t i = lane; i < cnt; i += 32) l[i] = s_keys[i];
            __syncwarp();
        }
        else {
            // rare: more runs on one va_block than the shared-memory sorter holds -- odd-even transposition in place
            for (uint32_t pass = 0; pass < cnt; pass++) {
                for (uint32_t i = (pass & 1) + 2 * lane; i + 1 < cnt; i += 64) {
                    const uint32_t x = l[i], y = l[i + 1];
                    if (x > y) { l[i] = y; l[i + 1] = x; }
                }
                __threadfence_block();
                __syncwarp();
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// One thread per state slot; the slot's run list is sorted.
__global__ void __launch_bounds__(kTpbThreads)
k_service_blocks_tpb(service_kparams prm, const uvmtier_ordered_fault_t *__restrict__ ordered,
                     const uvmtier_fault_group_t *__restrict__ groups, uvmtier_fault_batch_t *__restrict__ batches,
                     block_state_ptrs st, const uint32_t *__restrict__ count, const uint32_t *__restrict__ offset,
                     const uint32_t *__restrict__ list, uint16_t *__restrict__ entry_flags,
                     uvmtier_service_group_t *__restrict__ service, uvmtier_copy_run_t *__restrict__ runs,
                     uint32_t *__restrict__ counters)
{
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= prm.n_slots)
        return;
    const uint32_t cnt = count[slot];
    if (!cnt)
        return;
    const uint32_t off = offset[slot];
    uint32_t gres[16], c0[16], c1[16], p0[16], p1[16], p2[16];
    {
        const uint4 *g = (const uint4 *)(st.gpu_resident + (size_t)slot * 16), *a = (const uint4 *)(st.cpu_resident[0] + (size_t)slot * 16),
                    *b = (const uint4 *)(st.cpu_resident[1] + (size_t)slot * 16), *x = (const uint4 *)(st.gpu_pte[0] + (size_t)slot * 16),
                    *y = (const uint4 *)(st.gpu_pte[1] + (size_t)slot * 16), *z = (const uint4 *)(st.gpu_pte[2] + (size_t)slot * 16);
#pragma unroll
        for (int q = 0; q < 4; q++) {
            uint4 v;
            v = g[q]; gres[4 * q] = v.x; gres[4 * q + 1] = v.y; gres[4 * q + 2] = v.z; gres[4 * q + 3] = v.w;
            v = a[q]; c0[4 * q] = v.x; c0[4 * q + 1] = v.y; c0[4 * q + 2] = v.z; c0[4 * q + 3] = v.w;
            v = b[q]; c1[4 * q] = v.x; c1[4 * q + 1] = v.y; c1[4 * q + 2] = v.z; c1[4 * q + 3] = v.w;
            v = x[q]; p0[4 * q] = v.x; p0[4 * q + 1] = v.y; p0[4 * q + 2] = v.z; p0[4 * q + 3] = v.w;
            v = y[q]; p1[4 * q] = v.x; p1[4 * q + 1] = v.y; p1[4 * q + 2] = v.z; p1[4 * q + 3] = v.w;
            v = z[q]; p2[4 * q] = v.x; p2[4 * q + 1] = v.y; p2[4 * q + 2] = v.z; p2[4 * q + 3] = v.w;
        }
    }
    const uint32_t logical = st.logical_prot[slot], pref = st.preferred_location[slot];
    uvmtier_prefetch_info_t pi = st.prefetch_info[slot];

    for (uint32_t k = 0; k < cnt; k++) {
        const uint32_t t = list[off + k];
        const uint32_t b = t / prm.B, base = b * prm.B;
        const uvmtier_fault_group_t grp = groups[t];
        const uint32_t first = grp.first_fault_index, n = grp.block_faults;

        // ---- service_fault_batch_block_locked :1433-1580 ----
        uint32_t newres[16];
#pragma unroll
        for (int w = 0; w < 16; w++) newres[w] = 0;
        uint32_t page_faults = 0, dup_sum = 0, ip_sum = 0, prev_page = 0xffffffffu;
        bool prev_fatal = false, prev_ip = false;
        for (uint32_t i = 0; i < n; i++) {
            const uvmtier_ordered_fault_t e = ordered[base + first + i];
            const uint32_t acc = e.fault_access_type & 7u;
            const uint32_t page = (uint32_t)((e.fault_address - grp.block_start) >> 12);
            const bool dup = i > 0 && page == prev_page;
            uint32_t service_acc = acc, fl = 0;
            bool entry_fatal = false, entry_ip = false, done = false;
            dup_sum += dup ? e.num_instances : (uint32_t)e.num_instances - 1u;   // :1200-1203
            if (dup) {
                fl |= UVMTIER_SF_DUPLICATE;
                if (prev_ip) {
                    entry_ip = true;
                    ip_sum += e.num_instances;
                }
                if (!prev_fatal)
                    done = true;
            }
            if (!done && svc_prot_of(acc) > logical) {                // check_fault_access_permissions :1293-1360
                if (acc == ACC_PREFETCH) {
                    entry_ip = true;
                    ip_sum += e.num_instances;
                    done = true;
                }
                else {
                    uint32_t cancel = CANCEL_ALL;
                    service_acc = ACC_COUNT;
                    if (acc > ACC_READ) {
                        cancel = CANCEL_WRITE_AND_ATOMIC;
                        if (e.access_type_mask & (1u << ACC_READ)) {
                            if (svc_prot_of(ACC_READ) <= logical)
                                service_acc = ACC_READ;
                            else
                                cancel = CANCEL_ALL;
                        }
                    }
                    entry_fatal = true;
                    fl |= UVMTIER_SF_FATAL | (REASON_INVALID_PERMISSIONS << UVMTIER_SF_REASON_SHIFT) | (cancel << UVMTIER_SF_CANCEL_SHIFT);
                    if (service_acc == ACC_COUNT)
                        done = true;
                }
            }
            if (!done) {
                const uint32_t prot = svc_prot_of(service_acc), wi = page >> 5, bit = 1u << (page & 31u);
                uint32_t word = 0;
#pragma unroll
                for (int w = 0; w < 16; w++) {
                    const uint32_t pw = prot == PROT_RWA ? p2[w] : (prot == PROT_RW ? p1[w] : p0[w]);
                    word |= (uint32_t)w == wi ? pw : 0u;
                }
                if (word & bit) {
                    fl |= UVMTIER_SF_AUTHORIZED;                      // :1518-1522
                }
                else {
#pragma unroll
                    for (int w = 0; w < 16; w++) newres[w] |= (uint32_t)w == wi ? bit : 0u;
                    page_faults++;
                    fl |= UVMTIER_SF_SERVICED | (service_acc << UVMTIER_SF_ACCESS_SHIFT);
                }
            }
            if (entry_ip)
                fl |= UVMTIER_SF_INVALID_PREFETCH;
            entry_flags[base + first + i] = (uint16_t)fl;
            prev_fatal = entry_fatal;
            prev_ip = entry_ip;
            prev_page = page;
        }
        if (dup_sum) atomicAdd(&batches[b].num_duplicate_faults, dup_sum);
        if (ip_sum) atomicAdd(&batches[b].num_invalid_prefetch_faults, ip_sum);

        // ---- uvm_va_block_service_locked ----
        uint32_t hint = UVMTIER_ID_INVALID, n_pf = 0, n0 = 0, n1 = 0, nt = 0, run_off = 0, run_cnt = 0;
        int rf = 0, ro = 0;
        if (page_faults) {
            uint32_t pf[16];
            uint32_t occupied = 0;
#pragma unroll
            for (int w = 0; w < 16; w++) occupied |= gres[w] | c0[w] | c1[w];
            if (prm.pf_enable)
                hint = pf_get_hint_thread(newres, gres, occupied == 0u, prm.gpu_id, pref, prm.pf_threshold, prm.pf_min_faults, &pi, pf);
            if (hint != UVMTIER_ID_INVALID) {                         // uvm_va_block_get_prefetch_hint :11856-11876
#pragma unroll
                for (int w = 0; w < 16; w++) {
                    n_pf += __popc(pf[w]);
                    newres[w] |= pf[w];
                }
            }
            else {
#pragma unroll
                for (int w = 0; w < 16; w++) pf[w] = 0;
            }
            // region = span of new_residency (with or without the hint it is first..last set bit)
            rf = 512; ro = 0;
#pragma unroll
            for (int w = 15; w >= 0; w--)
                if (newres[w]) rf = 32 * w + __ffs((int)newres[w]) - 1;
#pragma unroll
            for (int w = 0; w < 16; w++)
                if (newres[w]) ro = 32 * w + 32 - __clz((int)newres[w]);
            // block_copy_resident_pages :4660-4847
            uint32_t from0[16], from1[16], touch[16];
#pragma unroll
            for (int w = 0; w < 16; w++) {
                uint32_t todo = newres[w] & ~gres[w];
                from0[w] = todo & c0[w];
                todo &= ~from0[w];
                from1[w] = todo & c1[w];
                touch[w] = todo & ~from1[w];
                n0 += __popc(from0[w]); n1 += __popc(from1[w]); nt += __popc(touch[w]);
            }
            const uint32_t r0 = n0 ? tpb_runs(from0, pf, slot, 0u, NULL, 0u, 0u) : 0u;
            const uint32_t r1 = n1 ? tpb_runs(from1, pf, slot, 1u, NULL, 0u, 0u) : 0u;
            const uint32_t r2 = nt ? tpb_runs(touch, pf, slot, UVMTIER_RUN_SRC_ZERO, NULL, 0u, 0u) : 0u;
            run_cnt = r0 + r1 + r2;
            if (run_cnt) {
#ifdef UVMTIER_SIMT_EMU
                run_off = atomicAdd(&counters[0], run_cnt);
#else
                // One atomic per converged group of lanes instead of one per thread (a single counter would
                // serialise a million atomics).  The lanes of `act` execute this together; lanes of a warp that are
                // elsewhere in their own block's loop are simply not part of it.
                const uint32_t act = __activemask();
                const int lane = threadIdx.x & 31;
                uint32_t before = 0, total = 0;
                for (int src = 0; src < 32; src++) {
                    const uint32_t v = __shfl_sync(act, run_cnt, src);
                    if ((act >> src) & 1u) {
                        if (src < lane) before += v;
                        total += v;
                    }
                }
                const int leader = __ffs((int)act) - 1;
                uint32_t basep = 0;
                if (lane == leader)
                    basep = atomicAdd(&counters[0], total);
                basep = __shfl_sync(act, basep, leader);
                run_off = basep + before;
#endif
                if (r0) tpb_runs(from0, pf, slot, 0u, runs, run_off, prm.run_cap);
                if (r1) tpb_runs(from1, pf, slot, 1u, runs, run_off + r0, prm.run_cap);
                if (r2) tpb_runs(touch, pf, slot, UVMTIER_RUN_SRC_ZERO, runs, run_off + r0 + r1, prm.run_cap);
            }
            // uvm_va_block_make_resident_finish :5042-5084 and the mapping step (:11019-11066, :12063-12080)
#pragma unroll
            for (int w = 0; w < 16; w++) {
                gres[w] |= newres[w];
                c0[w] &= ~newres[w];
                c1[w] &= ~newres[w];
                p0[w] |= newres[w];
                if (logical == PROT_RWA) {
                    p1[w] |= newres[w];
                    p2[w] |= newres[w];
                }
            }
        }
        // ---- record ----
        {
            uint4 *rec = (uint4 *)&service[t];
            rec[0] = make_uint4(slot, (uint32_t)rf | ((uint32_t)ro << 16), page_faults | (n_pf << 16), n0 | (n1 << 16));
            rec[1] = make_uint4(nt | (hint << 16), run_off, run_cnt | (b << 16), t - base);
#pragma unroll
            for (int q = 0; q < 4; q++)
                rec[2 + q] = page_faults ? make_uint4(newres[4 * q], newres[4 * q + 1], newres[4 * q + 2], newres[4 * q + 3])
                                         : make_uint4(0, 0, 0, 0);
        }
    }
    {
        uint4 *g = (uint4 *)(st.gpu_resident + (size_t)slot * 16), *a = (uint4 *)(st.cpu_resident[0] + (size_t)slot * 16),
              *bq = (uint4 *)(st.cpu_resident[1] + (size_t)slot * 16), *x = (uint4 *)(st.gpu_pte[0] + (size_t)slot * 16),
              *y = (uint4 *)(st.gpu_pte[1] + (size_t)slot * 16), *z = (uint4 *)(st.gpu_pte[2] + (size_t)slot * 16);
#pragma unroll
        for (int q = 0; q < 4; q++) {
            g[q] = make_uint4(gres[4 * q], gres[4 * q + 1], gres[4 * q + 2], gres[4 * q + 3]);
            a[q] = make_uint4(c0[4 * q], c0[4 * q + 1], c0[4 * q + 2], c0[4 * q + 3]);
            bq[q] = make_uint4(c1[4 * q], c1[4 * q + 1], c1[4 * q + 2], c1[4 * q + 3]);
            x[q] = make_uint4(p0[4 * q], p0[4 * q + 1], p0[4 * q + 2], p0[4 * q + 3]);
            y[q] = make_uint4(p1[4 * q], p1[4 * q + 1], p1[4 * q + 2], p1[4 * q + 3]);
            z[q] = make_uint4(p2[4 * q], p2[4 * q + 1], p2[4 * q + 2], p2[4 * q + 3]);
        }
        st.prefetch_info[slot] = pi;
    }
}

} // namespace uvmtier_k
