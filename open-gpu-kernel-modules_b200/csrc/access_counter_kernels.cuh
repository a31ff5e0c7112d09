// access_counter_kernels.cuh -- access-counter notifications: fetch + parse + sort (one CTA per batch) and the
// per-va_block servicing (one warp per va_block).  Device code only (host side: access_counters.cu); also compiled by
// tests/simt for the CPU emulator.
//
// Reference (UVM = kernel-open/nvidia-uvm): uvm_gpu_access_counters.c fetch_access_counter_buffer_entries :862-940,
// preprocess_notifications :974-992 (comparators :798-860), expand_notification_block :1287-1354,
// service_notifications_in_block :1356-1427, service_va_block_locked :1079-1249; packet format clc365.h:33-80, parser
// uvm_turing_access_counter_buffer.c:105-198.
// The two halves live in different translation units (fault_kernels.cuh defines kernels with external linkage and can
// be included by one .cu only): define UVMTIER_AC_PREPROCESS and / or UVMTIER_AC_SERVICE before including.
#pragma once

#ifdef UVMTIER_AC_PREPROCESS
#include "fault_kernels.cuh"
#endif
#ifdef UVMTIER_AC_SERVICE
#include "prefetch.cuh"
#include "service_kernels.cuh"
#endif

namespace uvmtier_k {

struct ac_kparams {
    uint32_t B;              // max_batch_size
    uint32_t va_bits;
    uint32_t ring_entries, get, n_entries;
    uint32_t tbl_n, idx_bits, clear_valid;
};

// uvm_hal_turing_access_counter_buffer_parse_entry: the fields of one NVC365 packet (8 words)
__device__ __forceinline__ void ac_decode(const uint32_t *w, uint32_t va_bits, uvmtier_access_counter_entry_t *e)
{
    // ADDR = words 2..3, sign-extended to the canonical form (uvm_parent_gpu_canonical_address, uvm_gpu.c:428-459)
    uint64_t addr = (uint64_t)w[2] | ((uint64_t)w[3] << 32);
    const uint32_t sh = 64u - va_bits;
    addr = (uint64_t)((int64_t)(addr << sh) >> sh);
    e->address = addr;
    // INST = bits 12.. of word 0 and word 1: the 4 KB page number of the instance block
    e->instance_ptr = (((uint64_t)(w[0] >> 12)) | ((uint64_t)w[1] << 20)) << 12;
    const uint32_t ap = (w[0] >> 10) & 3u;                              // INST_APERTURE: 0 VID, 2 / 3 SYS
    e->instance_aperture = ap == 0 ? 10u : 9u;                          // UVM_APERTURE_VID / UVM_APERTURE_SYS
    e->mmu_engine_id = (w[7] >> 20) & 0x1ffu;
    e->ve_id = (e->mmu_engine_id - 384u) & 0xffu;                       // get_ve_id, graphics engine (uvm_hopper_fault_buffer.c:31-44)
    e->counter_value = w[6] & 0xffffu;
    e->sub_granularity = w[4];
    e->bank = (w[0] >> 2) & 0xfu;
    e->tag = w[5] & 0xfffffu;
    e->valid = w[7] >> 31;
}

#ifdef UVMTIER_AC_PREPROCESS
// One CTA per batch, one thread per ring entry of the batch; blockDim.x = T = the power of two >= B.
template <int T>
__device__ __forceinline__ void ac_fetch_preprocess_body(const ac_kparams &prm, uvmtier_access_counter_packet_t *ring,
                                                         const uvmtier_instance_ptr_entry_t *tbl, uvmtier_ordered_notification_t *ordered,
                                                         uvmtier_fault_group_t *groups, uvmtier_fault_batch_t *batches,
                                                         uvmtier_access_counter_entry_t *entries, uint64_t *s_sort0, uint64_t *s_sort1,
                                                         uint64_t *s_addr, uint32_t *s_sub, uint16_t *s_ctx, uint16_t *s_gstart,
                                                         uvmtier_instance_ptr_entry_t *s_rows)
{
    __shared__ uint64_t s_inst0;
    __shared__ uint32_t s_warp_count[32], s_first_invalid, s_single, s_ve0, s_ap0;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, W = T / 32;
    const uint32_t b = blockIdx.x, base = b * prm.B;
    const uint32_t in_batch = prm.n_entries - base < prm.B ? prm.n_entries - base : prm.B;
    for (uint32_t k = tid; k < prm.tbl_n; k += T)
        s_rows[k] = tbl[k];
    if (tid == 0) { s_first_invalid = in_batch; s_single = 1; }
    __syncthreads();

    // fetch: entries up to the first one whose VALID bit is clear (:893-907, mode BATCH_READY)
    uvmtier_access_counter_entry_t e;
    e.valid = 0; e.address = 0; e.instance_ptr = 0; e.instance_aperture = 0; e.ve_id = 0; e.sub_granularity = 0;
    uint32_t ridx = 0;
    const bool in = (uint32_t)tid < in_batch;
    if (in) {
        ridx = (prm.get + base + (uint32_t)tid) % prm.ring_entries;
        const uint4 lo = ((const uint4 *)ring[ridx].w)[0], hi = ((const uint4 *)ring[ridx].w)[1];
        const uint32_t w[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
        ac_decode(w, prm.va_bits, &e);
        if (!e.valid)
            atomicMin(&s_first_invalid, (uint32_t)tid);
    }
    if (tid == 0) { s_inst0 = e.instance_ptr; s_ve0 = e.ve_id; s_ap0 = e.instance_aperture; }
    __syncthreads();
    const uint32_t n = s_first_invalid;
    const bool live = (uint32_t)tid < n;
    if (live && prm.clear_valid)
        ring[ridx].w[7] &= 0x7fffffffu;                                 // parse_entry clears VALID (:196-197)
    if (entries && in) {
        if (!live) e.valid = 0;
        entries[base + tid] = e;
    }

    // translate_notifications_instance_ptrs (:942-972): (instance_ptr, ve_id) -> table row; a failed translation
    // leaves va_space NULL, which sorts first and is ignored afterwards
    int ctx = -1;
    uint32_t rank = 0;
    if (live) {
        for (uint32_t k = 0; k < prm.tbl_n; k++) {
            const uvmtier_instance_ptr_entry_t r = s_rows[k];
            if (r.instance_ptr == e.instance_ptr && r.aperture == e.instance_aperture &&
                (r.ve_id == UVMTIER_VE_ID_ANY || r.ve_id == e.ve_id)) {
                ctx = (int)k;
                rank = (uint32_t)r.sort_rank + 1u;
                break;
            }
        }
        // is_single_instance_ptr (:925-930): every entry has the (instance_ptr, ve_id) of the first one
        if (e.instance_ptr != s_inst0 || e.ve_id != s_ve0 || e.instance_aperture != s_ap0)
            s_single = 0;                                               // benign race: everybody writes 0
    }
    s_ctx[tid] = (uint16_t)(ctx < 0 ? UVMTIER_AC_CTX_NONE : ctx);
    s_addr[tid] = e.address;
    s_sub[tid] = e.sub_granularity;

    // key = rank | page number | notification_cache index: cmp_sort_notifications_by_va_space_gpu_address (:834-851)
    // with ties in cache order (the kernel sort() is not stable; same canonical choice as the fault path)
    const uint32_t ib = prm.idx_bits;
    uint64_t key = ~0ull;
    if (live)
        key = ((uint64_t)rank << (ib + kPageBits)) | (((e.address >> 12) & ((1ull << kPageBits) - 1)) << ib) | (uint64_t)tid;
    key = fault_sort_keys<T>(key, tid, s_sort0, s_sort1);
    __syncthreads();
    s_sort0[tid] = key;
    __syncthreads();

    // outputs + va_block runs (service_notifications_batch :1540-1560: same va_space, same gpu, address <= block end)
    const bool have = (uint32_t)tid < n;
    const uint32_t idxmask = (1u << ib) - 1u;
    uint32_t idx = 0;
    bool boundary = false;
    if (have) {
        idx = (uint32_t)key & idxmask;
        uvmtier_ordered_notification_t o;
        o.address = s_addr[idx];
        o.index = (uint16_t)idx;
        o.ctx = s_ctx[idx];
        o.sub_granularity = s_sub[idx];
        ordered[base + tid] = o;
        if (o.ctx != UVMTIER_AC_CTX_NONE)
            boundary = tid == 0 || (s_sort0[tid - 1] >> (ib + 9)) != (key >> (ib + 9));
    }
    const uint32_t bb = __ballot_sync(0xffffffffu, boundary);
    if (lane == 0)
        s_warp_count[warp] = (uint32_t)__popc(bb);
    __syncthreads();
    uint32_t gidx = (uint32_t)__popc(bb & ((1u << lane) - 1u)), ngroups = 0;
    for (int ww = 0; ww < W; ww++) {
        const uint32_t c = s_warp_count[ww];
        if (ww < warp) gidx += c;
        ngroups += c;
    }
    if (boundary)
        s_gstart[gidx] = (uint16_t)tid;
    __syncthreads();
    if (boundary) {
        const uint32_t next = gidx + 1 < ngroups ? s_gstart[gidx + 1] : n;
        uvmtier_fault_group_t g;
        g.block_start = s_addr[idx] & ~((uint64_t)UVMTIER_VA_BLOCK_SIZE - 1);
        g.ctx = s_ctx[idx];
        g.first_fault_index = (uint16_t)tid;
        g.block_faults = (uint16_t)(next - (uint32_t)tid);
        g.flags = 0;
        groups[base + gidx] = g;
    }
    if (tid == 0) {
        uvmtier_fault_batch_t bt;
        bt.num_cached_faults = n;
        bt.num_coalesced_faults = n;
        bt.is_single_instance_ptr = s_single;
        bt.max_utlb_id = 0;
        bt.status = NV_OK;
        bt.num_duplicate_faults = 0;
        bt.num_invalid_prefetch_faults = 0;
        bt.num_block_groups = ngroups;
        batches[b] = bt;
    }
}

UVMTIER_KERNEL void __launch_bounds__(kFaultMaxThreads)
k_access_counters_fetch_preprocess(ac_kparams prm, uvmtier_access_counter_packet_t *ring, const uvmtier_instance_ptr_entry_t *tbl,
                                   uvmtier_ordered_notification_t *ordered, uvmtier_fault_group_t *groups,
                                   uvmtier_fault_batch_t *batches, uvmtier_access_counter_entry_t *entries)
{
    // one set of arrays for every CTA size (static shared memory of the instantiations would add up)
    __shared__ uint64_t s_sort0[kFaultMaxThreads], s_sort1[kFaultMaxThreads], s_addr[kFaultMaxThreads];
    __shared__ uint32_t s_sub[kFaultMaxThreads];
    __shared__ uint16_t s_ctx[kFaultMaxThreads], s_gstart[kFaultMaxThreads];
    __shared__ uvmtier_instance_ptr_entry_t s_rows[kFaultMaxTableRows];
#define AC_BODY(T) ac_fetch_preprocess_body<T>(prm, ring, tbl, ordered, groups, batches, entries, s_sort0, s_sort1, s_addr, s_sub, \
                                               s_ctx, s_gstart, s_rows)
    switch (blockDim.x) {
        case 32: AC_BODY(32); break;
        case 64: AC_BODY(64); break;
        case 128: AC_BODY(128); break;
        case 256: AC_BODY(256); break;
        case 512: AC_BODY(512); break;
        default: AC_BODY(1024); break;
    }
#undef AC_BODY
}
#endif // UVMTIER_AC_PREPROCESS

#ifdef UVMTIER_AC_SERVICE
struct ac_service_cfg {
    uint32_t granularity_2m;     // tracking granularity is 2 MB (anything else is ignored, :1303-1306)
    uint32_t cpu_chunk_2m;       // CPU chunks are 2 MB: a notification expands to its sub-granularity regions
};

// One warp per va_block of `slots[0 .. *n_list)`; the block's runs (list[offset[slot] ..], batch order) one after the
// other.  Masks: one 32-bit word per lane, lanes 16-31 mirror lanes 0-15 (as k_service_blocks / k_make_resident).
UVMTIER_KERNEL void __launch_bounds__(kSvcThreads)
k_access_counters_blocks(service_kparams prm, ac_service_cfg cfg, const uvmtier_ordered_notification_t *__restrict__ ordered,
                         const uvmtier_fault_group_t *__restrict__ groups, block_state_ptrs st, const uint2 *__restrict__ cw,
                         const uint32_t *__restrict__ offset, const uint32_t *__restrict__ list, const uint32_t *__restrict__ slots,
                         const uint32_t *__restrict__ n_list, uvmtier_service_group_t *__restrict__ service,
                         uvmtier_copy_run_t *__restrict__ runs, uint32_t *__restrict__ counters)
{
    const int lane = threadIdx.x & 31, l = lane & 15, warp = threadIdx.x >> 5;
    const uint32_t n_warps = gridDim.x * (kSvcThreads / 32), n = *n_list;
    for (uint32_t q = blockIdx.x * (kSvcThreads / 32) + warp; q < n; q += n_warps) {
        const uint32_t slot = slots[q], cnt = cw[slot].x, off = offset[slot];
        const size_t w = (size_t)slot * 16 + l;
        uint32_t gres = st.gpu_resident[w], c0 = st.cpu_resident[0][w], c1 = st.cpu_resident[1][w];
        uint32_t pte0 = st.gpu_pte[0][w], pte1 = st.gpu_pte[1][w], pte2 = st.gpu_pte[2][w];
        const uint32_t logical = st.logical_prot[slot], pref = st.preferred_location[slot];
        uvmtier_prefetch_info_t pi = st.prefetch_info[slot];
        for (uint32_t k = 0; k < cnt; k++) {
            const uint32_t t = list[off + k], b = svc_batch_of(prm, t), base = b * prm.B;
            const uvmtier_fault_group_t grp = groups[t];
            // expand_notification_block for every notification of the run (:1387-1400)
            uint32_t accessed = 0;
            for (uint32_t i = 0; cfg.granularity_2m && i < grp.block_faults; i++) {
                const uvmtier_ordered_notification_t o = ordered[base + grp.first_fault_index + i];
                const uint32_t page = (uint32_t)((o.address - grp.block_start) >> 12) & 511u;
                const int wl = (int)(page >> 5);
                const uint32_t bit = page & 31u;
                const uint32_t cpu = c0 | c1;
                // uvm_va_block_page_get_closest_resident (:1307): the GPU itself -> stale; nowhere -> nothing to move
                const bool on_gpu = (__shfl_sync(HW_FULL, gres, wl) >> bit) & 1u, on_cpu = (__shfl_sync(HW_FULL, cpu, wl) >> bit) & 1u;
                if (on_gpu || !on_cpu)
                    continue;
                if (!cfg.cpu_chunk_2m) {
                    accessed |= l == wl ? 1u << bit : 0u;                 // :1325-1326
                }
                else {
                    // sub-granularity region r = pages [16 r, 16 r + 16) (:1335-1346), then only what the CPU holds
                    accessed |= ((o.sub_granularity >> (2 * l)) & 1u ? 0x0000ffffu : 0u) |
                                ((o.sub_granularity >> (2 * l + 1)) & 1u ? 0xffff0000u : 0u);
                    accessed &= cpu;                                      // :1351
                }
            }
            // service_va_block_locked: only while the GPU maps something in the block (:1103-1104); pages already
            // on the GPU are stale (:1122-1123)
            if (!hw_any(pte0, lane))
                accessed = 0;
            uint32_t newres = accessed & ~gres;
            const uint32_t page_count = hw_weight(newres);
            // ---- uvm_va_block_service_locked, as in k_service_blocks ----
            uint32_t hint = UVMTIER_ID_INVALID, pf = 0, n_pf = 0, n0 = 0, n1 = 0, nt = 0, run_off = 0, run_cnt = 0;
            int rf = 0, ro = 0;
            if (page_count) {                                             // uniform across the warp
                rf = hw_first_bit(newres, lane);
                ro = hw_last_bit(newres, lane) + 1;
                const bool block_empty = !hw_any(gres | c0 | c1, lane);
                if (prm.pf_enable) {
                    const uint32_t flags = UVMTIER_PF_HAS_RESIDENT_MASK | UVMTIER_PF_GPU_VA_SPACE_REGISTERED |
                                           (block_empty ? UVMTIER_PF_BLOCK_RESIDENT_EMPTY : 0u);
                    pf = pf_get_hint_block(newres, gres, 0u, 0u, false, grp.block_start, 512, rf, ro, prm.gpu_id, pref, flags,
                                           16u, prm.pf_threshold, prm.pf_min_faults, &pi, &hint, lane);
                }
                if (hint != UVMTIER_ID_INVALID) {                         // uvm_va_block_get_prefetch_hint :11856-11876
                    n_pf = hw_weight(pf);
                    newres |= pf;
                    rf = hw_first_bit(newres, lane);
                    ro = hw_last_bit(newres, lane) + 1;
                }
                else {
                    pf = 0;
                }
                uint32_t todo = newres & ~gres;
                const uint32_t from0 = todo & c0;
                todo &= ~from0;
                const uint32_t from1 = todo & c1;
                const uint32_t touch = todo & ~from1;
                n0 = hw_weight(from0); n1 = hw_weight(from1); nt = hw_weight(touch);
                const uint32_t r0 = svc_emit_runs(from0, pf, slot, 0u, NULL, 0u, 0u, lane);
                const uint32_t r1 = svc_emit_runs(from1, pf, slot, 1u, NULL, 0u, 0u, lane);
                const uint32_t r2 = svc_emit_runs(touch, pf, slot, UVMTIER_RUN_SRC_ZERO, NULL, 0u, 0u, lane);
                run_cnt = r0 + r1 + r2;
                if (lane == 0 && run_cnt)
                    run_off = atomicAdd(&counters[0], run_cnt);
                run_off = __shfl_sync(HW_FULL, run_off, 0);
                svc_emit_runs(from0, pf, slot, 0u, runs, run_off, prm.run_cap, lane);
                svc_emit_runs(from1, pf, slot, 1u, runs, run_off + r0, prm.run_cap, lane);
                svc_emit_runs(touch, pf, slot, UVMTIER_RUN_SRC_ZERO, runs, run_off + r0 + r1, prm.run_cap, lane);
                const uint32_t arriving = newres & ~gres;                 // unmapped before they move (:4919-4926)
                pte0 &= ~arriving; pte1 &= ~arriving; pte2 &= ~arriving;
                gres |= newres;
                c0 &= ~newres;
                c1 &= ~newres;
                pte0 |= newres;                                           // service_finish: map the GPU
                if (logical == PROT_RWA) {
                    pte1 |= newres;
                    pte2 |= newres;
                }
            }
            if (lane < 16) {
                uint32_t *rec = (uint32_t *)&service[t];
                if (l < 8) {
                    uint32_t v = 0;
                    switch (l) {
                        case 0: v = slot; break;
                        case 1: v = (uint32_t)rf | ((uint32_t)ro << 16); break;
                        case 2: v = page_count | (n_pf << 16); break;
                        case 3: v = n0 | (n1 << 16); break;
                        case 4: v = nt | (hint << 16); break;
                        case 5: v = run_off; break;
                        case 6: v = run_cnt | ((b + prm.batch_base) << 16); break;
                        case 7: v = t - base; break;
                    }
                    rec[l] = v;
                }
                rec[8 + l] = page_count ? newres : 0u;
            }
        }
        if (lane < 16) {
            st.gpu_resident[w] = gres; st.cpu_resident[0][w] = c0; st.cpu_resident[1][w] = c1;
            st.gpu_pte[0][w] = pte0; st.gpu_pte[1][w] = pte1; st.gpu_pte[2][w] = pte2;
            if (st.gpu_evicted)
                st.gpu_evicted[w] &= ~gres;
            if (st.read_duplicated)
                st.read_duplicated[w] &= c0 | c1;
            if (l == 0)
                st.prefetch_info[slot] = pi;
        }
    }
}

#endif // UVMTIER_AC_SERVICE

} // namespace uvmtier_k
