// service_lists.cuh -- front half of the servicing pass: regroup the va_block runs of all batches by va_block.
//
// The reference walks the batches one after the other, so a va_block sees its runs in batch order.  Here every
// state slot gets the list of its runs (position t = batch * B + run of the group array), ascending:
//   k_service_count     one thread per run: slot of the run; ONE 64-bit atomic per run adds (1, entries) to the
//                       slot's packed (run count, entry count) and returns the run's position in the slot's list
//   k_service_alloc     one thread per slot: list space with one atomic per warp -- lists need not be in slot
//                       order, so no prefix scan over the table is needed
//   k_service_fill      one thread per run: list[offset[slot] + position] = t (no atomics)
//   k_service_sort_*    positions were handed out in arrival order: sort each list, and sort the slots into
//                       classes for the servicing kernels (see k_service_sort_lists)
#pragma once

#include "service_kernels.cuh"

namespace uvmtier_k {

// cw[slot] = (number of runs, number of entries) of the slot, packed little-endian in one 64-bit word.
__global__ void __launch_bounds__(kSvcThreads)
k_service_count(service_kparams prm, const uvmtier_fault_group_t *__restrict__ groups,
                const uvmtier_fault_batch_t *__restrict__ batches, const uvmtier_instance_ptr_entry_t *__restrict__ tbl,
                unsigned long long *__restrict__ cw, uint32_t *__restrict__ batch_unmapped, uint2 *__restrict__ slotpos)
{
    const uint32_t total = prm.n_batches * prm.B, stride = gridDim.x * blockDim.x;
    // two runs per thread and iteration: the loads of both are in flight before either is used (the kernel is a
    // chain of dependent loads: batch header, run, table row, atomic)
    for (uint32_t t0 = blockIdx.x * blockDim.x + threadIdx.x; t0 < total; t0 += 2 * stride) {
        uint32_t t[2], b[2];
        uint32_t status[2], ngroups[2];
        uvmtier_fault_group_t grp[2];
        bool in[2];
#pragma unroll
        for (int u = 0; u < 2; u++) {
            t[u] = t0 + (uint32_t)u * stride;
            in[u] = t[u] < total;
            b[u] = in[u] ? svc_batch_of(prm, t[u]) : 0u;
            status[u] = batches[b[u]].status;
            ngroups[u] = batches[b[u]].num_block_groups;
            grp[u] = groups[in[u] ? t[u] : 0u];
        }
#pragma unroll
        for (int u = 0; u < 2; u++) {
            if (!in[u])
                continue;
            uint2 sp = make_uint2(0xffffffffu, 0u);
            if (status[u] == NV_OK && t[u] - b[u] * prm.B < ngroups[u] && !(grp[u].flags & 1u)) {   // flags bit 0: fatal fault
                const uint32_t slot = svc_slot(prm, tbl[grp[u].ctx].va_space, grp[u].block_start);  // type, skipped (:2294-2303)
                if (slot == 0xffffffffu) {
                    batch_unmapped[b[u]] = 1u;
                }
                else {
                    const unsigned long long old = atomicAdd(&cw[slot], ((unsigned long long)grp[u].block_faults << 32) | 1ull);
                    sp = make_uint2(slot, (uint32_t)old);
                }
            }
            slotpos[t[u]] = sp;
        }
    }
}

// Also counts the light slots (fewer than heavy_min entries) by number of runs: hist[kLightBins - 1 - min(runs,
// kLightBins - 1)], and gives every light slot its rank inside its bin (light_rank[slot]), so that
// k_service_sort_lists can line them up with the longest lists first.  cursor[0] = list space handed out,
// cursor[1] = number of light slots.
constexpr uint32_t kLightBins = 256;
__global__ void __launch_bounds__(256)
k_service_alloc(uint32_t n_slots, const uint2 *__restrict__ cw, uint32_t heavy_min, uint32_t *__restrict__ offset,
                uint32_t *__restrict__ cursor, uint32_t *__restrict__ hist, uint32_t *__restrict__ light_rank,
                const uvmtier_thrashing_block_t *__restrict__ thr_blocks = NULL)
{
    __shared__ uint32_t s_hist[kLightBins], s_cta_base[kLightBins];
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    const uint2 cwv = slot < n_slots ? cw[slot] : make_uint2(0u, 0u);
    const uint32_t c = cwv.x;
    s_hist[threadIdx.x] = 0;
    __syncthreads();
    // the slot's rank inside its bin = this CTA's share of the bin (one global atomic per CTA and bin: thousands of
    // slots share a few dozen bins, per-slot global atomics on them serialise) + the slot's rank inside the CTA
    // va_blocks the thrashing detector tracks page by page take the warp-per-block kernel, which asks for hints
    const bool is_light = c && cwv.y < heavy_min && !(thr_blocks && thr_blocks[slot].tracked);
    const uint32_t bin = kLightBins - 1u - min(c, kLightBins - 1u);
    uint32_t rank = 0;
    if (is_light)
        rank = atomicAdd(&s_hist[bin], 1u);
    __syncthreads();
    if (s_hist[threadIdx.x]) {
        s_cta_base[threadIdx.x] = atomicAdd(&hist[threadIdx.x], s_hist[threadIdx.x]);
        atomicAdd(&cursor[1], s_hist[threadIdx.x]);
    }
    __syncthreads();
    if (is_light)
        light_rank[slot] = s_cta_base[bin] + rank;
    uint32_t incl = c;
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t y = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += y;
    }
    const uint32_t tot = __shfl_sync(0xffffffffu, incl, 31);
    uint32_t base = 0;
    if (lane == 31 && tot)
        base = atomicAdd(cursor, tot);
    base = __shfl_sync(0xffffffffu, base, 31);
    if (slot < n_slots)
        offset[slot] = base + incl - c;
}

__global__ void __launch_bounds__(kSvcThreads)
k_service_fill(uint32_t total, const uint2 *__restrict__ slotpos, const uint32_t *__restrict__ offset, uint32_t *__restrict__ list)
{
    const uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t t0 = blockIdx.x * blockDim.x + threadIdx.x; t0 < total; t0 += 4 * stride) {
        uint2 sp[4];
        uint32_t off[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const uint32_t t = t0 + (uint32_t)u * stride;
            sp[u] = t < total ? slotpos[t] : make_uint2(0xffffffffu, 0u);
        }
#pragma unroll
        for (int u = 0; u < 4; u++) off[u] = sp[u].x != 0xffffffffu ? offset[sp[u].x] : 0u;
#pragma unroll
        for (int u = 0; u < 4; u++)
            if (sp[u].x != 0xffffffffu)
                list[off[u] + sp[u].y] = t0 + (uint32_t)u * stride;
    }
}

// ------------------------------------------------------------------------------------------------
// Sorts every slot's run list ascending and sorts the slots themselves into three classes for the kernels that
// follow:
//   lists of more than kSortWarpKeys runs are left to k_service_sort_big   -> big[0] = how many, big[1..] = slots
//   slots whose runs hold heavy_min or more entries go to k_service_blocks -> heavy[0] = how many, heavy[1..]
//   everything else is serviced by k_service_blocks_grp.
// One warp per slot: up to 32 runs in registers, up to kSortWarpKeys in the warp's slice of shared memory.
constexpr uint32_t kSortWarpKeys = 512;
constexpr int kSortThreads = 256;
static_assert(kLightBins == (uint32_t)kSortThreads, "one thread per bin in the histogram scan");
__global__ void __launch_bounds__(kSortThreads)
k_service_sort_lists(uint32_t n_slots, const uint2 *__restrict__ cw, const uint32_t *__restrict__ offset,
                     uint32_t *__restrict__ list, uint32_t heavy_min, uint32_t *__restrict__ big, uint32_t *__restrict__ heavy,
                     const uint32_t *__restrict__ hist, const uint32_t *__restrict__ light_rank, uint32_t *__restrict__ light,
                     const uvmtier_thrashing_block_t *__restrict__ thr_blocks = NULL)
{
    __shared__ uint32_t s_all[(kSortThreads / 32) * kSortWarpKeys];
    __shared__ uint32_t s_base[kLightBins], s_part[kSortThreads / 32];
    const int lane = threadIdx.x & 31;
    uint32_t *s_keys = s_all + (threadIdx.x >> 5) * kSortWarpKeys;
    const uint32_t n_warps = gridDim.x * (kSortThreads / 32);
    {
        // where each bin of the light list starts: exclusive scan of the histogram (kLightBins == kSortThreads)
        const uint32_t v = hist[threadIdx.x];
        uint32_t x = v;
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, x, d);
            if (lane >= d) x += y;
        }
        if (lane == 31) s_part[threadIdx.x >> 5] = x;
        __syncthreads();
        uint32_t before = 0;
        for (uint32_t w = 0; w < (threadIdx.x >> 5); w++) before += s_part[w];
        s_base[threadIdx.x] = before + x - v;
        __syncthreads();
    }
    // the next slot's counters and list start are loaded while this slot's list is sorted
    uint32_t slot = blockIdx.x * (kSortThreads / 32) + (threadIdx.x >> 5);
    uint2 c = slot < n_slots ? cw[slot] : make_uint2(0u, 0u);
    uint32_t off = slot < n_slots ? offset[slot] : 0u;
    // lane 0: the slot's place in its work list comes back from an atomic; the store that needs it waits until the
    // next slot is taken up, so the sort below does not stall on it
    uint32_t *pend_at = NULL;
    uint32_t pend_slot = 0;
    for (; slot < n_slots;) {
        const uint32_t this_slot = slot, cnt = c.x, this_off = off;
        const uint2 this_c = c;
        slot += n_warps;
        c = slot < n_slots ? cw[slot] : make_uint2(0u, 0u);
        off = slot < n_slots ? offset[slot] : 0u;
        if (pend_at) {
            *pend_at = pend_slot;
            pend_at = NULL;
        }
        if (!cnt)
            continue;
        if (lane == 0) {
            if (this_c.y >= heavy_min || (thr_blocks && thr_blocks[this_slot].tracked)) {
                pend_at = &heavy[1u + atomicAdd(&heavy[0], 1u)];
            }
            else {
                const uint32_t bin = kLightBins - 1u - min(cnt, kLightBins - 1u);
                pend_at = &light[s_base[bin] + light_rank[this_slot]];
            }
            pend_slot = this_slot;
        }
        if (cnt < 2)
            continue;
        uint32_t *l = list + this_off;
        if (cnt <= 32) {
            uint32_t key = (uint32_t)lane < cnt ? l[lane] : 0xffffffffu;
#pragma unroll
            for (int k = 2; k <= 32; k <<= 1)
#pragma unroll
                for (int j = k >> 1; j > 0; j >>= 1) {
                    const uint32_t other = __shfl_xor_sync(0xffffffffu, key, j);
                    const bool up = (lane & k) == 0, keep_min = ((lane & j) == 0) == up;
                    key = (keep_min == (key < other)) ? key : other;
                }
            if ((uint32_t)lane < cnt)
                l[lane] = key;
        }
        else if (cnt <= 64) {
            // two keys per lane (elements lane and lane + 32): the bitonic network of 64, distance 32 inside the lane
            uint32_t k0 = l[lane], k1 = (uint32_t)lane + 32u < cnt ? l[lane + 32] : 0xffffffffu;
#pragma unroll
            for (int k = 2; k <= 64; k <<= 1)
#pragma unroll
                for (int j = k >> 1; j > 0; j >>= 1) {
                    if (j == 32) {
                        const uint32_t lo = min(k0, k1), hi = max(k0, k1);
                        k0 = lo; k1 = hi;
                    }
                    else {
                        const uint32_t o0 = __shfl_xor_sync(0xffffffffu, k0, j), o1 = __shfl_xor_sync(0xffffffffu, k1, j);
                        // element lane + 32 has bit 5 set: in the k = 32 round its direction is descending
                        const bool up0 = k == 64 || (lane & k) == 0, up1 = k == 64 || (k < 32 && (lane & k) == 0);
                        const bool low = (lane & j) == 0;
                        k0 = ((low == up0) == (k0 < o0)) ? k0 : o0;
                        k1 = ((low == up1) == (k1 < o1)) ? k1 : o1;
                    }
                }
            l[lane] = k0;
            if ((uint32_t)lane + 32u < cnt)
                l[lane + 32] = k1;
        }
        else if (cnt <= kSortWarpKeys) {
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
        else if (lane == 0) {
            big[1u + atomicAdd(&big[0], 1u)] = this_slot;
        }
    }
    if (pend_at)
        *pend_at = pend_slot;
}

// Long lists (a va_block that many batches touch): one CTA per list, bitonic sort in shared memory up to
// kSortBigKeys runs; beyond that an odd-even transposition sort in place (slow, but such a list needs more than
// kSortBigKeys batches that all fault on the same 2 MB).
constexpr uint32_t kSortBigKeys = 8192;
constexpr int kSortBigThreads = 512;
__global__ void __launch_bounds__(kSortBigThreads)
k_service_sort_big(const uint2 *__restrict__ cw, const uint32_t *__restrict__ offset, uint32_t *__restrict__ list,
                   const uint32_t *__restrict__ big)
{
    __shared__ uint32_t s_keys[kSortBigKeys];
    const uint32_t n_big = big[0];
    for (uint32_t q = blockIdx.x; q < n_big; q += gridDim.x) {
        const uint32_t slot = big[1u + q], cnt = cw[slot].x;
        uint32_t *l = list + offset[slot];
        if (cnt <= kSortBigKeys) {
            uint32_t P = 1024;
            while (P < cnt) P <<= 1;
            for (uint32_t i = threadIdx.x; i < P; i += kSortBigThreads) s_keys[i] = i < cnt ? l[i] : 0xffffffffu;
            __syncthreads();
            for (uint32_t k = 2; k <= P; k <<= 1)
                for (uint32_t j = k >> 1; j > 0; j >>= 1) {
                    for (uint32_t i = threadIdx.x; i < P / 2; i += kSortBigThreads) {
                        const uint32_t a = 2 * i - (i & (j - 1)), b = a + j;
                        const uint32_t x = s_keys[a], y = s_keys[b];
                        const bool up = (a & k) == 0;
                        if ((x > y) == up) { s_keys[a] = y; s_keys[b] = x; }
                    }
                    __syncthreads();
                }
            for (uint32_t i = threadIdx.x; i < cnt; i += kSortBigThreads) l[i] = s_keys[i];
            __syncthreads();
        }
        else {
            for (uint32_t pass = 0; pass < cnt; pass++) {
                for (uint32_t i = (pass & 1) + 2 * threadIdx.x; i + 1 < cnt; i += 2 * kSortBigThreads) {
                    const uint32_t x = l[i], y = l[i + 1];
                    if (x > y) { l[i] = y; l[i + 1] = x; }
                }
                __threadfence_block();
                __syncthreads();
            }
        }
    }
}

} // namespace uvmtier_k
