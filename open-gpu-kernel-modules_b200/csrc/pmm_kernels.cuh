// pmm_kernels.cuh -- which HBM root chunk to evict: the reference's two eviction lists as one stamp per va_block.
// Device code only (host side: pmm.cu); also compiled by tests/simt for the CPU emulator.
//
// Reference (UVM = kernel-open/nvidia-uvm): uvm_pmm_gpu.c root_chunk_update_eviction_list :1430-1448 (every move goes
// to the tail of va_block_used / va_block_unused), pick_root_chunk_to_evict :1460-1500 (head of unused, else head of
// used), chunk_update_lists_locked :627-651 (a freshly allocated root chunk joins va_block_used); the callers are
// block_mark_memory_used / block_clear_resident_processor (uvm_va_block.c:3547-3603).
//
// stamp[slot]: 0 = no root chunk; else (UVMTIER_PMM_USED if on va_block_used) | sequence number of the last move.
// The head of a list is its smallest stamp, and every unused stamp is smaller than every used one, so
// "pick k times" = the k smallest non-zero stamps.
#pragma once

#include "service_kernels.cuh"

namespace uvmtier_k {

constexpr int kPmmThreads = 256;
constexpr int kPmmSelectThreads = 1024;

// One servicing pass: record i = (batch b, run g) at b * B + g.  A serviced run is a make_resident to the GPU that
// leaves pages there -> block_mark_memory_used -> tail of va_block_used; later runs of the same va_block win.
UVMTIER_KERNEL void __launch_bounds__(kPmmThreads)
k_pmm_note_service(uint32_t n_batches, uint32_t B, const uvmtier_fault_batch_t *__restrict__ batches,
                   const uvmtier_service_group_t *__restrict__ service, unsigned long long *__restrict__ stamp,
                   uint32_t n_slots, unsigned long long clock, unsigned long long *__restrict__ pages_in = NULL)
{
    const uint64_t total = (uint64_t)n_batches * B;
    unsigned long long moved = 0;                                         // pages the runs bring into HBM (optional tally)
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t b = (uint32_t)(i / B), g = (uint32_t)(i - (uint64_t)b * B);
        if (g >= batches[b].num_block_groups)
            continue;
        const uint32_t slot = service[i].slot;
        if (slot >= n_slots || service[i].page_fault_count == 0)
            continue;
        atomicMax(&stamp[slot], (unsigned long long)(UVMTIER_PMM_USED | (clock + i + 1)));
        moved += (unsigned long long)service[i].migrated[0] + service[i].migrated[1] + service[i].first_touch;
    }
    if (pages_in && moved)
        atomicAdd(pages_in, moved);
}

// One make_resident call (distinct va_blocks).  The state table is the one the call left behind.
UVMTIER_KERNEL void __launch_bounds__(kPmmThreads)
k_pmm_note_make_resident(uint32_t n_requests, const uvmtier_make_resident_req_t *__restrict__ reqs,
                         const uvmtier_make_resident_result_t *__restrict__ results, const uint32_t *__restrict__ gpu_resident,
                         unsigned long long *__restrict__ stamp, uint32_t n_slots, unsigned long long clock,
                         unsigned long long *__restrict__ pages_out = NULL)
{
    for (uint32_t q = blockIdx.x * blockDim.x + threadIdx.x; q < n_requests; q += gridDim.x * blockDim.x) {
        const uvmtier_make_resident_req_t rq = reqs[q];
        if (results[q].slot == 0xffffffffu || rq.slot >= n_slots)         // rejected request
            continue;
        uint32_t any = 0;
        for (int w = 0; w < 16; w++)
            any |= gpu_resident[(size_t)rq.slot * 16 + w];
        const unsigned long long seq = clock + q + 1;
        if (pages_out && results[q].migrated[2])                          // pages that left HBM (optional tally)
            atomicAdd(pages_out, (unsigned long long)results[q].migrated[2]);
        if (rq.dest_tier == UVMTIER_TIER_HBM) {
            if (any)                                                      // make_resident_finish :5076-5077
                stamp[rq.slot] = UVMTIER_PMM_USED | seq;
        }
        else if (!any) {
            if (rq.flags & UVMTIER_MR_EVICT_GPU_PAGES)
                stamp[rq.slot] = 0;                                       // evicted: the chunk goes back to the allocator
            else if (stamp[rq.slot] != 0 && results[q].migrated[2] != 0)
                stamp[rq.slot] = seq;                                     // block_clear_resident_processor :3579-3603
        }
    }
}

// needed = va_blocks with runs in these batches and no root chunk (counted once each: `seen` is a zeroed bitmap of
// n_slots bits), in_use = allocated root chunks.
UVMTIER_KERNEL void __launch_bounds__(kPmmThreads)
k_pmm_count_needed(service_kparams prm, const uvmtier_fault_group_t *__restrict__ groups,
                   const uvmtier_fault_batch_t *__restrict__ batches, const uvmtier_instance_ptr_entry_t *__restrict__ tbl,
                   const unsigned long long *__restrict__ stamp, uint32_t *__restrict__ seen, uint32_t *__restrict__ out)
{
    const uint64_t total = (uint64_t)prm.n_batches * prm.B, nth = (uint64_t)gridDim.x * blockDim.x;
    const uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t needed = 0, in_use = 0;
    for (uint64_t i = tid; i < total; i += nth) {
        const uint32_t b = (uint32_t)(i / prm.B), g = (uint32_t)(i - (uint64_t)b * prm.B);
        if (g >= batches[b].num_block_groups)
            continue;
        const uvmtier_fault_group_t grp = groups[i];
        if (grp.flags & 1u)
            continue;
        const uint32_t slot = svc_slot(prm, tbl[grp.ctx].va_space, grp.block_start);
        if (slot == 0xffffffffu || stamp[slot] != 0)
            continue;
        const uint32_t bit = 1u << (slot & 31);
        if (!(atomicOr(&seen[slot >> 5], bit) & bit))
            needed++;
    }
    for (uint64_t s = tid; s < prm.n_slots; s += nth)
        in_use += stamp[s] != 0;
    if (needed)
        atomicAdd(&out[0], needed);
    if (in_use)
        atomicAdd(&out[1], in_use);
}

// Pages resident on the GPU, on CPU node 0 (DRAM) and on CPU node 1 (CXL): out[0..2] (the caller zeroes them).
UVMTIER_KERNEL void __launch_bounds__(kPmmThreads)
k_state_count_pages(uint32_t n_slots, const uint32_t *__restrict__ gpu_resident, const uint32_t *__restrict__ cpu0,
                    const uint32_t *__restrict__ cpu1, uint32_t *__restrict__ out)
{
    const uint64_t words = (uint64_t)n_slots * 16;
    uint32_t g = 0, a = 0, b = 0;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < words; i += (uint64_t)gridDim.x * blockDim.x) {
        g += (uint32_t)__popc(gpu_resident[i]);
        a += (uint32_t)__popc(cpu0[i]);
        b += (uint32_t)__popc(cpu1[i]);
    }
    for (int d = 16; d > 0; d >>= 1) {
        g += __shfl_xor_sync(0xffffffffu, g, d);
        a += __shfl_xor_sync(0xffffffffu, a, d);
        b += __shfl_xor_sync(0xffffffffu, b, d);
    }
    if ((threadIdx.x & 31) == 0) {
        if (g) atomicAdd(&out[0], g);
        if (a) atomicAdd(&out[1], a);
        if (b) atomicAdd(&out[2], b);
    }
}

// The k-th smallest non-zero stamp by most-significant-digit radix selection (8 passes of 8 bits), one CTA.
// sel[0..1] = that stamp (0 when k = 0 or nothing is allocated), sel[2] = how many stamps are <= it.
UVMTIER_KERNEL void __launch_bounds__(kPmmSelectThreads)
k_pmm_select(const unsigned long long *__restrict__ stamp, uint32_t n_slots, uint32_t k, uint32_t *__restrict__ sel)
{
    __shared__ uint32_t hist[256];
    __shared__ unsigned long long s_prefix;
    __shared__ uint32_t s_left, s_total;
    if (threadIdx.x == 0) { s_prefix = 0; s_total = 0; }
    __syncthreads();
    uint32_t mine = 0;
    for (uint32_t i = threadIdx.x; i < n_slots; i += blockDim.x)
        mine += stamp[i] != 0;
    if (mine)
        atomicAdd(&s_total, mine);
    __syncthreads();
    const uint32_t want = k < s_total ? k : s_total;
    if (want == 0) {
        if (threadIdx.x == 0) { sel[0] = 0; sel[1] = 0; sel[2] = 0; }
        return;
    }
    if (threadIdx.x == 0)
        s_left = want;
    for (int pass = 7; pass >= 0; pass--) {
        for (int d = threadIdx.x; d < 256; d += blockDim.x)
            hist[d] = 0;
        __syncthreads();
        const unsigned long long prefix = s_prefix;
        for (uint32_t i = threadIdx.x; i < n_slots; i += blockDim.x) {
            const unsigned long long s = stamp[i];
            const bool match = pass == 7 ? true : (s >> (8 * (pass + 1))) == prefix;
            if (s != 0 && match)
                atomicAdd(&hist[(uint32_t)(s >> (8 * pass)) & 255u], 1u);
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            uint32_t left = s_left, d = 0;
            while (hist[d] < left) {                                      // the digit in which the left-th smallest falls
                left -= hist[d];
                d++;
            }
            s_left = left;
            s_prefix = (prefix << 8) | d;
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        sel[0] = (uint32_t)s_prefix;
        sel[1] = (uint32_t)(s_prefix >> 32);
        sel[2] = want;
    }
}

// One eviction request per va_block whose stamp is at most the selected one.
UVMTIER_KERNEL void __launch_bounds__(kPmmThreads)
k_pmm_emit_evictions(const unsigned long long *__restrict__ stamp, uint32_t n_slots, const uint32_t *__restrict__ sel,
                     uint32_t dest_tier, uvmtier_make_resident_req_t *__restrict__ requests, uint32_t *__restrict__ count)
{
    const unsigned long long limit = (unsigned long long)sel[0] | ((unsigned long long)sel[1] << 32);
    for (uint32_t s = blockIdx.x * blockDim.x + threadIdx.x; s < n_slots; s += gridDim.x * blockDim.x) {
        const unsigned long long v = stamp[s];
        if (v == 0 || v > limit)
            continue;
        const uint32_t at = atomicAdd(count, 1u);
        uvmtier_make_resident_req_t rq;
        rq.slot = s;
        rq.region_first = 0;
        rq.region_outer = UVMTIER_PAGES_PER_VA_BLOCK;
        rq.dest_tier = (uint8_t)dest_tier;
        rq.cause = UVMTIER_CAUSE_EVICTION;
        rq.flags = UVMTIER_MR_EVICT_GPU_PAGES;
        rq.mask_row = 0;
        requests[at] = rq;
    }
}

} // namespace uvmtier_k
