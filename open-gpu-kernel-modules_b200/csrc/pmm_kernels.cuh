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
constexpr uint32_t kNoChunk = 0xffffffffu;

// HBM root chunks behind the va_blocks (uvmtier_pmm_enable_chunk_table): chunk_of_slot[slot] = index of the 2 MB chunk of
// the HBM arena that backs the slot (kNoChunk = none), free_bits = one bit per chunk of the arena (set = free).  A slot
// takes a chunk when its stamp leaves 0 (block_populate_gpu_chunk, UVM/uvm_va_block.c:2974-3049) and gives it back when
// an eviction empties it (evict_root_chunk, UVM/uvm_pmm_gpu.c:1290-1380).  Which chunk a slot gets is arbitrary.
struct pmm_chunk_table {
    uint32_t *chunk_of_slot;                   // NULL = the arena is VA-sized, slot s lives at s * 2 MB
    uint32_t *free_bits;
    uint32_t n_words;                          // words of free_bits
    uint32_t *errors;                          // counts allocations that found no free chunk
};

__device__ __forceinline__ void pmm_chunk_take(const pmm_chunk_table &t, uint32_t slot)
{
    if (!t.chunk_of_slot)
        return;
    const uint32_t start = (slot * 2654435761u) % t.n_words;
    for (uint32_t i = 0; i < t.n_words; i++) {
        const uint32_t w = start + i < t.n_words ? start + i : start + i - t.n_words;
        uint32_t v = t.free_bits[w];
        while (v) {
            const uint32_t bit = v & (0u - v);
            const uint32_t old = atomicAnd(&t.free_bits[w], ~bit);
            if (old & bit) {
                t.chunk_of_slot[slot] = w * 32u + (uint32_t)(__ffs((int)bit) - 1);
                return;
            }
            v = old & ~bit;
        }
    }
    atomicAdd(t.errors, 1u);                                              // the caller let HBM overflow
}

__device__ __forceinline__ void pmm_chunk_give_back(const pmm_chunk_table &t, uint32_t slot)
{
    if (!t.chunk_of_slot)
        return;
    const uint32_t c = t.chunk_of_slot[slot];
    if (c == kNoChunk)
        return;
    t.chunk_of_slot[slot] = kNoChunk;
    atomicOr(&t.free_bits[c >> 5], 1u << (c & 31));
}

// One servicing pass: record i = (batch b, run g) at b * B + g.  A serviced run is a make_resident to the GPU that
// leaves pages there -> block_mark_memory_used -> tail of va_block_used; later runs of the same va_block win.
UVMTIER_KERNEL void __launch_bounds__(kPmmThreads)
k_pmm_note_service(uint32_t n_batches, uint32_t B, const uvmtier_fault_batch_t *__restrict__ batches,
                   const uvmtier_service_group_t *__restrict__ service, unsigned long long *__restrict__ stamp,
                   uint32_t n_slots, unsigned long long clock, unsigned long long *__restrict__ pages_in = NULL,
                   pmm_chunk_table chunks = pmm_chunk_table{NULL, NULL, 0, NULL})
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
        if (atomicMax(&stamp[slot], (unsigned long long)(UVMTIER_PMM_USED | (clock + i + 1))) == 0)
            pmm_chunk_take(chunks, slot);                                 // first GPU pages of the va_block: it needs a root chunk
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
                         unsigned long long *__restrict__ pages_out = NULL, const uint32_t *__restrict__ n_requests_dev = NULL,
                         pmm_chunk_table chunks = pmm_chunk_table{NULL, NULL, 0, NULL},
                         unsigned long long *__restrict__ dram_stamp = NULL, unsigned long long wave = 0,
                         const uint32_t *__restrict__ cpu0_resident = NULL)
{
    if (n_requests_dev)
        n_requests = min(n_requests, *n_requests_dev);
    for (uint32_t q = blockIdx.x * blockDim.x + threadIdx.x; q < n_requests; q += gridDim.x * blockDim.x) {
        const uvmtier_make_resident_req_t rq = reqs[q];
        if (results[q].slot == 0xffffffffu || rq.slot >= n_slots)         // rejected request
            continue;
        uint32_t any = 0;
        for (int w = 0; w < 16; w++)
            any |= gpu_resident[(size_t)rq.slot * 16 + w];
        const unsigned long long seq = clock + q + 1;
        if (pages_out && rq.dest_tier != UVMTIER_TIER_HBM) {               // pages the call copied to a CPU node (optional tally)
            const unsigned long long moved = (unsigned long long)results[q].migrated[0] + results[q].migrated[1] + results[q].migrated[2];
            if (moved)
                atomicAdd(pages_out, moved);
        }
        if (rq.dest_tier == UVMTIER_TIER_HBM) {
            if (any) {                                                    // make_resident_finish :5076-5077
                if (stamp[rq.slot] == 0)
                    pmm_chunk_take(chunks, rq.slot);
                stamp[rq.slot] = UVMTIER_PMM_USED | seq;
            }
        }
        else if (!any) {
            if (rq.flags & UVMTIER_MR_EVICT_GPU_PAGES) {
                stamp[rq.slot] = 0;                                       // evicted: the chunk goes back to the allocator
                pmm_chunk_give_back(chunks, rq.slot);
            }
            else if (stamp[rq.slot] != 0 && results[q].migrated[2] != 0)
                stamp[rq.slot] = seq;                                     // block_clear_resident_processor :3579-3603
        }
        if (dram_stamp && rq.dest_tier != UVMTIER_TIER_HBM) {
            // when the va_block's pages last arrived in DRAM from HBM (the demotion order); 0 once DRAM holds none of them
            uint32_t on_dram = 0;
            for (int w = 0; w < 16; w++)
                on_dram |= cpu0_resident[(size_t)rq.slot * 16 + w];
            if (!on_dram)
                dram_stamp[rq.slot] = 0;
            else if (rq.dest_tier == UVMTIER_TIER_DRAM && results[q].migrated[2] != 0)
                dram_stamp[rq.slot] = wave;
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

// ---- the tiered loop's device-side planning (uvmtier_service_fault_batches_tiered_device) ----
// plan words (device): see pmm.cu
constexpr int kPlanWords = 64;

// touched[slot] bit = the va_block has runs in these batches (bitmap zeroed by the caller).
UVMTIER_KERNEL void __launch_bounds__(kPmmThreads)
k_pmm_mark_touched(service_kparams prm, const uvmtier_fault_group_t *__restrict__ groups,
                   const uvmtier_fault_batch_t *__restrict__ batches, const uvmtier_instance_ptr_entry_t *__restrict__ tbl,
                   uint32_t *__restrict__ touched)
{
    const uint64_t total = (uint64_t)prm.n_batches * prm.B;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t b = (uint32_t)(i / prm.B), g = (uint32_t)(i - (uint64_t)b * prm.B);
        if (g >= batches[b].num_block_groups)
            continue;
        const uvmtier_fault_group_t grp = groups[i];
        if (grp.flags & 1u)
            continue;
        const uint32_t slot = svc_slot(prm, tbl[grp.ctx].va_space, grp.block_start);
        if (slot != 0xffffffffu)
            atomicOr(&touched[slot >> 5], 1u << (slot & 31));
    }
}

// Per va_block: the keys and weights of the two selections of a round, and the totals.
//   eviction: key = stamp (0 = holds no chunk), weight = 1 unless this chunk of batches faults on the block (evicting such
//             a block frees nothing: it needs a chunk again at once), aux = its GPU pages (what will arrive in DRAM)
//   demotion: key = 1 + dram_stamp for a block with DRAM pages and no chunk that this chunk of batches does not touch
//             (0 otherwise), weight = its DRAM pages
// tot[0] needed (touched, no chunk)  [1] chunks in use  [2] GPU pages  [3] DRAM pages  [4] CXL pages
UVMTIER_KERNEL void __launch_bounds__(kPmmThreads)
k_pmm_slot_keys(uint32_t n_slots, const unsigned long long *__restrict__ stamp, const unsigned long long *__restrict__ dram_stamp,
                const uint32_t *__restrict__ touched, const uint32_t *__restrict__ gpu_resident, const uint32_t *__restrict__ cpu0,
                const uint32_t *__restrict__ cpu1, uint32_t *__restrict__ e_weight, uint32_t *__restrict__ e_aux,
                unsigned long long *__restrict__ d_key, uint32_t *__restrict__ d_weight, uint32_t *__restrict__ tot)
{
    uint32_t needed = 0, in_use = 0, gp_all = 0, dp_all = 0, cp_all = 0;
    for (uint32_t s = blockIdx.x * blockDim.x + threadIdx.x; s < n_slots; s += gridDim.x * blockDim.x) {
        uint32_t gp = 0, dp = 0, cp = 0;
        const uint4 *g4 = (const uint4 *)(gpu_resident + (size_t)s * 16), *a4 = (const uint4 *)(cpu0 + (size_t)s * 16),
                    *b4 = (const uint4 *)(cpu1 + (size_t)s * 16);
        for (int q = 0; q < 4; q++) {
            const uint4 g = g4[q], a = a4[q], b = b4[q];
            gp += __popc(g.x) + __popc(g.y) + __popc(g.z) + __popc(g.w);
            dp += __popc(a.x) + __popc(a.y) + __popc(a.z) + __popc(a.w);
            cp += __popc(b.x) + __popc(b.y) + __popc(b.z) + __popc(b.w);
        }
        const bool t = (touched[s >> 5] >> (s & 31)) & 1u;
        const unsigned long long st = stamp[s];
        in_use += st != 0;
        needed += t && st == 0;
        gp_all += gp; dp_all += dp; cp_all += cp;
        e_weight[s] = (st != 0 && !t) ? 1u : 0u;
        e_aux[s] = gp;
        const bool cand = dp != 0 && st == 0 && !t;
        d_key[s] = cand ? dram_stamp[s] + 1ull : 0ull;
        d_weight[s] = dp;
    }
    for (int d = 16; d > 0; d >>= 1) {
        needed += __shfl_xor_sync(0xffffffffu, needed, d);
        in_use += __shfl_xor_sync(0xffffffffu, in_use, d);
        gp_all += __shfl_xor_sync(0xffffffffu, gp_all, d);
        dp_all += __shfl_xor_sync(0xffffffffu, dp_all, d);
        cp_all += __shfl_xor_sync(0xffffffffu, cp_all, d);
    }
    if ((threadIdx.x & 31) == 0) {
        if (needed) atomicAdd(&tot[0], needed);
        if (in_use) atomicAdd(&tot[1], in_use);
        if (gp_all) atomicAdd(&tot[2], gp_all);
        if (dp_all) atomicAdd(&tot[3], dp_all);
        if (cp_all) atomicAdd(&tot[4], cp_all);
    }
}

// Weighted selection, one CTA: the smallest threshold T with  sum of weight[i] over {key[i] != 0, key[i] <= T}  >= *target
// (most-significant-digit radix walk, 8 passes of 8 bits); every key when the weights do not add up to the target.
// sel[0..1] = T (0 when *target = 0: nothing selected), sel[2] = keys selected, sel[3] = their weights, sel[4] = their aux,
// sel[5] = 1 when the target was not reached.
UVMTIER_KERNEL void __launch_bounds__(kPmmSelectThreads)
k_pmm_select_weighted(const unsigned long long *__restrict__ key, const uint32_t *__restrict__ weight,
                      const uint32_t *__restrict__ aux, uint32_t n, const uint32_t *__restrict__ target, uint32_t *__restrict__ sel)
{
    // weights add up to at most 512 * 65536 = 2^25: 32-bit bins.  Stamps of one table share their high bytes, so most of
    // a pass's keys fall into one bin: every thread first adds up consecutive keys of the same digit, then the lanes of a
    // warp that ended on the same digit are merged, and only then the shared-memory atomics happen.
    __shared__ uint32_t hist[256];
    __shared__ unsigned long long s_prefix;
    __shared__ uint32_t s_left, s_total, s_cnt, s_w, s_aux;
    const int lane = threadIdx.x & 31;
    const uint32_t want = *target;
    if (threadIdx.x == 0) { s_prefix = 0; s_total = 0; s_cnt = 0; s_w = 0; s_aux = 0; }
    __syncthreads();
    if (want == 0) {
        if (threadIdx.x < 6) sel[threadIdx.x] = 0;
        return;
    }
    uint32_t mine = 0;
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x)
        if (key[i] != 0)
            mine += weight[i];
    mine = __reduce_add_sync(0xffffffffu, mine);
    if (lane == 0 && mine)
        atomicAdd(&s_total, mine);
    __syncthreads();
    unsigned long long T = ~0ull;
    const bool short_of = s_total < want;
    if (!short_of) {
        if (threadIdx.x == 0)
            s_left = want;
        for (int pass = 7; pass >= 0; pass--) {
            for (int d = threadIdx.x; d < 256; d += blockDim.x)
                hist[d] = 0;
            __syncthreads();
            const unsigned long long prefix = s_prefix;
            uint32_t cur = 0xffffffffu, sum = 0;                          // the digit this thread is adding up, and its sum
            for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) {
                const unsigned long long k = key[i];
                const bool match = pass == 7 ? true : (k >> (8 * (pass + 1))) == prefix;
                const uint32_t w = weight[i];
                if (k != 0 && match && w) {
                    const uint32_t d = (uint32_t)(k >> (8 * pass)) & 255u;
                    if (d != cur) {
                        if (sum) atomicAdd(&hist[cur], sum);
                        cur = d;
                        sum = 0;
                    }
                    sum += w;
                }
            }
            // lanes that ended on the same digit: the lowest of them adds for all
            uint32_t total = 0, first = 32;
            for (int src = 0; src < 32; src++) {
                const uint32_t d = __shfl_sync(0xffffffffu, cur, src), v = __shfl_sync(0xffffffffu, sum, src);
                if (d == cur && v) {
                    total += v;
                    if (first == 32) first = (uint32_t)src;
                }
            }
            if (sum && first == (uint32_t)lane)
                atomicAdd(&hist[cur], total);
            __syncthreads();
            if (threadIdx.x == 0) {
                uint32_t left = s_left, d = 0;
                while (hist[d] < left) {                                  // the digit in which the running sum reaches the target
                    left -= hist[d];
                    d++;
                }
                s_left = left;
                s_prefix = (prefix << 8) | d;
            }
            __syncthreads();
        }
        T = s_prefix;
    }
    uint32_t c = 0, w = 0, a = 0;
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) {
        const unsigned long long k = key[i];
        if (k != 0 && k <= T) {
            c++;
            w += weight[i];
            a += aux ? aux[i] : 0u;
        }
    }
    c = __reduce_add_sync(0xffffffffu, c);
    w = __reduce_add_sync(0xffffffffu, w);
    a = __reduce_add_sync(0xffffffffu, a);
    if (lane == 0 && c) { atomicAdd(&s_cnt, c); atomicAdd(&s_w, w); atomicAdd(&s_aux, a); }
    __syncthreads();
    if (threadIdx.x == 0) {
        sel[0] = (uint32_t)T; sel[1] = (uint32_t)(T >> 32);
        sel[2] = s_cnt; sel[3] = s_w; sel[4] = s_aux; sel[5] = short_of ? 1u : 0u;
    }
}

// One request per va_block whose key is non-zero and at most the selected threshold.  evict != 0: an eviction (whole
// block, cause EVICTION, UVMTIER_MR_EVICT_GPU_PAGES) to the tier *dest_tier names; else a move of the block's DRAM pages
// to the CXL node (page mask = row `slot` of the cpu_resident[0] array).
UVMTIER_KERNEL void __launch_bounds__(kPmmThreads)
k_pmm_emit_selected(const unsigned long long *__restrict__ key, uint32_t n_slots, const uint32_t *__restrict__ sel, int evict,
                    const uint32_t *__restrict__ dest_tier, uvmtier_make_resident_req_t *__restrict__ requests,
                    uint32_t *__restrict__ count)
{
    const unsigned long long limit = (unsigned long long)sel[0] | ((unsigned long long)sel[1] << 32);
    if (limit == 0)
        return;
    for (uint32_t s = blockIdx.x * blockDim.x + threadIdx.x; s < n_slots; s += gridDim.x * blockDim.x) {
        const unsigned long long v = key[s];
        if (v == 0 || v > limit)
            continue;
        const uint32_t at = atomicAdd(count, 1u);
        uvmtier_make_resident_req_t rq;
        rq.slot = s;
        rq.region_first = 0;
        rq.region_outer = UVMTIER_PAGES_PER_VA_BLOCK;
        rq.dest_tier = (uint8_t)(evict ? *dest_tier : UVMTIER_TIER_CXL);
        rq.cause = (uint8_t)(evict ? UVMTIER_CAUSE_EVICTION : UVMTIER_CAUSE_API_MIGRATE);
        rq.flags = (uint16_t)(evict ? UVMTIER_MR_EVICT_GPU_PAGES : UVMTIER_MR_HAS_PAGE_MASK);
        rq.mask_row = evict ? 0u : s;
        requests[at] = rq;
    }
}

// The three decisions of a round, one thread each time (step 0: how many chunks must go; 1: how many DRAM pages must be
// demoted to make room for what the victims bring; 2: where the victims go, and the tallies).
// plan[0..4] = tot of k_pmm_slot_keys, [8] eviction target, [9] demotion target, [10] destination tier of the evictions,
// [16..21] eviction selection, [24..29] demotion selection, [32..] tallies: waves, to DRAM, to CXL, blocks demoted,
// overcommitted passes, peak chunks in use, chunks in use after the pass.
UVMTIER_KERNEL void k_pmm_plan(uint32_t *__restrict__ plan, int step, uint32_t capacity_chunks, unsigned long long dram_capacity_pages,
                               int demote)
{
    if (threadIdx.x != 0 || blockIdx.x != 0)
        return;
    const uint32_t needed = plan[0], in_use = plan[1], dram_pages = plan[3];
    if (step == 0) {
        const unsigned long long want = (unsigned long long)in_use + needed;
        plan[8] = want > capacity_chunks ? (uint32_t)(want - capacity_chunks) : 0u;
    }
    else if (step == 1) {
        const unsigned long long arriving = plan[16 + 4], total = (unsigned long long)dram_pages + arriving;
        plan[9] = (demote && plan[16 + 2] && total > dram_capacity_pages) ? (uint32_t)(total - dram_capacity_pages) : 0u;
    }
    else {
        const uint32_t k = plan[16 + 2], untouched = plan[16 + 3], m = plan[24 + 2];
        const unsigned long long after = (unsigned long long)dram_pages - plan[24 + 3] + plan[16 + 4];
        const uint32_t dest = after <= dram_capacity_pages ? UVMTIER_TIER_DRAM : UVMTIER_TIER_CXL;
        plan[10] = dest;
        if (k) {
            plan[32] += 1;
            plan[dest == UVMTIER_TIER_DRAM ? 33 : 34] += k;
        }
        plan[35] += m;
        // victims this chunk of batches faults on come straight back
        const uint32_t after_chunks = in_use - k + needed + (k - untouched);
        if (after_chunks > capacity_chunks)
            plan[36] += 1;
        if (after_chunks > plan[37])
            plan[37] = after_chunks;
        plan[38] = after_chunks;
    }
}

} // namespace uvmtier_k
