// migrate_kernels.cuh -- K3: SM-driven page gather/scatter copies between tiers (sm_100a).
// Replaces the copy primitive of block_copy_pages / ce_hal->memcopy (UVM/uvm_va_block.c:3863-3991,
// UVM/uvm_volta_ce.c:119-183) and the CPU kmap+memcpy page loop (:3972-3984).
//
// Work unit = one 4 KB piece (a 4 KB page, or 1/16 of a 64 KB page, or 1/512 of a 2 MB page).  A persistent
// one-warp CTA owns a ring of kStages shared-memory stages of kPiecesPerStage pieces (several CTAs per SM).
// ONE elected thread per CTA drives its pipeline with the TMA engine:
//   global(src tier) --cp.async.bulk--> shared  (completion: mbarrier complete_tx)
//   shared --cp.async.bulk--> global(dst tier)   (completion: bulk async-group)
// so loads of later stages are in flight while earlier stages drain; no thread touches the payload.  The src /
// dst pointers may be HBM, host-mapped (DRAM / CXL, zero-copy over PCIe) or peer HBM over NVLink: the access
// path is picked by the address, the kernel is the same.
#pragma once

#include <stdint.h>

#include "uvmtier.h"

namespace uvmtier_k {

constexpr uint32_t kPiece = 4096;            // bytes per TMA bulk op
constexpr uint32_t kPiecesPerStage = 4;      // 16 KB per stage
constexpr uint32_t kStages = 4;              // 64 KB of shared memory per CTA
constexpr uint32_t kCtasPerSm = 3;           // three independent pipelines (issuing threads) per SM
constexpr uint32_t kStageBytes = kPiece * kPiecesPerStage;
constexpr uint32_t kCopyThreads = 32;

struct copy_list_params {
    const uint8_t *src_base;
    uint8_t *dst_base;
    const uint64_t *src_off;     // per page; NULL = page i at i * page size (one contiguous range)
    const uint64_t *dst_off;
    uint64_t n_pages;
    uint32_t pieces_per_page;    // page_size / 4096
    // device-resident overrides (NULL = use the values above): the page count, and the index at which this copy's
    // segment starts in src_off / dst_off -- lets a caller enqueue copies whose size is decided by an earlier kernel
    const uint32_t *n_pages_dev;
    const uint32_t *list_base_dev;
};

__device__ __forceinline__ void copy_params_resolve(copy_list_params &p)
{
    if (p.n_pages_dev)
        p.n_pages = *p.n_pages_dev;
    if (p.list_base_dev) {
        const uint32_t b = *p.list_base_dev;
        if (p.src_off) p.src_off += b;
        if (p.dst_off) p.dst_off += b;
    }
}

// ---- thin PTX wrappers ----
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE;\n"
        "bra WAIT_LOOP;\n"
        "DONE:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_1d(void *smem_dst, const void *gmem_src, uint32_t bytes, uint64_t *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void tma_store_1d(void *gmem_dst, const void *smem_src, uint32_t bytes)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst), "r"(smem_u32(smem_src)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void tma_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void tma_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
template <int N> __device__ __forceinline__ void tma_wait_all() { asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory"); }

// Source / destination address of 4 KB piece u of the page list.
__device__ __forceinline__ void piece_addr(const copy_list_params &p, uint64_t u, const uint8_t **src, uint8_t **dst)
{
    const uint64_t page = u / p.pieces_per_page;
    const uint64_t within = (u - page * p.pieces_per_page) * kPiece;
    const uint64_t lin = page * p.pieces_per_page * kPiece;
    *src = p.src_base + (p.src_off ? p.src_off[page] : lin) + within;
    *dst = p.dst_base + (p.dst_off ? p.dst_off[page] : lin) + within;
}

// Persistent TMA pipeline.  kCtasPerSm CTAs of one warp per SM; CTA c takes items c, c + grid, ... (an item =
// kPiecesPerStage consecutive pieces = one shared-memory stage), so neighbouring CTAs stream neighbouring memory.
// All 32 lanes fetch page offsets (one piece each, 8 items at a time) so that the list reads cost one memory
// latency per 32 pieces; lane 0 alone talks to the TMA engine.
__global__ void __launch_bounds__(kCopyThreads, kCtasPerSm) k_copy_pages_tma(copy_list_params p)
{
    extern __shared__ __align__(128) uint8_t smem[];
    copy_params_resolve(p);
    __shared__ __align__(8) uint64_t full[kStages];
    __shared__ uint8_t *s_dst[kStages][kPiecesPerStage];
    const int lane = threadIdx.x;
    const uint32_t ppp_shift = 31 - __clz((int)p.pieces_per_page);        // pieces_per_page is 1, 16 or 512
    const uint64_t n_pieces = p.n_pages << ppp_shift;
    const uint64_t n_items = (n_pieces + kPiecesPerStage - 1) / kPiecesPerStage;
    const uint64_t first = blockIdx.x, step = gridDim.x;
    if (first >= n_items)
        return;
    const uint64_t my_items = (n_items - first + step - 1) / step;
    const bool big = p.pieces_per_page >= kPiecesPerStage;               // a stage never straddles two pages

    if (lane == 0) {
        for (uint32_t s = 0; s < kStages; s++)
            mbar_init(&full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncwarp();

    const uint8_t *my_src = NULL;
    uint8_t *my_dst = NULL;
    auto fetch = [&](uint64_t k0) {                    // lanes: piece (lane & 3) of item k0 + (lane >> 2)
        const uint64_t k = k0 + (uint64_t)(lane >> 2);
        const uint64_t u = ((first + k * step) * kPiecesPerStage) + (uint64_t)(lane & 3);
        if (k < my_items && u < n_pieces) {
            const uint64_t page = u >> ppp_shift;
            const uint64_t within = (u & (p.pieces_per_page - 1)) * kPiece;
            const uint64_t lin = page << (ppp_shift + 12);                 // kPiece = 4096
            my_src = p.src_base + (p.src_off ? p.src_off[page] : lin) + within;
            my_dst = p.dst_base + (p.dst_off ? p.dst_off[page] : lin) + within;
        }
    };
    auto load = [&](uint64_t k) {                      // all lanes call; lane 0 issues
        if ((k & 7) == 0)
            fetch(k);
        const uint32_t s = (uint32_t)(k % kStages);
        const uint64_t u0 = (first + k * step) * kPiecesPerStage;
        const uint32_t cnt = (uint32_t)min((uint64_t)kPiecesPerStage, n_pieces - u0);
        const uint8_t *src[kPiecesPerStage];
        uint8_t *dst[kPiecesPerStage];
#pragma unroll
        for (uint32_t j = 0; j < kPiecesPerStage; j++) {
            src[j] = (const uint8_t *)__shfl_sync(0xffffffffu, (unsigned long long)my_src, (int)((k & 7) * 4 + j));
            dst[j] = (uint8_t *)__shfl_sync(0xffffffffu, (unsigned long long)my_dst, (int)((k & 7) * 4 + j));
        }
        if (lane == 0) {
            mbar_expect_tx(&full[s], cnt * kPiece);
            if (big) {
                tma_load_1d(smem + (size_t)s * kStageBytes, src[0], cnt * kPiece, &full[s]);
                s_dst[s][0] = dst[0];
            }
            else {
#pragma unroll
                for (uint32_t j = 0; j < kPiecesPerStage; j++)
                    if (j < cnt) {
                        tma_load_1d(smem + (size_t)s * kStageBytes + j * kPiece, src[j], kPiece, &full[s]);
                        s_dst[s][j] = dst[j];
                    }
            }
        }
    };
    auto store = [&](uint64_t k) {                     // lane 0 only
        const uint32_t s = (uint32_t)(k % kStages);
        const uint64_t u0 = (first + k * step) * kPiecesPerStage;
        const uint32_t cnt = (uint32_t)min((uint64_t)kPiecesPerStage, n_pieces - u0);
        mbar_wait(&full[s], (uint32_t)((k / kStages) & 1));
        if (big) {
            tma_store_1d(s_dst[s][0], smem + (size_t)s * kStageBytes, cnt * kPiece);
        }
        else {
            for (uint32_t j = 0; j < cnt; j++)
                tma_store_1d(s_dst[s][j], smem + (size_t)s * kStageBytes + j * kPiece, kPiece);
        }
        tma_commit();
    };

    const uint64_t prologue = min(my_items, (uint64_t)(kStages - 1));
    for (uint64_t k = 0; k < prologue; k++)
        load(k);
    for (uint64_t k = 0; k < my_items; k++) {
        if (lane == 0)
            store(k);
        const uint64_t nxt = k + kStages - 1;          // reuses the stage whose stores were committed at k - 1
        if (nxt < my_items) {
            if (lane == 0)
                tma_wait_read<1>();                    // all but the newest store group have left shared memory
            load(nxt);
        }
    }
    if (lane == 0)
        tma_wait_all<0>();                             // writes performed before the kernel ends
}

// Comparison kernel: every thread moves 16-byte vectors with plain loads/stores (the pre-TMA way).
__global__ void __launch_bounds__(256) k_copy_pages_ldst(copy_list_params p)
{
    copy_params_resolve(p);
    const uint64_t n_pieces = p.n_pages * p.pieces_per_page;
    for (uint64_t u = blockIdx.x; u < n_pieces; u += gridDim.x) {
        const uint8_t *src;
        uint8_t *dst;
        piece_addr(p, u, &src, &dst);
        const uint4 v = *((const uint4 *)src + threadIdx.x);    // 256 threads x 16 B = 4 KB
        *((uint4 *)dst + threadIdx.x) = v;
    }
}

// Zero-fill of 4 KB pieces (first-touch population, UVM/uvm_va_block.c:3011).
__global__ void __launch_bounds__(256) k_zero_pages(uint8_t *base, const uint64_t *off, uint64_t n_pages, uint32_t pieces_per_page,
                                                    const uint32_t *n_pages_dev = NULL, const uint32_t *list_base_dev = NULL)
{
    if (n_pages_dev) n_pages = *n_pages_dev;
    if (list_base_dev) off += *list_base_dev;
    const uint64_t n_pieces = n_pages * pieces_per_page;
    for (uint64_t u = blockIdx.x; u < n_pieces; u += gridDim.x) {
        const uint64_t page = u / pieces_per_page;
        uint8_t *dst = base + off[page] + (u - page * pieces_per_page) * kPiece;
        *((uint4 *)dst + threadIdx.x) = make_uint4(0, 0, 0, 0);
    }
}

// Copy runs are executed per (source, destination) pair.  pair = src * 3 + dst, src 0 DRAM, 1 CXL, 2 HBM, 3..6 the HBM
// of peer GPU 0..3 of the state table, 7 zero fill; dst 0 HBM, 1 DRAM, 2 CXL (UVMTIER_RUN_SRC_* / UVMTIER_RUN_DST_*).
constexpr uint32_t kRunPairs = 24;
struct run_pair_bases { uint32_t base[kRunPairs]; };
__device__ __forceinline__ uint32_t run_pair_of(const uvmtier_copy_run_t &run)
{
    const uint32_t src = (uint32_t)run.src;
    const uint32_t dst = run.dst > 2u ? 2u : (uint32_t)run.dst;
    return src * 3u + dst;
}

// Pass 1: pages per pair (sizes the page list and places each pair's segment in it).
__global__ void __launch_bounds__(256) k_count_run_pages(const uvmtier_copy_run_t *runs, const uint32_t *n_runs_dev, uint32_t run_cap,
                                                         uint32_t *pair_count)
{
    __shared__ uint32_t s_count[kRunPairs];
    if (threadIdx.x < kRunPairs) s_count[threadIdx.x] = 0;
    __syncthreads();
    const uint32_t n_runs = min(*n_runs_dev, run_cap);
    for (uint32_t r = blockIdx.x * blockDim.x + threadIdx.x; r < n_runs; r += gridDim.x * blockDim.x) {
        const uvmtier_copy_run_t run = runs[r];
        if (run.num_pages)
            atomicAdd(&s_count[run_pair_of(run)], (uint32_t)run.num_pages);
    }
    __syncthreads();
    if (threadIdx.x < kRunPairs && s_count[threadIdx.x])
        atomicAdd(&pair_count[threadIdx.x], s_count[threadIdx.x]);
}

// Pass 2: expands the runs into one page-offset list, pair after pair: for run r, page j the offset
// slot * 2 MB + (first_page + j) * 4 KB (the same offset in every tier's arena).  One thread per run; a run's pages
// are appended to its pair's segment with one atomicAdd.
// With a chunk table (HBM arena smaller than the VA, pmm_kernels.cuh) an HBM offset is chunk_of_slot[slot] * 2 MB + ... and
// the source and destination offsets of a page differ: list_dst is then a second array (else list_dst == list).
// bases_dev != NULL: the segment starts are read from the device (k_pair_bases) instead of `bases`.
__global__ void __launch_bounds__(256) k_expand_runs(const uvmtier_copy_run_t *runs, const uint32_t *n_runs_dev, uint32_t run_cap,
                                                     run_pair_bases bases, uint32_t *pair_cursor, uint64_t *list, uint32_t list_cap,
                                                     uint64_t *list_dst = NULL, const uint32_t *chunk_of_slot = NULL,
                                                     const uint32_t *bases_dev = NULL)
{
    const uint32_t n_runs = min(*n_runs_dev, run_cap);
    for (uint32_t r = blockIdx.x * blockDim.x + threadIdx.x; r < n_runs; r += gridDim.x * blockDim.x) {
        const uvmtier_copy_run_t run = runs[r];
        if (!run.num_pages)
            continue;
        const uint32_t pair = run_pair_of(run);
        const uint32_t at = (bases_dev ? bases_dev[pair] : bases.base[pair]) + atomicAdd(&pair_cursor[pair], (uint32_t)run.num_pages);
        const uint64_t in_block = (uint64_t)run.first_page * UVMTIER_PAGE_SIZE;
        const uint64_t base = (uint64_t)run.slot * UVMTIER_VA_BLOCK_SIZE + in_block;
        uint64_t hbm = base;
        if (chunk_of_slot) {
            // a slot without a chunk cannot happen unless the caller let HBM overflow (counted where chunks are handed
            // out, and the call fails); chunk 0 keeps the copy inside the arena
            const uint32_t c = chunk_of_slot[run.slot];
            hbm = (uint64_t)(c == 0xffffffffu ? 0u : c) * UVMTIER_VA_BLOCK_SIZE + in_block;
        }
        const uint64_t src0 = run.src == UVMTIER_RUN_SRC_HBM ? hbm : base, dst0 = run.dst == UVMTIER_RUN_DST_HBM ? hbm : base;
        for (uint32_t j = 0; j < run.num_pages; j++)
            if (at + j < list_cap) {
                list[at + j] = src0 + (uint64_t)j * UVMTIER_PAGE_SIZE;
                if (list_dst)
                    list_dst[at + j] = dst0 + (uint64_t)j * UVMTIER_PAGE_SIZE;
            }
    }
}

// Segment starts of the pairs from their page counts (one thread); bases[kRunPairs] = the total.
__global__ void k_pair_bases(const uint32_t *pair_count, uint32_t *bases)
{
    if (threadIdx.x || blockIdx.x)
        return;
    uint32_t total = 0;
    for (uint32_t p = 0; p < kRunPairs; p++) {
        bases[p] = total;
        total += pair_count[p];
    }
    bases[kRunPairs] = total;
}

// One 64-bit value per page: xor-rotate-multiply over the page's 64-bit words in order.
__global__ void __launch_bounds__(256) k_checksum_pages(const uint8_t *base, const uint64_t *off, uint64_t n_pages, uint32_t page_size,
                                                        uint64_t *out)
{
    __shared__ uint64_t s_part[8];
    for (uint64_t pg = blockIdx.x; pg < n_pages; pg += gridDim.x) {
        const uint64_t *src = (const uint64_t *)(base + off[pg]);
        const uint32_t words = page_size / 8;
        uint64_t h = 0;
        for (uint32_t i = threadIdx.x; i < words; i += 256) {
            uint64_t v = src[i] + 0x9E3779B97F4A7C15ull * (uint64_t)(i + 1);
            v ^= v >> 29;
            v *= 0xBF58476D1CE4E5B9ull;
            v ^= v >> 32;
            h += v;                                    // commutative combine of position-salted words
        }
        for (int d = 16; d > 0; d >>= 1) h += __shfl_xor_sync(0xffffffffu, h, d);
        if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = h;
        __syncthreads();
        if (threadIdx.x == 0) {
            uint64_t t = 0;
            for (int w = 0; w < 8; w++) t += s_part[w];
            out[pg] = t;
        }
        __syncthreads();
    }
}

} // namespace uvmtier_k
