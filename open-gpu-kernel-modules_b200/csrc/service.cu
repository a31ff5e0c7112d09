// service.cu -- C-ABI entry points of the per-va_block servicing pass (stage C/D/E) and the block state table.
// Reference: UVM/uvm_gpu_replayable_faults.c:2232-2377 (service_fault_batch) and callees; see service_kernels.cuh.
//
// Working space in HBM (one buffer owned by the context, grown on demand and reused; or the fixed array of the
// no-alloc entry point): per slot the packed (runs, entries) counter and the start of its run list (12 B x
// n_slots), per run its (slot, position) pair and its place in a list (12 B per run), one flag per batch and two
// short work lists.
// Algorithmic bytes per serviced va_block run: 16 (run) + 16 per entry + 2 per entry flag + 96 record
// + the block's state (6 masks x 64 B + 8 B) read and written once per touched block, + 8 per copy run.
#include "common.cuh"

#include <stdlib.h>

#include "service_kernels.cuh"
#include "service_lists.cuh"
#include "service_grp.cuh"

using namespace uvmtier_k;

// A slot whose runs hold this many entries or more is serviced by a whole warp (entries filtered 32 at a time)
// instead of a small group of lanes (entries one after the other).
constexpr uint32_t kHeavyMinEntries = 96;

// Carves the working space for one call out of `base`.
struct service_scratch_view {
    uint32_t *cw, *cursor, *hist, *big, *heavy, *batch_unmapped;   // zeroed at the start of every call
    uint32_t *offset, *light, *light_rank, *slotpos, *list;                           // fully rewritten before use
    size_t zero_words, words;
};
static service_scratch_view scratch_view(uint32_t *base, uint32_t n_slots, uint32_t n_batches, size_t total)
{
    service_scratch_view v;
    const size_t slots4 = ((size_t)n_slots + 3) & ~(size_t)3;            // keeps every array 16-byte aligned
    const size_t total4 = (total + 3) & ~(size_t)3;
    const size_t big_cap = ((total / kSortWarpKeys + 2) + 3) & ~(size_t)3;
    const size_t heavy_cap = ((total / kHeavyMinEntries + 2) + 3) & ~(size_t)3;
    const size_t batches4 = ((size_t)n_batches + 3) & ~(size_t)3;
    uint32_t *p = base;
    v.cw = p; p += 2 * slots4;
    v.cursor = p; p += 4;
    v.hist = p; p += kLightBins;
    v.big = p; p += big_cap;
    v.heavy = p; p += heavy_cap;
    v.batch_unmapped = p; p += batches4;
    v.zero_words = (size_t)(p - base);
    v.offset = p; p += slots4;
    v.light = p; p += slots4;
    v.light_rank = p; p += slots4;
    v.slotpos = p; p += 2 * total4;
    v.list = p; p += total4;
    v.words = (size_t)(p - base);
    return v;
}

static uint32_t service_lanes_per_block()
{
    static int lanes = -1;
    if (lanes < 0) {
        const char *e = getenv("UVMTIER_SERVICE_LANES");
        const int v = e ? atoi(e) : 4;
        lanes = (v == 1 || v == 2 || v == 4 || v == 8 || v == 16) ? v : 4;
    }
    return (uint32_t)lanes;
}

template <int G, int M>
static NV_STATUS launch_grp(uvmtier_ctx_t *ctx, const service_kparams &prm, const uvmtier_ordered_fault_t *d_ordered,
                            const uvmtier_fault_group_t *d_groups, uvmtier_fault_batch_t *d_batches, const block_state_ptrs &bs,
                            const service_scratch_view &v, uint32_t heavy_min, uint16_t *d_entry_flags,
                            uvmtier_service_group_t *d_service, uvmtier_copy_run_t *d_runs, uint32_t *d_counters)
{
    const size_t threads = (size_t)prm.n_slots * G;
    const uint32_t grid = (uint32_t)((threads + kGrpThreads - 1) / kGrpThreads);
    UVMTIER_LAUNCH((k_service_blocks_grp<G, M>), grid, kGrpThreads, 0, ctx->stream, prm, d_ordered, d_groups, d_batches, bs,
                   (const uint2 *)v.cw, v.offset, v.list, v.light, v.cursor + 1, d_entry_flags, d_service, d_runs, d_counters);
    return NV_OK;
}

// The whole pass on the context's stream, working space at `base` (v.words words).
static NV_STATUS service_pass(uvmtier_ctx_t *ctx, uint32_t *base, uint32_t n_batches, const uvmtier_ordered_fault_t *d_ordered,
                              const uvmtier_fault_group_t *d_groups, uvmtier_fault_batch_t *d_batches, const uvmtier_va_layout_t *lay,
                              const uvmtier_block_state_t *state, uint16_t *d_entry_flags, uvmtier_service_group_t *d_service,
                              uvmtier_copy_run_t *d_runs, uint32_t run_cap, uint32_t *d_counters)
{
    const uint32_t B = ctx->faults->gp.max_batch_size;
    const uint32_t n_slots = lay->n_spaces * lay->blocks_per_space;
    const size_t total = (size_t)n_batches * B;
    const service_scratch_view v = scratch_view(base, n_slots, n_batches, total);

    service_kparams prm;
    prm.va_base = lay->va_base;
    prm.B = B;
    prm.n_batches = n_batches;
    prm.blocks_per_space = lay->blocks_per_space;
    prm.n_spaces = lay->n_spaces;
    prm.gpu_id = lay->gpu_id;
    prm.n_slots = n_slots;
    prm.pf_enable = ctx->cfg.uvm_perf_prefetch_enable;
    prm.pf_threshold = ctx->cfg.uvm_perf_prefetch_threshold;
    prm.pf_min_faults = ctx->cfg.uvm_perf_prefetch_min_faults;
    prm.run_cap = run_cap;
    prm.B_shift = svc_log2_or_none(B);

    block_state_ptrs bs;
    bs.gpu_resident = (uint32_t *)state->gpu_resident;
    bs.cpu_resident[0] = (uint32_t *)state->cpu_resident[0];
    bs.cpu_resident[1] = (uint32_t *)state->cpu_resident[1];
    bs.gpu_pte[0] = (uint32_t *)state->gpu_pte[0];
    bs.gpu_pte[1] = (uint32_t *)state->gpu_pte[1];
    bs.gpu_pte[2] = (uint32_t *)state->gpu_pte[2];
    bs.prefetch_info = state->prefetch_info;
    bs.logical_prot = state->logical_prot;
    bs.preferred_location = state->preferred_location;
    bs.gpu_evicted = (uint32_t *)state->gpu_evicted;
    bs.read_duplicated = (uint32_t *)state->read_duplicated;

    // The working space still holds the previous call's numbers.  The packed per-slot counters, the list cursor,
    // the two work lists' lengths and the per-batch flags are accumulated into and must start at zero (one memset:
    // they are adjacent); offsets, (slot, position) pairs and run lists are fully rewritten before they are read.
    UVMTIER_CUDA(cudaMemsetAsync(base, 0, v.zero_words * sizeof(uint32_t), ctx->stream));
    UVMTIER_CUDA(cudaMemsetAsync(d_counters, 0, 2 * sizeof(uint32_t), ctx->stream));
    UVMTIER_CUDA(cudaMemsetAsync(d_entry_flags, 0, total * sizeof(uint16_t), ctx->stream));

    const uvmtier_instance_ptr_entry_t *d_tbl = ctx->faults->d_table;
    const uint2 *cw = (const uint2 *)v.cw;
    const size_t cap = (size_t)ctx->sm_count * 8;
    size_t grid = (total + kSvcThreads - 1) / kSvcThreads;
    if (grid > cap)
        grid = cap;
    UVMTIER_LAUNCH(k_service_count, (int)grid, kSvcThreads, 0, ctx->stream, prm, d_groups, d_batches, d_tbl,
                   (unsigned long long *)v.cw, v.batch_unmapped, (uint2 *)v.slotpos);
    // UVMTIER_SERVICE_HEAVY_MIN moves the group / warp split up (for experiments; 0xffffffff = no block on a warp).
    // It cannot go below kHeavyMinEntries: that is what bounds the length of the heavy list.
    static const char *heavy_env = getenv("UVMTIER_SERVICE_HEAVY_MIN");
    const uint32_t heavy_min = heavy_env ? (uint32_t)strtoul(heavy_env, NULL, 0) : kHeavyMinEntries;
    const uint32_t list_heavy_min = heavy_min < kHeavyMinEntries ? kHeavyMinEntries : heavy_min;
    UVMTIER_LAUNCH(k_service_alloc, (n_slots + 255) / 256, 256, 0, ctx->stream, n_slots, cw, list_heavy_min, v.offset, v.cursor, v.hist,
                   v.light_rank);
    UVMTIER_LAUNCH(k_service_fill, (int)grid, kSvcThreads, 0, ctx->stream, (uint32_t)total, (const uint2 *)v.slotpos, v.offset, v.list);
    size_t sgrid = ((size_t)n_slots + (kSortThreads / 32) - 1) / (kSortThreads / 32);
    if (sgrid > cap)
        sgrid = cap;
    UVMTIER_LAUNCH(k_service_sort_lists, (int)sgrid, kSortThreads, 0, ctx->stream, n_slots, cw, v.offset, v.list, list_heavy_min,
                   v.big, v.heavy, v.hist, v.light_rank, v.light);
    UVMTIER_LAUNCH(k_service_sort_big, ctx->sm_count, kSortBigThreads, 0, ctx->stream, cw, v.offset, v.list, v.big);

    // lanes per va_block x CTAs per SM the kernel is compiled for; UVMTIER_SERVICE_LANES / UVMTIER_SERVICE_MINCTAS
    // pick another instantiation (for experiments)
    static const char *minctas_env = getenv("UVMTIER_SERVICE_MINCTAS");
    const int minctas = minctas_env ? atoi(minctas_env) : 0;
#define GRP_ARGS ctx, prm, d_ordered, d_groups, d_batches, bs, v, list_heavy_min, d_entry_flags, d_service, d_runs, d_counters
    NV_STATUS st = NV_OK;
    switch (service_lanes_per_block()) {
        case 1: st = launch_grp<1, 1>(GRP_ARGS); break;
        case 2: st = minctas == 4 ? launch_grp<2, 4>(GRP_ARGS) : (minctas == 3 ? launch_grp<2, 3>(GRP_ARGS) : launch_grp<2, 2>(GRP_ARGS)); break;
        case 8: st = minctas == 8 ? launch_grp<8, 8>(GRP_ARGS) : (minctas == 6 ? launch_grp<8, 6>(GRP_ARGS) : launch_grp<8, 5>(GRP_ARGS)); break;
        case 16: st = launch_grp<16, 8>(GRP_ARGS); break;
        default: st = minctas == 7 ? launch_grp<4, 7>(GRP_ARGS) : minctas == 6 ? launch_grp<4, 6>(GRP_ARGS) : (minctas == 5 ? launch_grp<4, 5>(GRP_ARGS) : launch_grp<4, 4>(GRP_ARGS)); break;
    }
#undef GRP_ARGS
    if (st != NV_OK)
        return st;
    UVMTIER_LAUNCH(k_service_blocks, ctx->sm_count * 2, kSvcThreads, 0, ctx->stream, prm, d_ordered, d_groups, d_batches, bs, cw,
                   v.offset, v.list, v.heavy, d_entry_flags, d_service, d_runs, d_counters);
    UVMTIER_LAUNCH(k_service_unmapped, (prm.n_batches + 127) / 128, 128, 0, ctx->stream, prm, d_ordered, d_groups, d_batches, d_tbl,
                   v.batch_unmapped, d_entry_flags, d_service);
    return NV_OK;
}

static NV_STATUS service_check_args(uvmtier_ctx_t *ctx, uint32_t n_batches, const void *d_ordered, const void *d_groups,
                                    const void *d_batches, const uvmtier_va_layout_t *lay, const uvmtier_block_state_t *state,
                                    const void *d_entry_flags, const void *d_service, const void *d_runs, uint32_t run_cap,
                                    const void *d_counters)
{
    if (!ctx || !lay || !state)
        return NV_ERR_INVALID_ARGUMENT;
    if (!ctx->faults)
        return NV_ERR_INVALID_STATE;
    if (!n_batches)
        return NV_OK;
    if (!d_ordered || !d_groups || !d_batches || !d_entry_flags || !d_service || !d_counters || (run_cap && !d_runs))
        return NV_ERR_INVALID_ARGUMENT;
    const size_t n_slots64 = (size_t)lay->n_spaces * lay->blocks_per_space;
    if (!ctx->faults->gp.max_batch_size || n_slots64 == 0 || n_slots64 > kMaxSlots ||
        (size_t)n_batches * ctx->faults->gp.max_batch_size > 0xffffffffu)
        return NV_ERR_INVALID_ARGUMENT;
    return NV_OK;
}

void uvmtier_service_scratch_destroy(uvmtier_ctx_t *ctx)
{
    uvmtier_service_scratch *s = &ctx->svc;
    if (s->d_base) cudaFree(s->d_base);
    memset(s, 0, sizeof(*s));
}

extern "C" {

NV_STATUS uvmtier_block_state_alloc(uvmtier_ctx_t *ctx, const uvmtier_va_layout_t *lay, uvmtier_block_state_t *out)
{
    if (!ctx || !lay || !out || !lay->n_spaces || !lay->blocks_per_space)
        return NV_ERR_INVALID_ARGUMENT;
    const size_t n = (size_t)lay->n_spaces * lay->blocks_per_space;
    if (n > kMaxSlots)
        return NV_ERR_INVALID_ARGUMENT;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    memset(out, 0, sizeof(*out));
    // one allocation: 6 mask arrays, prefetch_info, preferred_location, logical_prot, then (64-byte aligned) gpu_evicted
    const size_t mask_bytes = n * 64;
    const size_t evicted_at = (6 * mask_bytes + n * 4 + n * 2 + n + 63) & ~(size_t)63;
    char *base = NULL;
    UVMTIER_CUDA(cudaMalloc(&base, evicted_at + mask_bytes));
    out->gpu_resident = (uvmtier_page_mask_t *)base;
    out->cpu_resident[0] = (uvmtier_page_mask_t *)(base + mask_bytes);
    out->cpu_resident[1] = (uvmtier_page_mask_t *)(base + 2 * mask_bytes);
    out->gpu_pte[0] = (uvmtier_page_mask_t *)(base + 3 * mask_bytes);
    out->gpu_pte[1] = (uvmtier_page_mask_t *)(base + 4 * mask_bytes);
    out->gpu_pte[2] = (uvmtier_page_mask_t *)(base + 5 * mask_bytes);
    out->prefetch_info = (uvmtier_prefetch_info_t *)(base + 6 * mask_bytes);
    out->preferred_location = (uint16_t *)(base + 6 * mask_bytes + n * 4);
    out->logical_prot = (uint8_t *)(base + 6 * mask_bytes + n * 4 + n * 2);
    out->gpu_evicted = (uvmtier_page_mask_t *)(base + evicted_at);
    UVMTIER_CUDA(cudaMemsetAsync(base, 0, 6 * mask_bytes, ctx->stream));
    UVMTIER_CUDA(cudaMemsetAsync(base + evicted_at, 0, mask_bytes, ctx->stream));
    // prefetch_info = {UVMTIER_ID_INVALID (0x0101), 0}; preferred_location = 0x0101; logical_prot = 3
    UVMTIER_CUDA(cudaMemset2DAsync(out->prefetch_info, 4, 0x01, 2, n, ctx->stream));
    UVMTIER_CUDA(cudaMemset2DAsync((char *)out->prefetch_info + 2, 4, 0x00, 2, n, ctx->stream));
    UVMTIER_CUDA(cudaMemsetAsync(out->preferred_location, 0x01, n * 2, ctx->stream));
    UVMTIER_CUDA(cudaMemsetAsync(out->logical_prot, 3, n, ctx->stream));
    UVMTIER_CUDA(cudaStreamSynchronize(ctx->stream));
    return NV_OK;
}

NV_STATUS uvmtier_block_state_free(uvmtier_ctx_t *ctx, uvmtier_block_state_t *state)
{
    if (!ctx || !state)
        return NV_ERR_INVALID_ARGUMENT;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    if (state->gpu_resident)
        UVMTIER_CUDA(cudaFree(state->gpu_resident));
    if (state->read_duplicated)
        UVMTIER_CUDA(cudaFree(state->read_duplicated));
    memset(state, 0, sizeof(*state));
    return NV_OK;
}

NV_STATUS uvmtier_block_state_enable_read_duplication(uvmtier_ctx_t *ctx, const uvmtier_va_layout_t *lay,
                                                      uvmtier_block_state_t *state)
{
    if (!ctx || !lay || !state || !state->gpu_resident || !lay->n_spaces || !lay->blocks_per_space)
        return NV_ERR_INVALID_ARGUMENT;
    if (state->read_duplicated)
        return NV_OK;
    const size_t n = (size_t)lay->n_spaces * lay->blocks_per_space;
    if (n > kMaxSlots)
        return NV_ERR_INVALID_ARGUMENT;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    void *p = NULL;
    UVMTIER_CUDA(cudaMalloc(&p, n * 64));
    UVMTIER_CUDA(cudaMemsetAsync(p, 0, n * 64, ctx->stream));
    UVMTIER_CUDA(cudaStreamSynchronize(ctx->stream));
    state->read_duplicated = (uvmtier_page_mask_t *)p;
    return NV_OK;
}

NV_STATUS uvmtier_service_fault_batches_device(uvmtier_ctx_t *ctx, uint32_t n_batches,
                                               const uvmtier_ordered_fault_t *d_ordered,
                                               const uvmtier_fault_group_t *d_groups, uvmtier_fault_batch_t *d_batches,
                                               const uvmtier_va_layout_t *lay, const uvmtier_block_state_t *state,
                                               uint16_t *d_entry_flags, uvmtier_service_group_t *d_service,
                                               uvmtier_copy_run_t *d_runs, uint32_t run_cap, uint32_t *d_counters)
{
    const NV_STATUS ok = service_check_args(ctx, n_batches, d_ordered, d_groups, d_batches, lay, state, d_entry_flags, d_service,
                                            d_runs, run_cap, d_counters);
    if (ok != NV_OK || !n_batches)
        return ok;
    const uint32_t n_slots = lay->n_spaces * lay->blocks_per_space;
    const size_t total = (size_t)n_batches * ctx->faults->gp.max_batch_size;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    uvmtier_service_scratch *s = &ctx->svc;
    const size_t words = scratch_view(NULL, n_slots, n_batches, total).words;
    if (words > s->cap_words) {
        if (s->d_base) {
            UVMTIER_CUDA(cudaStreamSynchronize(ctx->stream));
            UVMTIER_CUDA(cudaFree(s->d_base));
            s->d_base = NULL;
            s->cap_words = 0;
        }
        UVMTIER_CUDA(cudaMalloc(&s->d_base, words * sizeof(uint32_t)));
        s->cap_words = words;
    }
    return service_pass(ctx, s->d_base, n_batches, d_ordered, d_groups, d_batches, lay, state, d_entry_flags, d_service, d_runs,
                        run_cap, d_counters);
}

} // extern "C"

// ------------------------------------------------------------------------------------------------
// Servicing pass for callers that must not allocate.
//
// Who: a fault handler is the one piece of code that still has to run when the GPU's memory is exhausted -- it is
// what frees memory, by evicting.  A servicing thread in that position cannot ask cudaMalloc for working space:
// the request may fail, or block behind the very allocator state the faults are waiting on.  The same holds for
// a host integration that pins its memory budget up front and forbids allocator calls afterwards.
//
// What: the pass above keeps its working space in a buffer it grows with cudaMalloc.  This variant keeps the same
// working space in one fixed array that is part of the library image.  Every GPU that loads the library carries
// the array (about 4.5 MB) until the process exits; the price of never allocating is a hard ceiling on the
// problem size.
constexpr uint32_t kNoAllocMaxSlots = 65536;        // va_block slots: 128 GB of VA
constexpr uint32_t kNoAllocMaxEntries = 262144;     // n_batches * max_batch_size
constexpr uint32_t kNoAllocMaxBatches = 4096;
// scratch_view() of the largest problem, with room for its alignment padding
constexpr size_t kNoAllocWords = (size_t)kNoAllocMaxSlots * 5 + kLightBins + (kNoAllocMaxEntries / kSortWarpKeys + 8) +
                                 (kNoAllocMaxEntries / kHeavyMinEntries + 8) + kNoAllocMaxBatches +
                                 (size_t)kNoAllocMaxEntries * 3 + 64;

static __device__ __align__(256) uint32_t g_service_scratch[kNoAllocWords];

// In host code `g_service_scratch` is only a handle for the symbol, not memory a kernel could use.  The device
// address is looked up once per GPU, with that GPU current, and kept here.
static uint32_t *g_service_scratch_addr[64];

extern "C" NV_STATUS uvmtier_service_fault_batches_noalloc_device(uvmtier_ctx_t *ctx, uint32_t n_batches,
                                                                  const uvmtier_ordered_fault_t *d_ordered,
                                                                  const uvmtier_fault_group_t *d_groups,
                                                                  uvmtier_fault_batch_t *d_batches,
                                                                  const uvmtier_va_layout_t *lay,
                                                                  const uvmtier_block_state_t *state,
                                                                  uint16_t *d_entry_flags, uvmtier_service_group_t *d_service,
                                                                  uvmtier_copy_run_t *d_runs, uint32_t run_cap,
                                                                  uint32_t *d_counters)
{
    if (ctx && lay && ctx->faults) {
        // the size ceiling is checked before anything else so that an oversized request is reported as such
        const size_t n_slots64 = (size_t)lay->n_spaces * lay->blocks_per_space;
        const size_t total = (size_t)n_batches * ctx->faults->gp.max_batch_size;
        if (n_slots64 > kNoAllocMaxSlots || total > kNoAllocMaxEntries || n_batches > kNoAllocMaxBatches) {
            char msg[256];
            snprintf(msg, sizeof(msg),
                     "uvmtier_service_fault_batches_noalloc_device: %zu va_block slots / %zu entries / %u batches exceed the "
                     "fixed scratch (%u slots, %u entries, %u batches); use uvmtier_service_fault_batches_device",
                     n_slots64, total, n_batches, kNoAllocMaxSlots, kNoAllocMaxEntries, kNoAllocMaxBatches);
            uvmtier_set_error(msg, cudaErrorInvalidValue, __FILE__, __LINE__);
            return NV_ERR_INVALID_ARGUMENT;
        }
    }
    const NV_STATUS ok = service_check_args(ctx, n_batches, d_ordered, d_groups, d_batches, lay, state, d_entry_flags, d_service,
                                            d_runs, run_cap, d_counters);
    if (ok != NV_OK || !n_batches)
        return ok;
    if (ctx->device < 0 || ctx->device >= 64)
        return NV_ERR_INVALID_ARGUMENT;
    const uint32_t n_slots = lay->n_spaces * lay->blocks_per_space;
    const size_t total = (size_t)n_batches * ctx->faults->gp.max_batch_size;
    if (scratch_view(NULL, n_slots, n_batches, total).words > kNoAllocWords)
        return NV_ERR_INVALID_ARGUMENT;                                   // cannot happen within the three limits above
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    if (!g_service_scratch_addr[ctx->device]) {
        uint32_t *p = NULL;
        UVMTIER_CUDA(cudaGetSymbolAddress((void **)&p, g_service_scratch));
        g_service_scratch_addr[ctx->device] = p;
    }
    return service_pass(ctx, g_service_scratch_addr[ctx->device], n_batches, d_ordered, d_groups, d_batches, lay, state,
                        d_entry_flags, d_service, d_runs, run_cap, d_counters);
}
