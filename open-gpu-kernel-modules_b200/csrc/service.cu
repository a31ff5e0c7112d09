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
#define UVMTIER_AC_SERVICE
#include "access_counter_kernels.cuh"

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

// The whole pass on the context's stream, working space at `base` (v.words words).
static NV_STATUS service_pass(uvmtier_ctx_t *ctx, uint32_t *base, uint32_t n_batches, const uvmtier_ordered_fault_t *d_ordered,
                              const uvmtier_fault_group_t *d_groups, uvmtier_fault_batch_t *d_batches, const uvmtier_va_layout_t *lay,
                              const uvmtier_block_state_t *state, uint16_t *d_entry_flags, uvmtier_service_group_t *d_service,
                              uvmtier_copy_run_t *d_runs, uint32_t run_cap, uint32_t *d_counters, uint32_t batch_base = 0,
                              bool zero_outputs = true, bool non_replayable = false, const svc_thrash_params *thrash = NULL)
{
    svc_thrash_params tp;
    memset(&tp, 0, sizeof(tp));
    if (thrash)
        tp = *thrash;
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
    prm.batch_base = batch_base;
    prm.non_replayable = non_replayable ? 1u : 0u;

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
    if (zero_outputs) {                                     // a chunked caller has done both for the whole trace
        UVMTIER_CUDA(cudaMemsetAsync(d_counters, 0, 2 * sizeof(uint32_t), ctx->stream));
        UVMTIER_CUDA(cudaMemsetAsync(d_entry_flags, 0, total * sizeof(uint16_t), ctx->stream));
    }

    const uvmtier_instance_ptr_entry_t *d_tbl = ctx->faults->d_table;
    const uint2 *cw = (const uint2 *)v.cw;
    const size_t cap = (size_t)ctx->sm_count * 8;
    size_t grid = (total + kSvcThreads - 1) / kSvcThreads;
    if (grid > cap)
        grid = cap;
    UVMTIER_LAUNCH(k_service_count, (int)grid, kSvcThreads, 0, ctx->stream, prm, d_groups, d_batches, d_tbl,
                   (unsigned long long *)v.cw, v.batch_unmapped, (uint2 *)v.slotpos);
    const uint32_t list_heavy_min = kHeavyMinEntries;
    UVMTIER_LAUNCH(k_service_alloc, (n_slots + 255) / 256, 256, 0, ctx->stream, n_slots, cw, list_heavy_min, v.offset, v.cursor, v.hist,
                   v.light_rank, (const uvmtier_thrashing_block_t *)tp.th.blocks);
    UVMTIER_LAUNCH(k_service_fill, (int)grid, kSvcThreads, 0, ctx->stream, (uint32_t)total, (const uint2 *)v.slotpos, v.offset, v.list);
    size_t sgrid = ((size_t)n_slots + (kSortThreads / 32) - 1) / (kSortThreads / 32);
    if (sgrid > cap)
        sgrid = cap;
    UVMTIER_LAUNCH(k_service_sort_lists, (int)sgrid, kSortThreads, 0, ctx->stream, n_slots, cw, v.offset, v.list, list_heavy_min,
                   v.big, v.heavy, v.hist, v.light_rank, v.light, (const uvmtier_thrashing_block_t *)tp.th.blocks);
    // Three independent branches from here: the light va_blocks (lane groups), the heavy ones (their long lists
    // are sorted first, then a warp each) and the runs outside every va_range.  They touch different va_blocks,
    // records and entry flags, and add to the batch counters atomically; the two small branches run on side streams
    // (created with the context) next to the big one.
    cudaStream_t s_heavy = ctx->aux_stream[0], s_unmapped = ctx->aux_stream[1];
    UVMTIER_CUDA(cudaEventRecord(ctx->aux_events[0], ctx->stream));
    UVMTIER_CUDA(cudaStreamWaitEvent(s_heavy, ctx->aux_events[0], 0));
    UVMTIER_CUDA(cudaStreamWaitEvent(s_unmapped, ctx->aux_events[0], 0));
    // a launch that fails between the fork and the join must not leave the side streams unjoined (a capturing stream would
    // be invalidated): remember the first failure, join, then report it
    NV_STATUS st = NV_OK;
#define SVC_SIDE_LAUNCH(kernel, grid, block, stream, ...)                                  \
    do {                                                                                   \
        kernel<<<(grid), (block), 0, (stream)>>>(__VA_ARGS__);                             \
        g_uvmtier_launches.fetch_add(1, std::memory_order_relaxed);                        \
        const cudaError_t _e = cudaGetLastError();                                         \
        if (_e != cudaSuccess && st == NV_OK) {                                            \
            uvmtier_set_error(#kernel, _e, __FILE__, __LINE__);                            \
            st = uvmtier_status_from_cuda(_e);                                             \
        }                                                                                  \
    } while (0)
    SVC_SIDE_LAUNCH(k_service_sort_big, ctx->sm_count, kSortBigThreads, s_heavy, cw, v.offset, v.list, v.big);
    SVC_SIDE_LAUNCH(k_service_blocks, ctx->sm_count * 2, kSvcThreads, s_heavy, prm, d_ordered, d_groups, d_batches, bs, cw, v.offset, v.list,
                    v.heavy, d_entry_flags, d_service, d_runs, d_counters, tp);
    SVC_SIDE_LAUNCH(k_service_unmapped, (prm.n_batches + 127) / 128, 128, s_unmapped, prm, d_ordered, d_groups, d_batches, d_tbl,
                    v.batch_unmapped, d_entry_flags, d_service);
#undef SVC_SIDE_LAUNCH
    UVMTIER_CUDA(cudaEventRecord(ctx->aux_events[1], s_heavy));
    UVMTIER_CUDA(cudaEventRecord(ctx->aux_events[2], s_unmapped));
    // light va_blocks: four lanes each
    {
        // one wave: as many CTAs as the GPU keeps resident (fewer when the table is small); v.cursor[2] is the work counter
        uint32_t grid = (uint32_t)(((size_t)n_slots * kGrpLanes + kGrpThreads - 1) / kGrpThreads);
        if (grid > (uint32_t)ctx->sm_count * kGrpMinCtas)
            grid = (uint32_t)ctx->sm_count * kGrpMinCtas;
        k_service_blocks_grp<kGrpLanes, kGrpMinCtas><<<grid, kGrpThreads, 0, ctx->stream>>>(
            prm, d_ordered, d_groups, d_batches, bs, cw, v.offset, v.list, v.light, v.cursor + 1, d_entry_flags, d_service, d_runs,
            d_counters, v.cursor + 2);
        g_uvmtier_launches.fetch_add(1, std::memory_order_relaxed);
        const cudaError_t le = cudaGetLastError();
        if (le != cudaSuccess && st == NV_OK) {                 // the side streams are joined below all the same
            uvmtier_set_error("k_service_blocks_grp", le, __FILE__, __LINE__);
            st = uvmtier_status_from_cuda(le);
        }
    }
    UVMTIER_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->aux_events[1], 0));
    UVMTIER_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->aux_events[2], 0));
    return st;
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
    if (state->n_peers) {                                   // the servicing passes know one GPU (uvmtier_block_state_add_peer)
        uvmtier_set_error("the servicing passes take one-GPU state tables; this one names other GPUs (uvmtier_block_state_add_peer)",
                          cudaErrorNotSupported, __FILE__, __LINE__);
        return NV_ERR_NOT_SUPPORTED;
    }
    if (!n_batches)
        return NV_OK;
    if (!d_ordered || !d_groups || !d_batches || !d_entry_flags || !d_service || !d_counters || (run_cap && !d_runs))
        return NV_ERR_INVALID_ARGUMENT;
    const size_t n_slots64 = (size_t)lay->n_spaces * lay->blocks_per_space;
    if (!ctx->faults->gp.max_batch_size || n_slots64 == 0 || n_slots64 > kMaxSlots ||
        (size_t)n_batches * ctx->faults->gp.max_batch_size > 0xffffffffu)
        return NV_ERR_INVALID_ARGUMENT;
    if (n_batches > 65536u)                                 // uvmtier_service_group_t.batch is 16 bits wide
        return NV_ERR_INVALID_ARGUMENT;
    return NV_OK;
}

// Grows the context's working space to what a pass over n_batches batches needs.
static NV_STATUS service_scratch_reserve(uvmtier_ctx_t *ctx, uint32_t n_slots, uint32_t n_batches, size_t total)
{
    uvmtier_service_scratch *s = &ctx->svc;
    const size_t words = scratch_view(NULL, n_slots, n_batches, total).words;
    if (words > s->cap_words) {
        if (s->graph_exec) {                              // the graph holds the old working space's address
            cudaGraphExecDestroy(s->graph_exec);
            s->graph_exec = NULL;
        }
        s->graph_key_seen = 0;
        if (s->d_base) {
            UVMTIER_CUDA(cudaStreamSynchronize(ctx->stream));
            UVMTIER_CUDA(cudaFree(s->d_base));
            s->d_base = NULL;
            s->cap_words = 0;
        }
        UVMTIER_CUDA(cudaMalloc(&s->d_base, words * sizeof(uint32_t)));
        s->cap_words = words;
    }
    return NV_OK;
}

void uvmtier_service_scratch_destroy(uvmtier_ctx_t *ctx)
{
    uvmtier_service_scratch *s = &ctx->svc;
    if (s->graph_exec) cudaGraphExecDestroy(s->graph_exec);
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
    for (uint32_t k = 0; k < state->n_peers && k < UVMTIER_MAX_PEER_GPUS; k++)
        if (state->peer_resident[k])
            UVMTIER_CUDA(cudaFree(state->peer_resident[k]));
    memset(state, 0, sizeof(*state));
    return NV_OK;
}

NV_STATUS uvmtier_block_state_add_peer(uvmtier_ctx_t *ctx, const uvmtier_va_layout_t *lay, uvmtier_block_state_t *state,
                                       uint32_t peer_gpu_id, int direct)
{
    if (!ctx || !lay || !state || !state->gpu_resident || !lay->n_spaces || !lay->blocks_per_space)
        return NV_ERR_INVALID_ARGUMENT;
    if (state->n_peers >= UVMTIER_MAX_PEER_GPUS)
        return NV_ERR_INSUFFICIENT_RESOURCES;
    // ascending uvm_gpu_id_t (1 .. 0xFFFF fit the table), none of them this GPU
    if (!peer_gpu_id || peer_gpu_id > 0xffffu || peer_gpu_id == lay->gpu_id ||
        (state->n_peers && state->peer_gpu_id[state->n_peers - 1] >= peer_gpu_id))
        return NV_ERR_INVALID_ARGUMENT;
    const size_t n = (size_t)lay->n_spaces * lay->blocks_per_space;
    if (n > kMaxSlots)
        return NV_ERR_INVALID_ARGUMENT;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    void *p = NULL;
    UVMTIER_CUDA(cudaMalloc(&p, n * 64));
    UVMTIER_CUDA(cudaMemsetAsync(p, 0, n * 64, ctx->stream));
    UVMTIER_CUDA(cudaStreamSynchronize(ctx->stream));
    const uint32_t k = state->n_peers++;
    state->peer_resident[k] = (uvmtier_page_mask_t *)p;
    state->peer_gpu_id[k] = (uint16_t)peer_gpu_id;
    if (direct)
        state->peer_direct |= 1u << k;
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
    const NV_STATUS st = service_scratch_reserve(ctx, n_slots, n_batches, total);
    if (st != NV_OK)
        return st;
    return service_pass(ctx, ctx->svc.d_base, n_batches, d_ordered, d_groups, d_batches, lay, state, d_entry_flags, d_service,
                        d_runs, run_cap, d_counters);
}

// The same pass with the thrashing heuristic's hint inside the fault loop (UVM/uvm_gpu_replayable_faults.c:1524-1545).
NV_STATUS uvmtier_service_fault_batches_thrashing_device(uvmtier_ctx_t *ctx, uint32_t n_batches,
                                                         const uvmtier_ordered_fault_t *d_ordered,
                                                         const uvmtier_fault_group_t *d_groups, uvmtier_fault_batch_t *d_batches,
                                                         const uvmtier_va_layout_t *lay, const uvmtier_block_state_t *state,
                                                         uvmtier_thrashing_t *thrashing, const uint64_t *d_batch_time_ns,
                                                         uint16_t *d_entry_flags, uvmtier_service_group_t *d_service,
                                                         uvmtier_copy_run_t *d_runs, uint32_t run_cap, uint32_t *d_counters,
                                                         uint32_t *d_thrashing_counters)
{
    const NV_STATUS ok = service_check_args(ctx, n_batches, d_ordered, d_groups, d_batches, lay, state, d_entry_flags, d_service,
                                            d_runs, run_cap, d_counters);
    if (ok != NV_OK)
        return ok;
    if (!thrashing || !d_batch_time_ns || !d_thrashing_counters || !thrashing->d_blocks ||
        (size_t)lay->n_spaces * lay->blocks_per_space != thrashing->n_slots || lay->gpu_id >= 32u)
        return NV_ERR_INVALID_ARGUMENT;
    if (!thrashing->d_page_throttle)
        return NV_ERR_INVALID_STATE;                            // uvmtier_thrashing_enable_hints first
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    UVMTIER_CUDA(cudaMemsetAsync(d_thrashing_counters, 0, 8 * sizeof(uint32_t), ctx->stream));
    if (!n_batches)
        return NV_OK;
    const uint32_t n_slots = lay->n_spaces * lay->blocks_per_space;
    const size_t total = (size_t)n_batches * ctx->faults->gp.max_batch_size;
    const NV_STATUS st = service_scratch_reserve(ctx, n_slots, n_batches, total);
    if (st != NV_OK)
        return st;
    svc_thrash_params tp;
    tp.th.blocks = thrashing->d_blocks;
    tp.th.page_word = (unsigned long long *)thrashing->d_page_word;
    tp.th.page_procs = thrashing->d_page_processors;
    tp.th.thrashing_pages = thrashing->d_thrashing_pages;
    tp.th.page_throttle = (unsigned long long *)thrashing->d_page_throttle;
    tp.th.page_hint = thrashing->d_page_hint;
    tp.th.page_throttled = thrashing->d_page_throttled;
    tp.th.page_deadline = (unsigned long long *)thrashing->d_page_deadline;
    tp.th.pinned_pages = thrashing->d_pinned_pages;
    tp.cfg = thrashing->d_hint_config;
    tp.batch_time = (const unsigned long long *)d_batch_time_ns;
    tp.tally = d_thrashing_counters;
    return service_pass(ctx, ctx->svc.d_base, n_batches, d_ordered, d_groups, d_batches, lay, state, d_entry_flags, d_service,
                        d_runs, run_cap, d_counters, 0, true, false, &tp);
}

} // extern "C"

// A servicing pass over one chunk of a longer trace (pmm.cu): d_* point at the chunk, records carry batch numbers
// from batch_base, the counters and entry flags are the caller's to zero.
NV_STATUS uvmtier_service_pass_chunk(uvmtier_ctx_t *ctx, uint32_t n_batches, const uvmtier_ordered_fault_t *d_ordered,
                                     const uvmtier_fault_group_t *d_groups, uvmtier_fault_batch_t *d_batches,
                                     const uvmtier_va_layout_t *lay, const uvmtier_block_state_t *state, uint16_t *d_entry_flags,
                                     uvmtier_service_group_t *d_service, uvmtier_copy_run_t *d_runs, uint32_t run_cap,
                                     uint32_t *d_counters, uint32_t batch_base, bool non_replayable)
{
    const NV_STATUS ok = service_check_args(ctx, n_batches, d_ordered, d_groups, d_batches, lay, state, d_entry_flags, d_service,
                                            d_runs, run_cap, d_counters);
    if (ok != NV_OK || !n_batches)
        return ok;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    const NV_STATUS st = service_scratch_reserve(ctx, lay->n_spaces * lay->blocks_per_space, n_batches,
                                                 (size_t)n_batches * ctx->faults->gp.max_batch_size);
    if (st != NV_OK)
        return st;
    return service_pass(ctx, ctx->svc.d_base, n_batches, d_ordered, d_groups, d_batches, lay, state, d_entry_flags, d_service, d_runs,
                        run_cap, d_counters, batch_base, false, non_replayable);
}

// Access-counter notifications, servicing (include/uvmtier.h section 1e): the list front end of the fault path
// (runs per va_block, in batch order), then a warp per va_block -- the light list first, then the heavy one.
extern "C" NV_STATUS uvmtier_access_counters_service_device(uvmtier_ctx_t *ctx, uint32_t n_batches,
                                                            const uvmtier_ordered_notification_t *d_ordered,
                                                            const uvmtier_fault_group_t *d_groups, uvmtier_fault_batch_t *d_batches,
                                                            const uvmtier_va_layout_t *lay, const uvmtier_block_state_t *state,
                                                            uvmtier_service_group_t *d_service, uvmtier_copy_run_t *d_runs,
                                                            uint32_t run_cap, uint32_t *d_counters)
{
    const NV_STATUS ok = service_check_args(ctx, n_batches, d_ordered, d_groups, d_batches, lay, state, d_service, d_service, d_runs,
                                            run_cap, d_counters);
    if (ok != NV_OK || !n_batches)
        return ok;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    const uint32_t B = ctx->faults->gp.max_batch_size;
    const uint32_t n_slots = lay->n_spaces * lay->blocks_per_space;
    const size_t total = (size_t)n_batches * B;
    const NV_STATUS rs = service_scratch_reserve(ctx, n_slots, n_batches, total);
    if (rs != NV_OK)
        return rs;
    const service_scratch_view v = scratch_view(ctx->svc.d_base, n_slots, n_batches, total);
    service_kparams prm;
    memset(&prm, 0, sizeof(prm));
    prm.va_base = lay->va_base; prm.B = B; prm.n_batches = n_batches; prm.blocks_per_space = lay->blocks_per_space;
    prm.n_spaces = lay->n_spaces; prm.gpu_id = lay->gpu_id; prm.n_slots = n_slots;
    prm.pf_enable = ctx->cfg.uvm_perf_prefetch_enable; prm.pf_threshold = ctx->cfg.uvm_perf_prefetch_threshold;
    prm.pf_min_faults = ctx->cfg.uvm_perf_prefetch_min_faults; prm.run_cap = run_cap; prm.B_shift = svc_log2_or_none(B);
    block_state_ptrs bs;
    bs.gpu_resident = (uint32_t *)state->gpu_resident;
    bs.cpu_resident[0] = (uint32_t *)state->cpu_resident[0]; bs.cpu_resident[1] = (uint32_t *)state->cpu_resident[1];
    bs.gpu_pte[0] = (uint32_t *)state->gpu_pte[0]; bs.gpu_pte[1] = (uint32_t *)state->gpu_pte[1]; bs.gpu_pte[2] = (uint32_t *)state->gpu_pte[2];
    bs.prefetch_info = state->prefetch_info; bs.logical_prot = state->logical_prot; bs.preferred_location = state->preferred_location;
    bs.gpu_evicted = (uint32_t *)state->gpu_evicted; bs.read_duplicated = (uint32_t *)state->read_duplicated;
    UVMTIER_CUDA(cudaMemsetAsync(ctx->svc.d_base, 0, v.zero_words * sizeof(uint32_t), ctx->stream));
    UVMTIER_CUDA(cudaMemsetAsync(d_counters, 0, 2 * sizeof(uint32_t), ctx->stream));
    const uint2 *cw = (const uint2 *)v.cw;
    const size_t cap = (size_t)ctx->sm_count * 8;
    size_t grid = (total + kSvcThreads - 1) / kSvcThreads;
    if (grid > cap)
        grid = cap;
    UVMTIER_LAUNCH(k_service_count, (int)grid, kSvcThreads, 0, ctx->stream, prm, d_groups, d_batches, ctx->faults->d_table,
                   (unsigned long long *)v.cw, v.batch_unmapped, (uint2 *)v.slotpos);
    UVMTIER_LAUNCH(k_service_alloc, (n_slots + 255) / 256, 256, 0, ctx->stream, n_slots, cw, kHeavyMinEntries, v.offset, v.cursor, v.hist,
                   v.light_rank);
    UVMTIER_LAUNCH(k_service_fill, (int)grid, kSvcThreads, 0, ctx->stream, (uint32_t)total, (const uint2 *)v.slotpos, v.offset, v.list);
    size_t sgrid = ((size_t)n_slots + (kSortThreads / 32) - 1) / (kSortThreads / 32);
    if (sgrid > cap)
        sgrid = cap;
    UVMTIER_LAUNCH(k_service_sort_lists, (int)sgrid, kSortThreads, 0, ctx->stream, n_slots, cw, v.offset, v.list, kHeavyMinEntries, v.big,
                   v.heavy, v.hist, v.light_rank, v.light);
    UVMTIER_LAUNCH(k_service_sort_big, ctx->sm_count, kSortBigThreads, 0, ctx->stream, cw, v.offset, v.list, v.big);
    ac_service_cfg cfg;
    cfg.granularity_2m = ctx->faults->ac_granularity_2m;
    cfg.cpu_chunk_2m = ctx->faults->ac_cpu_chunk_2m;
    UVMTIER_LAUNCH(k_access_counters_blocks, ctx->sm_count * 4, kSvcThreads, 0, ctx->stream, prm, cfg, d_ordered, d_groups, bs, cw,
                   v.offset, v.list, (const uint32_t *)v.light, (const uint32_t *)(v.cursor + 1), d_service, d_runs, d_counters);
    UVMTIER_LAUNCH(k_access_counters_blocks, ctx->sm_count * 2, kSvcThreads, 0, ctx->stream, prm, cfg, d_ordered, d_groups, bs, cw,
                   v.offset, v.list, (const uint32_t *)(v.heavy + 1), (const uint32_t *)v.heavy, d_service, d_runs, d_counters);
    return NV_OK;
}

// Everything uvmtier_service_replayable_faults_streamed puts on the three streams.  Streams, events and working
// space exist already (streamed_prepare), so this is pure enqueueing and can run under stream capture.
struct streamed_args {
    const uvmtier_fault_packet_t *h_ring;
    uvmtier_fault_packet_t *d_ring;
    uint32_t ring_entries, get, put;
    int clear_valid;
    uint32_t chunk_batches;
    uint32_t *d_disposition;
    uvmtier_ordered_fault_t *d_ordered;
    uvmtier_fault_group_t *d_groups;
    uvmtier_fault_batch_t *d_batches;
    uvmtier_va_layout_t lay;
    uvmtier_block_state_t state;
    uint16_t *d_entry_flags;
    uvmtier_service_group_t *d_service;
    uvmtier_copy_run_t *d_runs;
    uint32_t run_cap;
    uint32_t *d_counters;
    uvmtier_service_readback_t rb;
    uint32_t tunables[4];                                   // what the kernels' parameters are derived from
    uvmtier_fault_gpu_params_t gp;
    const void *d_table;
    uint32_t table_rows, table_generation;
};
static_assert(sizeof(streamed_args) <= sizeof(((uvmtier_service_scratch *)0)->graph_key), "graph key too small");

static NV_STATUS streamed_prepare(uvmtier_ctx_t *ctx, const streamed_args &a, uvmtier_chunk_plan *plan, bool back)
{
    const uint32_t n = a.put >= a.get ? a.put - a.get : a.ring_entries - a.get + a.put;
    plan->B = ctx->faults->gp.max_batch_size;
    const uint32_t n_batches = (n + plan->B - 1) / plan->B;
    uint32_t per = a.chunk_batches ? a.chunk_batches : (n_batches + 7) / 8;
    if ((n_batches + per - 1) / per > UVMTIER_MAX_COPY_CHUNKS)
        per = (n_batches + UVMTIER_MAX_COPY_CHUNKS - 1) / UVMTIER_MAX_COPY_CHUNKS;
    const uint32_t n_chunks = (n_batches + per - 1) / per;
    if (!ctx->copy_stream)
        UVMTIER_CUDA(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
    if (back && !ctx->d2h_stream)
        UVMTIER_CUDA(cudaStreamCreateWithFlags(&ctx->d2h_stream, cudaStreamNonBlocking));
    if (!ctx->aux_stream[2])
        UVMTIER_CUDA(cudaStreamCreateWithFlags(&ctx->aux_stream[2], cudaStreamNonBlocking));
    for (uint32_t c = 0; c <= n_chunks; c++) {
        if (!ctx->copy_events[c])
            UVMTIER_CUDA(cudaEventCreateWithFlags(&ctx->copy_events[c], cudaEventDisableTiming));
        if (!ctx->k1_events[c])
            UVMTIER_CUDA(cudaEventCreateWithFlags(&ctx->k1_events[c], cudaEventDisableTiming));
        if (back && !ctx->d2h_events[c])
            UVMTIER_CUDA(cudaEventCreateWithFlags(&ctx->d2h_events[c], cudaEventDisableTiming));
    }
    return service_scratch_reserve(ctx, a.lay.n_spaces * a.lay.blocks_per_space, per, (size_t)per * plan->B);
}

static NV_STATUS streamed_enqueue(uvmtier_ctx_t *ctx, const streamed_args &a, bool back)
{
    const uvmtier_fault_packet_t *h_ring = a.h_ring;
    uvmtier_fault_packet_t *d_ring = a.d_ring;
    const uint32_t ring_entries = a.ring_entries, get = a.get, put = a.put, chunk_batches = a.chunk_batches, run_cap = a.run_cap;
    const int clear_valid = a.clear_valid;
    uint32_t *d_disposition = a.d_disposition, *d_counters = a.d_counters;
    uvmtier_ordered_fault_t *d_ordered = a.d_ordered;
    uvmtier_fault_group_t *d_groups = a.d_groups;
    uvmtier_fault_batch_t *d_batches = a.d_batches;
    const uvmtier_va_layout_t *lay = &a.lay;
    const uvmtier_block_state_t *state = &a.state;
    uint16_t *d_entry_flags = a.d_entry_flags;
    uvmtier_service_group_t *d_service = a.d_service;
    uvmtier_copy_run_t *d_runs = a.d_runs;
    const uvmtier_service_readback_t *rb = &a.rb;
    UVMTIER_CUDA(cudaMemsetAsync(d_counters, 0, 2 * sizeof(uint32_t), ctx->stream));
    uvmtier_chunk_plan plan;
    NV_STATUS st = uvmtier_streamed_begin(ctx, ring_entries, get, put, chunk_batches, &plan);
    if (st != NV_OK)
        return st;
    UVMTIER_CUDA(cudaMemsetAsync(d_entry_flags, 0, (size_t)plan.n_batches * plan.B * sizeof(uint16_t), ctx->stream));
    for (uint32_t c = 0; c < plan.n_chunks; c++) {
        uint32_t e0, e1, first, last;
        // the chunk is preprocessed on a stream of its own, next to the servicing of the chunk before it
        cudaStream_t s_k1 = ctx->aux_stream[2];
        st = uvmtier_streamed_copy_chunk(ctx, plan, c, h_ring, d_ring, ring_entries, get, &e0, &e1, &first, &last, s_k1);
        if (st != NV_OK)
            return st;
        const uint32_t b0 = c * plan.per, nb = (e1 - e0 + plan.B - 1) / plan.B;
        st = uvmtier_fetch_preprocess_on_stream(ctx, s_k1, d_ring, ring_entries, first, last, clear_valid, d_disposition + e0,
                                                d_ordered + e0, d_groups + e0, d_batches + b0, NULL, NULL, b0);
        if (st != NV_OK)
            return st;
        UVMTIER_CUDA(cudaEventRecord(ctx->k1_events[c], s_k1));
        UVMTIER_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->k1_events[c], 0));
        st = service_pass(ctx, ctx->svc.d_base, nb, d_ordered + e0, d_groups + e0, d_batches + b0, lay, state, d_entry_flags + e0,
                          d_service + e0, d_runs, run_cap, d_counters, b0, false);
        if (st != NV_OK)
            return st;
        if (back) {
            UVMTIER_CUDA(cudaEventRecord(ctx->d2h_events[c], ctx->stream));
            UVMTIER_CUDA(cudaStreamWaitEvent(ctx->d2h_stream, ctx->d2h_events[c], 0));
            if (rb->h_batches)
                UVMTIER_CUDA(cudaMemcpyAsync(rb->h_batches + b0, d_batches + b0, (size_t)nb * sizeof(*d_batches),
                                             cudaMemcpyDeviceToHost, ctx->d2h_stream));
            if (rb->h_entry_flags)
                UVMTIER_CUDA(cudaMemcpyAsync(rb->h_entry_flags + e0, d_entry_flags + e0, (size_t)(e1 - e0) * sizeof(uint16_t),
                                             cudaMemcpyDeviceToHost, ctx->d2h_stream));
        }
    }
    if (back) {
        // the counters are final after the last chunk; the context's stream then waits for the copy-back stream, so
        // that uvmtier_synchronize covers it
        if (rb->h_counters) {
            UVMTIER_CUDA(cudaEventRecord(ctx->d2h_events[plan.n_chunks], ctx->stream));
            UVMTIER_CUDA(cudaStreamWaitEvent(ctx->d2h_stream, ctx->d2h_events[plan.n_chunks], 0));
            UVMTIER_CUDA(cudaMemcpyAsync(rb->h_counters, d_counters, 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->d2h_stream));
        }
        UVMTIER_CUDA(cudaEventRecord(ctx->d2h_events[plan.n_chunks], ctx->d2h_stream));
        UVMTIER_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->d2h_events[plan.n_chunks], 0));
    }
    return NV_OK;
}

extern "C" {

NV_STATUS uvmtier_service_replayable_faults_streamed(uvmtier_ctx_t *ctx, const uvmtier_fault_packet_t *h_ring,
                                                     uvmtier_fault_packet_t *d_ring, uint32_t ring_entries, uint32_t get,
                                                     uint32_t put, int clear_valid, uint32_t chunk_batches,
                                                     uint32_t *d_disposition, uvmtier_ordered_fault_t *d_ordered,
                                                     uvmtier_fault_group_t *d_groups, uvmtier_fault_batch_t *d_batches,
                                                     const uvmtier_va_layout_t *lay, const uvmtier_block_state_t *state,
                                                     uint16_t *d_entry_flags, uvmtier_service_group_t *d_service,
                                                     uvmtier_copy_run_t *d_runs, uint32_t run_cap, uint32_t *d_counters,
                                                     const uvmtier_service_readback_t *rb, uint32_t *out_get)
{
    if (!ctx || !h_ring || !d_ring || !ring_entries || get >= ring_entries || put >= ring_entries)
        return NV_ERR_INVALID_ARGUMENT;
    const uint32_t n = put >= get ? put - get : ring_entries - get + put;
    if (out_get)
        *out_get = put;
    if (!d_disposition && n)
        return NV_ERR_INVALID_ARGUMENT;
    const uint32_t B_all = ctx->faults ? ctx->faults->gp.max_batch_size : 0;
    const NV_STATUS ok = service_check_args(ctx, B_all ? (n + B_all - 1) / B_all : 0, d_ordered, d_groups, d_batches, lay, state,
                                            d_entry_flags, d_service, d_runs, run_cap, d_counters);
    if (ok != NV_OK || !n)
        return ok;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));

    streamed_args a;
    memset(&a, 0, sizeof(a));                               // padding too: the struct is compared bytewise
    a.h_ring = h_ring; a.d_ring = d_ring; a.ring_entries = ring_entries; a.get = get; a.put = put;
    a.clear_valid = clear_valid != 0; a.chunk_batches = chunk_batches;
    a.d_disposition = d_disposition; a.d_ordered = d_ordered; a.d_groups = d_groups; a.d_batches = d_batches;
    a.lay = *lay; a.state = *state;
    a.d_entry_flags = d_entry_flags; a.d_service = d_service; a.d_runs = d_runs; a.run_cap = run_cap; a.d_counters = d_counters;
    if (rb)
        a.rb = *rb;
    a.tunables[0] = ctx->cfg.uvm_perf_prefetch_enable; a.tunables[1] = ctx->cfg.uvm_perf_prefetch_threshold;
    a.tunables[2] = ctx->cfg.uvm_perf_prefetch_min_faults; a.tunables[3] = ctx->cfg.uvm_perf_fault_coalesce;
    a.gp = ctx->faults->gp; a.d_table = ctx->faults->d_table; a.table_rows = ctx->faults->n_rows;
    a.table_generation = ctx->faults->generation;
    const bool back = a.rb.h_batches || a.rb.h_entry_flags || a.rb.h_counters;
    uvmtier_chunk_plan plan;
    NV_STATUS st = streamed_prepare(ctx, a, &plan, back);
    if (st != NV_OK)
        return st;

    // A servicing loop calls this with the same buffers over and over; the per-chunk work is some twenty stream
    // operations, and enqueueing them costs more host time than the GPU needs to execute them.  When a call repeats
    // the previous call's arguments exactly, everything it enqueues is captured into a CUDA graph once and replayed
    // from then on (UVMTIER_GRAPHS=0 turns this off).  The graph is dropped when the arguments, the tunables or the
    // fault-buffer configuration (every uvmtier_fault_buffer_init is a new generation) change, or the working space moves.
    static const char *graphs_env = getenv("UVMTIER_GRAPHS");
    const bool graphs = !(graphs_env && atoi(graphs_env) == 0);
    uvmtier_service_scratch *s = &ctx->svc;
    const bool same = s->graph_key_seen > 0 && memcmp(s->graph_key, &a, sizeof(a)) == 0;
    if (!same) {
        if (s->graph_exec) {
            cudaGraphExecDestroy(s->graph_exec);
            s->graph_exec = NULL;
        }
        memset(s->graph_key, 0, sizeof(s->graph_key));
        memcpy(s->graph_key, &a, sizeof(a));
        s->graph_key_seen = 1;
        return streamed_enqueue(ctx, a, back);
    }
    s->graph_key_seen++;
    if (!graphs)
        return streamed_enqueue(ctx, a, back);
    if (!s->graph_exec) {
        cudaGraph_t graph = NULL;
        const uint64_t before = g_uvmtier_launches.load();
        UVMTIER_CUDA(cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeThreadLocal));
        st = streamed_enqueue(ctx, a, back);
        const cudaError_t ce = cudaStreamEndCapture(ctx->stream, &graph);
        s->graph_launches = g_uvmtier_launches.load() - before;
        g_uvmtier_launches.fetch_sub(s->graph_launches);    // nothing ran yet; the replay below counts them
        if (st != NV_OK || ce != cudaSuccess || !graph) {
            if (graph)
                cudaGraphDestroy(graph);
            cudaGetLastError();
            s->graph_key_seen = 0;                          // do not try again with these arguments right away
            return st != NV_OK ? st : streamed_enqueue(ctx, a, back);
        }
        const cudaError_t ie = cudaGraphInstantiate(&s->graph_exec, graph, 0);
        cudaGraphDestroy(graph);
        if (ie != cudaSuccess) {
            s->graph_exec = NULL;
            cudaGetLastError();
            return streamed_enqueue(ctx, a, back);
        }
    }
    UVMTIER_CUDA(cudaGraphLaunch(s->graph_exec, ctx->stream));
    g_uvmtier_launches.fetch_add(s->graph_launches, std::memory_order_relaxed);
    return NV_OK;
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

// Called by uvmtier_init with the context's GPU current: resolves the array's device address once per GPU.  This
// first touch is also what makes the CUDA runtime load this file's module (and with it reserve the array), so it
// happens while the caller is setting up, not in a servicing call made under memory pressure.
NV_STATUS uvmtier_service_prepare(uvmtier_ctx_t *ctx)
{
    if (ctx->device < 0 || ctx->device >= 64)
        return NV_OK;                                                     // the no-alloc entry point rejects such a context
    if (!g_service_scratch_addr[ctx->device]) {
        uint32_t *p = NULL;
        UVMTIER_CUDA(cudaGetSymbolAddress((void **)&p, g_service_scratch));
        g_service_scratch_addr[ctx->device] = p;
    }
    return NV_OK;
}

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
    if (!g_service_scratch_addr[ctx->device])                             // uvmtier_init looked it up; cannot be missing
        return NV_ERR_INVALID_STATE;
    return service_pass(ctx, g_service_scratch_addr[ctx->device], n_batches, d_ordered, d_groups, d_batches, lay, state,
                        d_entry_flags, d_service, d_runs, run_cap, d_counters);
}
