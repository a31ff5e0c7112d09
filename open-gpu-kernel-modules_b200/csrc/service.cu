// service.cu -- C-ABI entry points of the per-va_block servicing pass (stage C/D/E) and the block state table.
// Reference: UVM/uvm_gpu_replayable_faults.c:2232-2377 (service_fault_batch) and callees; see service_kernels.cuh.
//
// Scratch in HBM (owned by the context, grown on demand, reused by every call): per-slot run counters, their
// exclusive scan and fill cursors (3 x 4 B x n_slots), the per-block run lists (4 B per run) and one flag per
// batch.  Algorithmic bytes per serviced va_block run: 16 (run) + 16 per entry + 2 per entry flag + 96 record
// + the block's state (6 masks x 64 B + 8 B) read and written once per touched block, + 8 per copy run.
#include "common.cuh"

#include <stdlib.h>

#include "service_kernels.cuh"
#include "service_tpb.cuh"

using namespace uvmtier_k;

// Tail of the pass, shared by the two entry points: sort every slot's run list, service the va_blocks (one thread
// per block), then the runs that belong to no va_range.  UVMTIER_SERVICE_WARP_PER_BLOCK=1 selects the older
// warp-per-block kernel (kept for comparison; same results).
static NV_STATUS launch_block_service(uvmtier_ctx_t *ctx, const service_kparams &prm, const uvmtier_ordered_fault_t *d_ordered,
                                      const uvmtier_fault_group_t *d_groups, uvmtier_fault_batch_t *d_batches,
                                      const uvmtier_instance_ptr_entry_t *d_tbl, const block_state_ptrs &bs, uint32_t *d_count,
                                      uint32_t *d_offset, uint32_t *d_list, uint32_t *d_batch_unmapped, uint16_t *d_entry_flags,
                                      uvmtier_service_group_t *d_service, uvmtier_copy_run_t *d_runs, uint32_t *d_counters)
{
    static const bool warp_per_block = getenv("UVMTIER_SERVICE_WARP_PER_BLOCK") != NULL;
    const size_t cap = (size_t)ctx->sm_count * 8;
    if (warp_per_block) {
        size_t bgrid = ((size_t)prm.n_slots + (kSvcThreads / 32) - 1) / (kSvcThreads / 32);
        if (bgrid > cap)
            bgrid = cap;
        UVMTIER_LAUNCH(k_service_blocks, (int)bgrid, kSvcThreads, 0, ctx->stream, prm, d_ordered, d_groups, d_batches, bs, d_count,
                       d_offset, d_list, d_entry_flags, d_service, d_runs, d_counters);
    }
    else {
        static bool attr_set[64];
        if (!attr_set[ctx->device & 63]) {
            // opt in to the 16 KB of dynamic shared memory of the list sorter (below the 48 KB default: no-op, kept explicit)
            UVMTIER_CUDA(cudaFuncSetAttribute(k_service_sort_lists, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                              (int)(kSortSmemKeys * sizeof(uint32_t))));
            attr_set[ctx->device & 63] = true;
        }
        size_t sgrid = prm.n_slots < (size_t)ctx->sm_count * 32 ? prm.n_slots : (size_t)ctx->sm_count * 32;
        UVMTIER_LAUNCH(k_service_sort_lists, (int)sgrid, 32, kSortSmemKeys * sizeof(uint32_t), ctx->stream, prm.n_slots, d_count,
                       d_offset, d_list);
        const uint32_t tgrid = (prm.n_slots + kTpbThreads - 1) / kTpbThreads;
        UVMTIER_LAUNCH(k_service_blocks_tpb, tgrid, kTpbThreads, 0, ctx->stream, prm, d_ordered, d_groups, d_batches, bs, d_count,
                       d_offset, d_list, d_entry_flags, d_service, d_runs, d_counters);
    }
    UVMTIER_LAUNCH(k_service_unmapped, (prm.n_batches + 127) / 128, 128, 0, ctx->stream, prm, d_ordered, d_groups, d_batches, d_tbl,
                   d_batch_unmapped, d_entry_flags, d_service);
    return NV_OK;
}

static NV_STATUS scratch_reserve(uvmtier_ctx_t *ctx, uint32_t n_slots, uint32_t n_batches, size_t list_entries)
{
    uvmtier_service_scratch *s = &ctx->svc;
    if (n_slots > s->n_slots_cap) {
        if (s->d_count) { cudaStreamSynchronize(ctx->stream); cudaFree(s->d_count); s->d_count = NULL; s->n_slots_cap = 0; }
        // count | offset | cursor | sums(1024)
        UVMTIER_CUDA(cudaMalloc(&s->d_count, ((size_t)n_slots * 3 + 1024) * sizeof(uint32_t)));
        s->d_offset = s->d_count + n_slots;
        s->d_cursor = s->d_offset + n_slots;
        s->d_sums = s->d_cursor + n_slots;
        s->n_slots_cap = n_slots;
    }
    else if (s->d_count) {
        s->d_offset = s->d_count + s->n_slots_cap;
        s->d_cursor = s->d_offset + s->n_slots_cap;
        s->d_sums = s->d_cursor + s->n_slots_cap;
    }
    if (n_batches > s->n_batches_cap) {
        if (s->d_batch_unmapped) { cudaStreamSynchronize(ctx->stream); cudaFree(s->d_batch_unmapped); s->d_batch_unmapped = NULL; }
        UVMTIER_CUDA(cudaMalloc(&s->d_batch_unmapped, (size_t)n_batches * sizeof(uint32_t)));
        s->n_batches_cap = n_batches;
    }
    if (list_entries > s->list_cap) {
        if (s->d_list) { cudaStreamSynchronize(ctx->stream); cudaFree(s->d_list); s->d_list = NULL; }
        UVMTIER_CUDA(cudaMalloc(&s->d_list, list_entries * sizeof(uint32_t)));
        s->list_cap = list_entries;
    }
    return NV_OK;
}

void uvmtier_service_scratch_destroy(uvmtier_ctx_t *ctx)
{
    uvmtier_service_scratch *s = &ctx->svc;
    if (s->d_count) cudaFree(s->d_count);
    if (s->d_batch_unmapped) cudaFree(s->d_batch_unmapped);
    if (s->d_list) cudaFree(s->d_list);
    memset(s, 0, sizeof(*s));
}

extern "C" {

NV_STATUS uvmtier_block_state_alloc(uvmtier_ctx_t *ctx, const uvmtier_va_layout_t *lay, uvmtier_block_state_t *out)
{
    if (!ctx || !lay || !out || !lay->n_spaces || !lay->blocks_per_space)
        return NV_ERR_INVALID_ARGUMENT;
    const size_t n = (size_t)lay->n_spaces * lay->blocks_per_space;
    if (n > (size_t)1024 * kScanItems)
        return NV_ERR_INVALID_ARGUMENT;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    memset(out, 0, sizeof(*out));
    // one allocation: 6 mask arrays, prefetch_info, preferred_location, logical_prot
    const size_t mask_bytes = n * 64;
    char *base = NULL;
    UVMTIER_CUDA(cudaMalloc(&base, 6 * mask_bytes + n * 4 + n * 2 + n));
    out->gpu_resident = (uvmtier_page_mask_t *)base;
    out->cpu_resident[0] = (uvmtier_page_mask_t *)(base + mask_bytes);
    out->cpu_resident[1] = (uvmtier_page_mask_t *)(base + 2 * mask_bytes);
    out->gpu_pte[0] = (uvmtier_page_mask_t *)(base + 3 * mask_bytes);
    out->gpu_pte[1] = (uvmtier_page_mask_t *)(base + 4 * mask_bytes);
    out->gpu_pte[2] = (uvmtier_page_mask_t *)(base + 5 * mask_bytes);
    out->prefetch_info = (uvmtier_prefetch_info_t *)(base + 6 * mask_bytes);
    out->preferred_location = (uint16_t *)(base + 6 * mask_bytes + n * 4);
    out->logical_prot = (uint8_t *)(base + 6 * mask_bytes + n * 4 + n * 2);
    UVMTIER_CUDA(cudaMemsetAsync(base, 0, 6 * mask_bytes, ctx->stream));
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
    memset(state, 0, sizeof(*state));
    return NV_OK;
}

NV_STATUS uvmtier_service_fault_batches_device(uvmtier_ctx_t *ctx, uint32_t n_batches,
                                               const uvmtier_ordered_fault_t *d_ordered,
                                               const uvmtier_fault_group_t *d_groups, uvmtier_fault_batch_t *d_batches,
                                               const uvmtier_va_layout_t *lay, const uvmtier_block_state_t *state,
                                               uint16_t *d_entry_flags, uvmtier_service_group_t *d_service,
                                               uvmtier_copy_run_t *d_runs, uint32_t run_cap, uint32_t *d_counters)
{
    if (!ctx || !lay || !state)
        return NV_ERR_INVALID_ARGUMENT;
    if (!ctx->faults)
        return NV_ERR_INVALID_STATE;
    if (!n_batches)
        return NV_OK;
    if (!d_ordered || !d_groups || !d_batches || !d_entry_flags || !d_service || !d_counters || (run_cap && !d_runs))
        return NV_ERR_INVALID_ARGUMENT;
    const uint32_t B = ctx->faults->gp.max_batch_size;
    const size_t n_slots64 = (size_t)lay->n_spaces * lay->blocks_per_space;
    if (!B || n_slots64 == 0 || n_slots64 > (size_t)1024 * kScanItems)
        return NV_ERR_INVALID_ARGUMENT;
    const uint32_t n_slots = (uint32_t)n_slots64;
    const size_t total = (size_t)n_batches * B;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    NV_STATUS st = scratch_reserve(ctx, n_slots, n_batches, total);
    if (st != NV_OK)
        return st;
    uvmtier_service_scratch *s = &ctx->svc;

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

    // count + cursor are zeroed together (offset lies between them and is fully rewritten by the scan)
    UVMTIER_CUDA(cudaMemsetAsync(s->d_count, 0, (size_t)n_slots * sizeof(uint32_t), ctx->stream));
    UVMTIER_CUDA(cudaMemsetAsync(s->d_cursor, 0, (size_t)n_slots * sizeof(uint32_t), ctx->stream));
    UVMTIER_CUDA(cudaMemsetAsync(s->d_batch_unmapped, 0, (size_t)n_batches * sizeof(uint32_t), ctx->stream));
    UVMTIER_CUDA(cudaMemsetAsync(d_counters, 0, 2 * sizeof(uint32_t), ctx->stream));
    UVMTIER_CUDA(cudaMemsetAsync(d_entry_flags, 0, total * sizeof(uint16_t), ctx->stream));

    const uvmtier_instance_ptr_entry_t *d_tbl = ctx->faults->d_table;
    size_t grid = (total + kSvcThreads - 1) / kSvcThreads, cap = (size_t)ctx->sm_count * 8;
    if (grid > cap)
        grid = cap;
    UVMTIER_LAUNCH(k_service_count, (int)grid, kSvcThreads, 0, ctx->stream, prm, d_groups, d_batches, d_tbl, s->d_count,
                   s->d_batch_unmapped, s->d_list, 0, s->d_offset, s->d_cursor);
    const uint32_t n_sums = (n_slots + kScanItems - 1) / kScanItems;
    UVMTIER_LAUNCH(k_scan_sums, n_sums, 256, 0, ctx->stream, s->d_count, n_slots, s->d_sums);
    UVMTIER_LAUNCH(k_scan_top, 1, 1024, 0, ctx->stream, s->d_sums, n_sums);
    UVMTIER_LAUNCH(k_scan_final, n_sums, 256, 0, ctx->stream, s->d_count, n_slots, s->d_sums, s->d_offset);
    UVMTIER_LAUNCH(k_service_count, (int)grid, kSvcThreads, 0, ctx->stream, prm, d_groups, d_batches, d_tbl, s->d_count,
                   s->d_batch_unmapped, s->d_list, 1, s->d_offset, s->d_cursor);
    return launch_block_service(ctx, prm, d_ordered, d_groups, d_batches, d_tbl, bs, s->d_count, s->d_offset, s->d_list,
                                s->d_batch_unmapped, d_entry_flags, d_service, d_runs, d_counters);
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
// What: the pass above keeps its working space (per-slot run counters, their scan, fill cursors, the per-block
// run lists, one flag per batch) in buffers it grows with cudaMalloc.  This variant keeps the same working space
// in one fixed array that is part of the library image.  Every GPU that loads the library carries the array
// (about 1.8 MB) until the process exits; the price of never allocating is a hard ceiling on the problem size.
constexpr uint32_t kNoAllocMaxSlots = 65536;        // va_block slots: 128 GB of VA
constexpr uint32_t kNoAllocMaxEntries = 262144;     // n_batches * max_batch_size
constexpr uint32_t kNoAllocMaxBatches = 4096;
constexpr size_t kNoAllocWords = (size_t)kNoAllocMaxSlots * 3 + 1024 + kNoAllocMaxBatches + kNoAllocMaxEntries;

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
    if (!ctx || !lay || !state)
        return NV_ERR_INVALID_ARGUMENT;
    if (!ctx->faults)
        return NV_ERR_INVALID_STATE;
    if (!n_batches)
        return NV_OK;
    if (!d_ordered || !d_groups || !d_batches || !d_entry_flags || !d_service || !d_counters || (run_cap && !d_runs))
        return NV_ERR_INVALID_ARGUMENT;
    const uint32_t B = ctx->faults->gp.max_batch_size;
    const size_t n_slots64 = (size_t)lay->n_spaces * lay->blocks_per_space;
    const size_t total = (size_t)n_batches * B;
    if (!B || !n_slots64 || ctx->device < 0 || ctx->device >= 64)
        return NV_ERR_INVALID_ARGUMENT;
    if (n_slots64 > kNoAllocMaxSlots || total > kNoAllocMaxEntries || n_batches > kNoAllocMaxBatches) {
        char msg[256];
        snprintf(msg, sizeof(msg),
                 "uvmtier_service_fault_batches_noalloc_device: %zu va_block slots / %zu entries / %u batches exceed the "
                 "fixed scratch (%u slots, %u entries, %u batches); use uvmtier_service_fault_batches_device",
                 n_slots64, total, n_batches, kNoAllocMaxSlots, kNoAllocMaxEntries, kNoAllocMaxBatches);
        uvmtier_set_error(msg, cudaErrorInvalidValue, __FILE__, __LINE__);
        return NV_ERR_INVALID_ARGUMENT;
    }
    const uint32_t n_slots = (uint32_t)n_slots64;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    if (!g_service_scratch_addr[ctx->device]) {
        uint32_t *p = NULL;
        UVMTIER_CUDA(cudaGetSymbolAddress((void **)&p, g_service_scratch));
        g_service_scratch_addr[ctx->device] = p;
    }
    uint32_t *d_count = g_service_scratch_addr[ctx->device];
    uint32_t *d_offset = d_count + kNoAllocMaxSlots;
    uint32_t *d_cursor = d_offset + kNoAllocMaxSlots;
    uint32_t *d_sums = d_cursor + kNoAllocMaxSlots;
    uint32_t *d_batch_unmapped = d_sums + 1024;
    uint32_t *d_list = d_batch_unmapped + kNoAllocMaxBatches;

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

    // The array still holds the previous call's numbers: the counters, the cursors and the per-batch flags are
    // accumulated into and must start at zero; offsets, partial sums and lists are fully rewritten before use.
    UVMTIER_CUDA(cudaMemsetAsync(d_count, 0, (size_t)n_slots * sizeof(uint32_t), ctx->stream));
    UVMTIER_CUDA(cudaMemsetAsync(d_cursor, 0, (size_t)n_slots * sizeof(uint32_t), ctx->stream));
    UVMTIER_CUDA(cudaMemsetAsync(d_batch_unmapped, 0, (size_t)n_batches * sizeof(uint32_t), ctx->stream));
    UVMTIER_CUDA(cudaMemsetAsync(d_counters, 0, 2 * sizeof(uint32_t), ctx->stream));
    UVMTIER_CUDA(cudaMemsetAsync(d_entry_flags, 0, total * sizeof(uint16_t), ctx->stream));

    const uvmtier_instance_ptr_entry_t *d_tbl = ctx->faults->d_table;
    size_t grid = (total + kSvcThreads - 1) / kSvcThreads, cap = (size_t)ctx->sm_count * 8;
    if (grid > cap)
        grid = cap;
    UVMTIER_LAUNCH(k_service_count, (int)grid, kSvcThreads, 0, ctx->stream, prm, d_groups, d_batches, d_tbl, d_count,
                   d_batch_unmapped, d_list, 0, d_offset, d_cursor);
    const uint32_t n_sums = (n_slots + kScanItems - 1) / kScanItems;
    UVMTIER_LAUNCH(k_scan_sums, n_sums, 256, 0, ctx->stream, d_count, n_slots, d_sums);
    UVMTIER_LAUNCH(k_scan_top, 1, 1024, 0, ctx->stream, d_sums, n_sums);
    UVMTIER_LAUNCH(k_scan_final, n_sums, 256, 0, ctx->stream, d_count, n_slots, d_sums, d_offset);
    UVMTIER_LAUNCH(k_service_count, (int)grid, kSvcThreads, 0, ctx->stream, prm, d_groups, d_batches, d_tbl, d_count,
                   d_batch_unmapped, d_list, 1, d_offset, d_cursor);
    return launch_block_service(ctx, prm, d_ordered, d_groups, d_batches, d_tbl, bs, d_count, d_offset, d_list, d_batch_unmapped,
                                d_entry_flags, d_service, d_runs, d_counters);
}
