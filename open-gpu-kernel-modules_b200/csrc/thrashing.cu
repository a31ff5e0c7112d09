// thrashing.cu -- host side of section (1f) of include/uvmtier.h: the detection half of the thrashing heuristic
// (UVM/uvm_perf_thrashing.c:1230-1380) on flat tables.  Kernels: thrashing_kernels.cuh.
#include "common.cuh"

#include "thrashing_kernels.cuh"

using namespace uvmtier_k;

static size_t thrash_scratch_words(uint32_t n_slots) { return 4 + (size_t)n_slots + ((size_t)n_slots + 31) / 32; }

extern "C" {

NV_STATUS uvmtier_thrashing_create(uvmtier_ctx_t *ctx, const uvmtier_va_layout_t *lay, uint64_t lapse_ns, uint32_t threshold,
                                   uvmtier_thrashing_t *out)
{
    if (!ctx || !lay || !out || !lay->n_spaces || !lay->blocks_per_space || threshold > 7)
        return NV_ERR_INVALID_ARGUMENT;
    const size_t n = (size_t)lay->n_spaces * lay->blocks_per_space;
    if (n > kMaxSlots)
        return NV_ERR_INVALID_ARGUMENT;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    memset(out, 0, sizeof(*out));
    const size_t bytes[5] = { n * sizeof(uvmtier_thrashing_block_t), n * 512 * sizeof(uint64_t), n * 512 * sizeof(uint32_t),
                              n * 16 * sizeof(uint32_t), thrash_scratch_words((uint32_t)n) * sizeof(uint32_t) };
    void **ptrs[5] = { (void **)&out->d_blocks, (void **)&out->d_page_word, (void **)&out->d_page_processors,
                       (void **)&out->d_thrashing_pages, (void **)&out->d_scratch };
    for (int i = 0; i < 5; i++) {
        if (cudaMalloc(ptrs[i], bytes[i]) != cudaSuccess) {
            cudaGetLastError();
            for (int j = 0; j < i; j++) cudaFree(*ptrs[j]);
            memset(out, 0, sizeof(*out));
            return NV_ERR_NO_MEMORY;
        }
        UVMTIER_CUDA(cudaMemsetAsync(*ptrs[i], 0, bytes[i], ctx->stream));
    }
    UVMTIER_CUDA(cudaStreamSynchronize(ctx->stream));
    out->lapse_ns = lapse_ns ? lapse_ns : 500ull * 1000;       // UVM_PERF_THRASHING_LAPSE_USEC_DEFAULT (:264)
    out->threshold = threshold ? threshold : 3;                // UVM_PERF_THRASHING_THRESHOLD_DEFAULT (:246)
    out->n_slots = (uint32_t)n;
    return NV_OK;
}

NV_STATUS uvmtier_thrashing_destroy(uvmtier_ctx_t *ctx, uvmtier_thrashing_t *t)
{
    if (!ctx || !t)
        return NV_ERR_INVALID_ARGUMENT;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    UVMTIER_CUDA(cudaStreamSynchronize(ctx->stream));
    if (t->d_blocks) cudaFree(t->d_blocks);
    if (t->d_page_word) cudaFree(t->d_page_word);
    if (t->d_page_processors) cudaFree(t->d_page_processors);
    if (t->d_thrashing_pages) cudaFree(t->d_thrashing_pages);
    if (t->d_scratch) cudaFree(t->d_scratch);
    memset(t, 0, sizeof(*t));
    return NV_OK;
}

NV_STATUS uvmtier_thrashing_note_events_device(uvmtier_ctx_t *ctx, uvmtier_thrashing_t *t, uint32_t n_events,
                                               const uvmtier_perf_event_t *d_events, uint32_t *d_counters)
{
    if (!ctx || !t || !t->d_blocks || !t->d_page_word || !t->d_page_processors || !t->d_thrashing_pages || !t->d_scratch ||
        !t->n_slots || !d_counters || !t->threshold || t->threshold > 7 || (n_events && !d_events))
        return NV_ERR_INVALID_ARGUMENT;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    UVMTIER_CUDA(cudaMemsetAsync(d_counters, 0, 4 * sizeof(uint32_t), ctx->stream));
    if (!n_events)
        return NV_OK;
    // the segment counter and the one-bit-per-slot table; the segment list between them is overwritten
    UVMTIER_CUDA(cudaMemsetAsync(t->d_scratch, 0, 4 * sizeof(uint32_t), ctx->stream));
    UVMTIER_CUDA(cudaMemsetAsync(t->d_scratch + 4 + t->n_slots, 0, ((size_t)t->n_slots + 31) / 32 * sizeof(uint32_t), ctx->stream));
    uint32_t grid = (n_events + kThrashThreads - 1) / kThrashThreads;
    if (grid > (uint32_t)ctx->sm_count * 8)
        grid = (uint32_t)ctx->sm_count * 8;
    UVMTIER_LAUNCH(k_thrash_heads, (int)grid, kThrashThreads, 0, ctx->stream, n_events, d_events, t->n_slots, t->d_scratch, d_counters);
    // one warp per va_block with events; at most n_slots of them
    const uint64_t segs = n_events < t->n_slots ? n_events : t->n_slots;
    // (no cap on the grid: a warp's chain of events is latency bound, the block scheduler balances the chains)
    const uint64_t grid2 = (segs * 32 + kThrashThreads - 1) / kThrashThreads;
    UVMTIER_LAUNCH(k_thrash_events, (int)grid2, kThrashThreads, 0, ctx->stream, n_events, d_events, t->n_slots,
                   (unsigned long long)t->lapse_ns, t->threshold, t->d_blocks, (unsigned long long *)t->d_page_word,
                   t->d_page_processors, t->d_thrashing_pages, t->d_scratch, d_counters);
    return NV_OK;
}

} // extern "C"
