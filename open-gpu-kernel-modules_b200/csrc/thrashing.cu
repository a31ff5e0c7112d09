// thrashing.cu -- host side of section (1f) of include/uvmtier.h: the thrashing heuristic (UVM/uvm_perf_thrashing.c: the
// detector :1230-1380, the hint :1382-1780, pin expiry :1822-1906) on flat tables.  Kernels: thrashing_kernels.cuh,
// thrashing_hint.cuh.
#include "common.cuh"

#include <stddef.h>

#include "thrashing_kernels.cuh"

using namespace uvmtier_k;

static size_t thrash_scratch_words(uint32_t n_slots) { return 4 + 2 * (size_t)n_slots; }

static thrash_hint_tables thrash_tables(const uvmtier_thrashing_t *t)
{
    thrash_hint_tables th;
    th.blocks = t->d_blocks;
    th.page_word = (unsigned long long *)t->d_page_word;
    th.page_procs = t->d_page_processors;
    th.thrashing_pages = t->d_thrashing_pages;
    th.page_throttle = (unsigned long long *)t->d_page_throttle;
    th.page_hint = t->d_page_hint;
    th.page_throttled = t->d_page_throttled;
    th.page_deadline = (unsigned long long *)t->d_page_deadline;
    th.pinned_pages = t->d_pinned_pages;
    return th;
}

static bool thrash_ok(const uvmtier_thrashing_t *t)
{
    return t && t->d_blocks && t->d_page_word && t->d_page_processors && t->d_thrashing_pages && t->d_scratch && t->n_slots &&
           t->threshold && t->threshold <= 7;
}

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
    if (t->d_page_throttle) cudaFree(t->d_page_throttle);
    if (t->d_page_hint) cudaFree(t->d_page_hint);
    if (t->d_page_throttled) cudaFree(t->d_page_throttled);
    if (t->d_page_deadline) cudaFree(t->d_page_deadline);
    if (t->d_pinned_pages) cudaFree(t->d_pinned_pages);
    if (t->d_hint_config) cudaFree(t->d_hint_config);
    memset(t, 0, sizeof(*t));
    return NV_OK;
}

NV_STATUS uvmtier_thrashing_note_events_device(uvmtier_ctx_t *ctx, uvmtier_thrashing_t *t, uint32_t n_events,
                                               const uvmtier_perf_event_t *d_events, uint32_t *d_counters)
{
    if (!ctx || !thrash_ok(t) || !d_counters || (n_events && !d_events))
        return NV_ERR_INVALID_ARGUMENT;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    UVMTIER_CUDA(cudaMemsetAsync(d_counters, 0, 4 * sizeof(uint32_t), ctx->stream));
    if (!n_events)
        return NV_OK;
    // the segment counter and the one-bit-per-slot table; the segment list between them is overwritten
    UVMTIER_CUDA(cudaMemsetAsync(t->d_scratch, 0, 4 * sizeof(uint32_t), ctx->stream));
    UVMTIER_CUDA(cudaMemsetAsync(t->d_scratch + 4 + t->n_slots, 0xff, (size_t)t->n_slots * sizeof(uint32_t), ctx->stream));
    uint32_t grid = (n_events + kThrashThreads - 1) / kThrashThreads;
    if (grid > (uint32_t)ctx->sm_count * 8)
        grid = (uint32_t)ctx->sm_count * 8;
    UVMTIER_LAUNCH(k_thrash_heads_min, (int)grid, kThrashThreads, 0, ctx->stream, n_events, d_events, t->n_slots, t->d_scratch);
    UVMTIER_LAUNCH(k_thrash_heads, (int)grid, kThrashThreads, 0, ctx->stream, n_events, d_events, t->n_slots, t->d_scratch, d_counters);
    // one warp per va_block with events; at most n_slots of them
    const uint64_t segs = n_events < t->n_slots ? n_events : t->n_slots;
    // (no cap on the grid: a warp's chain of events is latency bound, the block scheduler balances the chains)
    const uint64_t grid2 = (segs * 32 + kThrashThreads - 1) / kThrashThreads;
    UVMTIER_LAUNCH(k_thrash_events, (int)grid2, kThrashThreads, 0, ctx->stream, n_events, d_events, t->n_slots,
                   (unsigned long long)t->lapse_ns, t->threshold, t->d_blocks, (unsigned long long *)t->d_page_word,
                   t->d_page_processors, t->d_thrashing_pages, t->d_scratch, d_counters, thrash_tables(t));
    return NV_OK;
}

void uvmtier_thrashing_default_hint_config(uvmtier_thrashing_hint_config_t *cfg, uint32_t gpu_mask, int fast_links)
{
    if (!cfg)
        return;
    memset(cfg, 0, sizeof(*cfg));
    cfg->lapse_ns = 500ull * 1000;             // uvm_perf_thrashing.c:246-310
    cfg->threshold = 3;
    cfg->pin_threshold = 10;
    cfg->nap_ns = 1 * cfg->lapse_ns;
    cfg->epoch_ns = 2000 * cfg->lapse_ns;
    cfg->pin_ns = 300 * cfg->lapse_ns;
    cfg->max_resets = 4;
    gpu_mask &= ~1u;
    const uint32_t all = gpu_mask | 1u;
    cfg->has_memory = all;
    for (uint32_t p = 0; p < 32; p++) {
        if (!((all >> p) & 1u))
            continue;
        const uint32_t me = 1u << p;
        if (fast_links >= 2) {
            cfg->accessible_from[p] = cfg->can_access[p] = cfg->has_fast_link[p] = cfg->has_native_atomics[p] = all;
            continue;
        }
        cfg->accessible_from[p] = p == 0 ? all : gpu_mask;      // who reaches p's memory
        cfg->can_access[p] = p == 0 ? me : all;                 // whose memory p reaches
        cfg->has_fast_link[p] = (fast_links == 1 && p != 0) ? (gpu_mask & ~me) : 0u;
        cfg->has_native_atomics[p] = me | cfg->has_fast_link[p];
    }
}

NV_STATUS uvmtier_thrashing_enable_hints(uvmtier_ctx_t *ctx, uvmtier_thrashing_t *t, const uvmtier_thrashing_hint_config_t *cfg)
{
    if (!ctx || !thrash_ok(t) || !cfg || cfg->pin_threshold > 255u)
        return NV_ERR_INVALID_ARGUMENT;
    if (t->d_page_throttle)
        return NV_ERR_INVALID_STATE;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    const size_t n = t->n_slots;
    const size_t bytes[6] = { n * 512 * sizeof(uint64_t), n * 512 * sizeof(uint32_t), n * 512 * sizeof(uint32_t), n * 512 * sizeof(uint64_t),
                              n * 16 * sizeof(uint32_t), sizeof(uvmtier_thrashing_hint_config_t) };
    void **ptrs[6] = { (void **)&t->d_page_throttle, (void **)&t->d_page_hint, (void **)&t->d_page_throttled, (void **)&t->d_page_deadline,
                       (void **)&t->d_pinned_pages, (void **)&t->d_hint_config };
    for (int i = 0; i < 6; i++) {
        if (cudaMalloc(ptrs[i], bytes[i]) != cudaSuccess) {
            cudaGetLastError();
            for (int j = 0; j < i; j++) { cudaFree(*ptrs[j]); *ptrs[j] = NULL; }
            *ptrs[i] = NULL;
            return NV_ERR_NO_MEMORY;
        }
        UVMTIER_CUDA(cudaMemsetAsync(*ptrs[i], 0, bytes[i], ctx->stream));
    }
    t->hint_config = *cfg;
    t->hint_config.lapse_ns = t->lapse_ns;
    t->hint_config.threshold = t->threshold;
    UVMTIER_CUDA(cudaMemcpyAsync(t->d_hint_config, &t->hint_config, sizeof(t->hint_config), cudaMemcpyHostToDevice, ctx->stream));
    UVMTIER_CUDA(cudaStreamSynchronize(ctx->stream));
    return NV_OK;
}

NV_STATUS uvmtier_thrashing_get_hints_device(uvmtier_ctx_t *ctx, uvmtier_thrashing_t *t, uint32_t n_queries,
                                             const uvmtier_thrashing_query_t *d_queries, uvmtier_thrashing_hint_t *d_hints,
                                             uint32_t *d_counters)
{
    if (!ctx || !thrash_ok(t) || !d_counters || (n_queries && (!d_queries || !d_hints)))
        return NV_ERR_INVALID_ARGUMENT;
    if (!t->d_page_throttle)
        return NV_ERR_INVALID_STATE;                // uvmtier_thrashing_enable_hints first
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    UVMTIER_CUDA(cudaMemsetAsync(d_counters, 0, 8 * sizeof(uint32_t), ctx->stream));
    if (!n_queries)
        return NV_OK;
    UVMTIER_CUDA(cudaMemsetAsync(d_hints, 0, (size_t)n_queries * sizeof(uvmtier_thrashing_hint_t), ctx->stream));
    UVMTIER_CUDA(cudaMemsetAsync(t->d_scratch, 0, 4 * sizeof(uint32_t), ctx->stream));
    UVMTIER_CUDA(cudaMemsetAsync(t->d_scratch + 4 + t->n_slots, 0xff, (size_t)t->n_slots * sizeof(uint32_t), ctx->stream));
    uint32_t grid = (n_queries + kThrashThreads - 1) / kThrashThreads;
    if (grid > (uint32_t)ctx->sm_count * 8)
        grid = (uint32_t)ctx->sm_count * 8;
    // the groups of queries: the query carries its slot where the event does, and has the event's size
    static_assert(sizeof(uvmtier_thrashing_query_t) == sizeof(uvmtier_perf_event_t) &&
                  offsetof(uvmtier_thrashing_query_t, slot) == offsetof(uvmtier_perf_event_t, slot), "k_thrash_heads reads .slot only");
    UVMTIER_LAUNCH(k_thrash_heads_min, (int)grid, kThrashThreads, 0, ctx->stream, n_queries, (const uvmtier_perf_event_t *)d_queries,
                   t->n_slots, t->d_scratch);
    UVMTIER_LAUNCH(k_thrash_heads, (int)grid, kThrashThreads, 0, ctx->stream, n_queries, (const uvmtier_perf_event_t *)d_queries,
                   t->n_slots, t->d_scratch, d_counters);
    const uint64_t segs = n_queries < t->n_slots ? n_queries : t->n_slots;
    const uint64_t grid2 = (segs + 127) / 128;
    UVMTIER_LAUNCH(k_thrash_hints, (int)grid2, 128, 0, ctx->stream, n_queries, d_queries, t->n_slots,
                   (const uvmtier_thrashing_hint_config_t *)t->d_hint_config, thrash_tables(t), (const uint32_t *)t->d_scratch, d_hints,
                   d_counters);
    return NV_OK;
}

NV_STATUS uvmtier_thrashing_unpin_expired_device(uvmtier_ctx_t *ctx, uvmtier_thrashing_t *t, uint64_t now_ns, uint32_t *d_counters)
{
    if (!ctx || !thrash_ok(t) || !d_counters)
        return NV_ERR_INVALID_ARGUMENT;
    if (!t->d_page_throttle)
        return NV_ERR_INVALID_STATE;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    UVMTIER_CUDA(cudaMemsetAsync(d_counters, 0, 8 * sizeof(uint32_t), ctx->stream));
    uint64_t grid = ((uint64_t)t->n_slots * 32 + 127) / 128;
    if (grid > (uint64_t)ctx->sm_count * 16)
        grid = (uint64_t)ctx->sm_count * 16;
    UVMTIER_LAUNCH(k_thrash_unpin_expired, (int)grid, 128, 0, ctx->stream, t->n_slots, (unsigned long long)now_ns, thrash_tables(t),
                   d_counters);
    return NV_OK;
}

} // extern "C"
