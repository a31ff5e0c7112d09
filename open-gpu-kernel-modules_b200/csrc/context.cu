// context.cu -- context lifetime, tunable validation, memory helpers of libuvmtier.
#include "common.cuh"

#include <stdlib.h>

std::atomic<uint64_t> g_uvmtier_launches{0};
static thread_local char g_last_error[512] = "";

void uvmtier_set_error(const char *what, cudaError_t err, const char *file, int line)
{
    snprintf(g_last_error, sizeof(g_last_error), "%s:%d: %s: %s", file, line, what, cudaGetErrorString(err));
    if (getenv("UVMTIER_VERBOSE"))
        fprintf(stderr, "uvmtier: %s\n", g_last_error);
}

void uvmtier_fault_state_destroy(uvmtier_ctx_t *ctx);
void uvmtier_tiers_destroy(uvmtier_ctx_t *ctx);
void uvmtier_service_scratch_destroy(uvmtier_ctx_t *ctx);
NV_STATUS uvmtier_service_prepare(uvmtier_ctx_t *ctx);

extern "C" {

const char *uvmtier_version(void) { return "uvmtier " UVMTIER_VERSION " sm_100a"; }
uint64_t uvmtier_kernel_launch_count(void) { return g_uvmtier_launches.load(); }
const char *uvmtier_last_error(void) { return g_last_error; }

void uvmtier_config_default(uvmtier_config_t *cfg)
{
    memset(cfg, 0, sizeof(*cfg));
    cfg->uvm_perf_prefetch_enable = 1;        // UVM/uvm_perf_prefetch.c:39
    cfg->uvm_perf_prefetch_threshold = 51;    // :42
    cfg->uvm_perf_prefetch_min_faults = 1;    // :51
    cfg->uvm_perf_fault_batch_count = 256;    // UVM/uvm_gpu_replayable_faults.c:72-78
    cfg->uvm_perf_fault_coalesce = 1;         // :115
}

// Same fallbacks as uvm_perf_prefetch_init (UVM/uvm_perf_prefetch.c:504-535) and
// fault_buffer_init_replayable_faults (UVM/uvm_gpu_replayable_faults.c:152-165).
static void validate_config(uvmtier_config_t *c)
{
    c->uvm_perf_prefetch_enable = c->uvm_perf_prefetch_enable != 0;
    if (c->uvm_perf_prefetch_threshold > 100)
        c->uvm_perf_prefetch_threshold = 51;
    if (c->uvm_perf_prefetch_min_faults < 1 || c->uvm_perf_prefetch_min_faults > 20)
        c->uvm_perf_prefetch_min_faults = 1;
    if (c->uvm_perf_fault_batch_count == 0)
        c->uvm_perf_fault_batch_count = 256;
    c->uvm_perf_fault_coalesce = c->uvm_perf_fault_coalesce != 0;
}

NV_STATUS uvmtier_init(int device, const uvmtier_config_t *cfg, uvmtier_ctx_t **out_ctx)
{
    if (!out_ctx)
        return NV_ERR_INVALID_ARGUMENT;
    *out_ctx = NULL;

    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        uvmtier_set_error("cudaGetDeviceCount", e == cudaSuccess ? cudaErrorNoDevice : e, __FILE__, __LINE__);
        return NV_ERR_GPU_NOT_FULL_POWER;   // no CPU fallback: fail loudly
    }
    if (device < 0 || device >= count)
        return NV_ERR_INVALID_ARGUMENT;
    UVMTIER_CUDA(cudaSetDevice(device));

    cudaDeviceProp prop;
    UVMTIER_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10) {
        snprintf(g_last_error, sizeof(g_last_error), "device %d is sm_%d%d; libuvmtier is built for sm_100a only",
                 device, prop.major, prop.minor);
        return NV_ERR_NOT_SUPPORTED;
    }

    uvmtier_ctx_t *ctx = (uvmtier_ctx_t *)calloc(1, sizeof(*ctx));
    if (!ctx)
        return NV_ERR_NO_MEMORY;
    ctx->device = device;
    ctx->sm_count = prop.multiProcessorCount;
    if (cfg)
        ctx->cfg = *cfg;
    else
        uvmtier_config_default(&ctx->cfg);
    validate_config(&ctx->cfg);

    e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) {
        uvmtier_set_error("cudaStreamCreateWithFlags", e, __FILE__, __LINE__);
        free(ctx);
        return uvmtier_status_from_cuda(e);
    }
    // Everything the servicing pass needs besides memory exists from here on: its two side streams, the events that
    // fork and join them, and the device address of the fixed working space of the no-alloc entry point (looking it
    // up loads the module that holds the array).  A servicing call then creates nothing, whichever comes first.
    for (int i = 0; i < 2 && e == cudaSuccess; i++)
        e = cudaStreamCreateWithFlags(&ctx->aux_stream[i], cudaStreamNonBlocking);
    for (int i = 0; i < 4 && e == cudaSuccess; i++)
        e = cudaEventCreateWithFlags(&ctx->aux_events[i], cudaEventDisableTiming);
    NV_STATUS ps = NV_OK;
    if (e != cudaSuccess) {
        uvmtier_set_error("side streams of the servicing pass", e, __FILE__, __LINE__);
        ps = uvmtier_status_from_cuda(e);
    }
    else {
        ps = uvmtier_service_prepare(ctx);
    }
    if (ps != NV_OK) {
        uvmtier_deinit(ctx);
        return ps;
    }
    *out_ctx = ctx;
    return NV_OK;
}

void uvmtier_deinit(uvmtier_ctx_t *ctx)
{
    if (!ctx)
        return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    uvmtier_fault_state_destroy(ctx);
    uvmtier_tiers_destroy(ctx);
    uvmtier_service_scratch_destroy(ctx);
    for (int i = 0; i < UVMTIER_MAX_EVENTS; i++)
        if (ctx->events[i])
            cudaEventDestroy(ctx->events[i]);
    for (int i = 0; i <= UVMTIER_MAX_COPY_CHUNKS; i++)
        if (ctx->copy_events[i])
            cudaEventDestroy(ctx->copy_events[i]);
    if (ctx->copy_stream) {
        cudaStreamSynchronize(ctx->copy_stream);
        cudaStreamDestroy(ctx->copy_stream);
    }
    for (int i = 0; i <= UVMTIER_MAX_COPY_CHUNKS; i++)
        if (ctx->d2h_events[i])
            cudaEventDestroy(ctx->d2h_events[i]);
    for (int i = 0; i < 4; i++)
        if (ctx->aux_events[i])
            cudaEventDestroy(ctx->aux_events[i]);
    for (int i = 0; i <= UVMTIER_MAX_COPY_CHUNKS; i++)
        if (ctx->k1_events[i])
            cudaEventDestroy(ctx->k1_events[i]);
    for (int i = 0; i < 3; i++)
        if (ctx->aux_stream[i]) {
            cudaStreamSynchronize(ctx->aux_stream[i]);
            cudaStreamDestroy(ctx->aux_stream[i]);
        }
    if (ctx->d2h_stream) {
        cudaStreamSynchronize(ctx->d2h_stream);
        cudaStreamDestroy(ctx->d2h_stream);
    }
    if (ctx->d_stage)
        cudaFree(ctx->d_stage);
    if (ctx->h_stage)
        cudaFreeHost(ctx->h_stage);
    cudaStreamDestroy(ctx->stream);
    free(ctx);
}

void uvmtier_get_config(const uvmtier_ctx_t *ctx, uvmtier_config_t *out) { *out = ctx->cfg; }
void *uvmtier_stream(uvmtier_ctx_t *ctx) { return (void *)ctx->stream; }

NV_STATUS uvmtier_synchronize(uvmtier_ctx_t *ctx)
{
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    UVMTIER_CUDA(cudaStreamSynchronize(ctx->stream));
    return NV_OK;
}

NV_STATUS uvmtier_mem_alloc_device(uvmtier_ctx_t *ctx, size_t bytes, void **out)
{
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    UVMTIER_CUDA(cudaMalloc(out, bytes ? bytes : 1));
    return NV_OK;
}
NV_STATUS uvmtier_mem_free_device(uvmtier_ctx_t *ctx, void *p)
{
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    UVMTIER_CUDA(cudaFree(p));
    return NV_OK;
}
NV_STATUS uvmtier_mem_alloc_host_pinned(uvmtier_ctx_t *ctx, size_t bytes, void **out)
{
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    UVMTIER_CUDA(cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocPortable | cudaHostAllocMapped));
    return NV_OK;
}
NV_STATUS uvmtier_mem_free_host_pinned(uvmtier_ctx_t *ctx, void *p)
{
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    UVMTIER_CUDA(cudaFreeHost(p));
    return NV_OK;
}
NV_STATUS uvmtier_memcpy_h2d(uvmtier_ctx_t *ctx, void *d, const void *h, size_t bytes)
{
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    UVMTIER_CUDA(cudaMemcpyAsync(d, h, bytes, cudaMemcpyHostToDevice, ctx->stream));
    UVMTIER_CUDA(cudaStreamSynchronize(ctx->stream));
    return NV_OK;
}
NV_STATUS uvmtier_memcpy_d2h(uvmtier_ctx_t *ctx, void *h, const void *d, size_t bytes)
{
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    UVMTIER_CUDA(cudaMemcpyAsync(h, d, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    UVMTIER_CUDA(cudaStreamSynchronize(ctx->stream));
    return NV_OK;
}

NV_STATUS uvmtier_event_record(uvmtier_ctx_t *ctx, uint32_t slot)
{
    if (!ctx || slot >= UVMTIER_MAX_EVENTS)
        return NV_ERR_INVALID_ARGUMENT;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    if (!ctx->events[slot])
        UVMTIER_CUDA(cudaEventCreate(&ctx->events[slot]));
    UVMTIER_CUDA(cudaEventRecord(ctx->events[slot], ctx->stream));
    return NV_OK;
}

NV_STATUS uvmtier_event_elapsed_ms(uvmtier_ctx_t *ctx, uint32_t a, uint32_t b, float *out_ms)
{
    if (!ctx || a >= UVMTIER_MAX_EVENTS || b >= UVMTIER_MAX_EVENTS || !out_ms || !ctx->events[a] || !ctx->events[b])
        return NV_ERR_INVALID_ARGUMENT;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    UVMTIER_CUDA(cudaEventSynchronize(ctx->events[b]));
    UVMTIER_CUDA(cudaEventElapsedTime(out_ms, ctx->events[a], ctx->events[b]));
    return NV_OK;
}

NV_STATUS uvmtier_memset_device(uvmtier_ctx_t *ctx, void *d, int value, size_t bytes)
{
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    UVMTIER_CUDA(cudaMemsetAsync(d, value, bytes, ctx->stream));
    return NV_OK;
}
NV_STATUS uvmtier_memcpy_h2d_async(uvmtier_ctx_t *ctx, void *d, const void *h, size_t bytes)
{
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    UVMTIER_CUDA(cudaMemcpyAsync(d, h, bytes, cudaMemcpyHostToDevice, ctx->stream));
    return NV_OK;
}
NV_STATUS uvmtier_memcpy_d2d_async(uvmtier_ctx_t *ctx, void *d_dst, const void *d_src, size_t bytes)
{
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    UVMTIER_CUDA(cudaMemcpyAsync(d_dst, d_src, bytes, cudaMemcpyDeviceToDevice, ctx->stream));
    return NV_OK;
}
NV_STATUS uvmtier_memcpy_d2h_async(uvmtier_ctx_t *ctx, void *h, const void *d, size_t bytes)
{
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    UVMTIER_CUDA(cudaMemcpyAsync(h, d, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    return NV_OK;
}

// Marsaglia KISS generator with the reference's constants (UVM/uvm_test_rng.c:37-62): every
// synthetic trace / page list of the harness is shaped from this stream so that inputs derive
// from the reference's own test RNG.
void uvmtier_test_rng_fill(uint32_t seed, size_t n, uint32_t *out)
{
    uint32_t z = 362436069u, w = 521288629u, jcong = 380116160u, jsr = seed;
    for (size_t i = 0; i < n; i++) {
        z = 36969u * (z & 65535u) + (z >> 16);
        w = 18000u * (w & 65535u) + (w >> 16);
        jcong = 69069u * jcong + 1234567u;
        jsr ^= jsr << 17;
        jsr ^= jsr >> 13;
        jsr ^= jsr << 5;
        out[i] = (((z << 16) + w) ^ jcong) + jsr;
    }
}

} // extern "C"

NV_STATUS uvmtier_stage_reserve(uvmtier_ctx_t *ctx, size_t d_bytes, size_t h_bytes)
{
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    if (d_bytes > ctx->d_stage_bytes) {
        if (ctx->d_stage) {
            UVMTIER_CUDA(cudaStreamSynchronize(ctx->stream));
            UVMTIER_CUDA(cudaFree(ctx->d_stage));
            ctx->d_stage = NULL;
            ctx->d_stage_bytes = 0;
        }
        size_t want = uvmtier_align_up(d_bytes + d_bytes / 4, 1 << 20);
        UVMTIER_CUDA(cudaMalloc(&ctx->d_stage, want));
        ctx->d_stage_bytes = want;
    }
    if (h_bytes > ctx->h_stage_bytes) {
        if (ctx->h_stage) {
            UVMTIER_CUDA(cudaStreamSynchronize(ctx->stream));
            UVMTIER_CUDA(cudaFreeHost(ctx->h_stage));
            ctx->h_stage = NULL;
            ctx->h_stage_bytes = 0;
        }
        size_t want = uvmtier_align_up(h_bytes + h_bytes / 4, 1 << 20);
        UVMTIER_CUDA(cudaHostAlloc(&ctx->h_stage, want, cudaHostAllocPortable));
        ctx->h_stage_bytes = want;
    }
    return NV_OK;
}

// The ops table (include/uvmtier.h): one slot per reference interface this library replaces.
extern "C" const uvmtier_ops_t uvmtier_gpu_ops = {
    uvmtier_fetch_preprocess_fault_batches,
    uvmtier_fetch_preprocess_fault_batches_device,
    uvmtier_service_fault_batches_device,
    uvmtier_perf_prefetch_get_hint_va_block,
    uvmtier_make_resident_device,
    uvmtier_copy_runs_device,
    uvmtier_copy_page_list,
    uvmtier_service_replayable_faults_streamed,
    uvmtier_access_counters_fetch_preprocess_device,
    uvmtier_access_counters_service_device,
    uvmtier_service_non_replayable_faults_device,
};
