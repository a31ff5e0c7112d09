// common.cuh -- shared host-side plumbing of libuvmtier (context, error mapping, launch counting).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <atomic>

#include "uvmtier.h"

#define UVMTIER_VERSION "0.1.0"
#define UVMTIER_FULL_MASK 0xffffffffu
#define UVMTIER_NUM_SMS_B200 148

// Replayable-fault state of a context (uvmtier_fault_buffer_init).
struct uvmtier_fault_state {
    uvmtier_fault_gpu_params_t gp;
    uvmtier_instance_ptr_entry_t *d_table;
    uint32_t n_rows;
    uint32_t threads;      // CTA size = next power of two >= max_batch_size (>= 32)
    uint32_t idx_bits;
    size_t smem_bytes;
};

// Scratch of the servicing pass (service.cu), grown on demand and reused.
struct uvmtier_service_scratch {
    uint32_t *d_base;
    size_t cap_words;
};

struct uvmtier_ctx {
    int device;
    int sm_count;
    cudaStream_t stream;
    uvmtier_config_t cfg;       // validated tunables
    // reusable staging buffers for the host-pointer entry points (grown on demand, never shrunk;
    // mirrors the reference's allocate-once batch context, UVM/uvm_gpu_replayable_faults.c:149-184)
    void *d_stage;
    size_t d_stage_bytes;
    void *h_stage;              // pinned
    size_t h_stage_bytes;
    cudaEvent_t events[UVMTIER_MAX_EVENTS];   // lazily created timing events
    cudaStream_t copy_stream;                 // lazily created: host->device chunks of the streamed fetch entry point
    cudaEvent_t copy_events[UVMTIER_MAX_COPY_CHUNKS + 1];
    struct uvmtier_fault_state *faults;   // lazily created by the fault entry points
    struct uvmtier_service_scratch svc;   // scratch of the servicing pass
    struct uvmtier_tiers *tiers;          // lazily created by the migration entry points
};

extern std::atomic<uint64_t> g_uvmtier_launches;
void uvmtier_set_error(const char *what, cudaError_t err, const char *file, int line);

// Maps a CUDA error to the reference's NV_STATUS convention (CUDA errors map to
// NV_ERR_GENERIC / NV_ERR_NO_MEMORY).
static inline NV_STATUS uvmtier_status_from_cuda(cudaError_t e)
{
    if (e == cudaSuccess)
        return NV_OK;
    if (e == cudaErrorMemoryAllocation)
        return NV_ERR_NO_MEMORY;
    if (e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver || e == cudaErrorInvalidDevice)
        return NV_ERR_GPU_NOT_FULL_POWER;
    return NV_ERR_GENERIC;
}

#define UVMTIER_CUDA(call)                                                     \
    do {                                                                       \
        cudaError_t _e = (call);                                               \
        if (_e != cudaSuccess) {                                               \
            uvmtier_set_error(#call, _e, __FILE__, __LINE__);                  \
            return uvmtier_status_from_cuda(_e);                               \
        }                                                                      \
    } while (0)

// Every kernel launch of the library goes through this: counts it and checks the launch error.
#define UVMTIER_LAUNCH(kernel, grid, block, smem, stream, ...)                 \
    do {                                                                       \
        kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__);            \
        g_uvmtier_launches.fetch_add(1, std::memory_order_relaxed);            \
        cudaError_t _e = cudaGetLastError();                                   \
        if (_e != cudaSuccess) {                                               \
            uvmtier_set_error(#kernel, _e, __FILE__, __LINE__);                \
            return uvmtier_status_from_cuda(_e);                               \
        }                                                                      \
    } while (0)

// Makes sure ctx->d_stage / h_stage hold at least `bytes`.
NV_STATUS uvmtier_stage_reserve(uvmtier_ctx_t *ctx, size_t d_bytes, size_t h_bytes);

static inline size_t uvmtier_align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
