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
    uint32_t ac_granularity_2m, ac_cpu_chunk_2m;   // access-counter configuration (uvmtier_access_counters_init)
    uint32_t generation;   // bumped by every uvmtier_fault_buffer_init (the table's contents may have changed)
    uint16_t *d_utlb_pending;   // optional output, uvmtier_fault_buffer_set_utlb_pending_output
    // batches of more than 1024 entries (k_fault_fetch_preprocess_big): entries per batch rounded up to a power of two
    // (0 = the one-thread-per-entry kernel is used) and the kernel's working space
    uint32_t big_T;
    unsigned char *d_big;
    size_t big_bytes;
};

// Scratch of the servicing pass (service.cu), grown on demand and reused.
struct uvmtier_service_scratch {
    uint32_t *d_base;
    size_t cap_words;
    // CUDA graph of the last uvmtier_service_replayable_faults_streamed call whose arguments repeated (service.cu)
    cudaGraphExec_t graph_exec;
    uint64_t graph_launches;                  // kernel launches inside the graph
    unsigned char graph_key[384];             // the arguments the graph (or the last eager call) was built from
    int graph_key_seen;                       // consecutive calls with graph_key
};

struct uvmtier_ctx;
// Chunked host->device streaming of a pinned ring (faults.cu), shared by the streamed entry points.
struct uvmtier_chunk_plan { uint32_t n, B, n_batches, per, n_chunks; };
NV_STATUS uvmtier_streamed_begin(struct uvmtier_ctx *ctx, uint32_t ring_entries, uint32_t get, uint32_t put,
                                 uint32_t chunk_batches, uvmtier_chunk_plan *plan);
// Copies chunk c into d_ring and makes the context's stream wait for it; [*first, *last) is its ring range.
NV_STATUS uvmtier_streamed_copy_chunk(struct uvmtier_ctx *ctx, const uvmtier_chunk_plan &plan, uint32_t c,
                                      const uvmtier_fault_packet_t *h_ring, uvmtier_fault_packet_t *d_ring,
                                      uint32_t ring_entries, uint32_t get, uint32_t *e0, uint32_t *e1, uint32_t *first,
                                      uint32_t *last, cudaStream_t waiter = NULL);
// uvmtier_fetch_preprocess_fault_batches_device on a stream of the caller's choice
NV_STATUS uvmtier_fetch_preprocess_on_stream(struct uvmtier_ctx *ctx, cudaStream_t stream, uvmtier_fault_packet_t *d_ring,
                                             uint32_t ring_entries, uint32_t get, uint32_t put, int clear_valid,
                                             uint32_t *d_disposition, uvmtier_ordered_fault_t *d_ordered,
                                             uvmtier_fault_group_t *d_groups, uvmtier_fault_batch_t *d_batches,
                                             uvmtier_fault_buffer_entry_t *d_entries, uint32_t *out_get,
                                             uint32_t utlb_batch_base = 0);

// uvmtier_service_fault_batches_device for one chunk of a longer trace (service.cu)
NV_STATUS uvmtier_service_pass_chunk(struct uvmtier_ctx *ctx, uint32_t n_batches, const uvmtier_ordered_fault_t *d_ordered,
                                     const uvmtier_fault_group_t *d_groups, uvmtier_fault_batch_t *d_batches,
                                     const uvmtier_va_layout_t *lay, const uvmtier_block_state_t *state, uint16_t *d_entry_flags,
                                     uvmtier_service_group_t *d_service, uvmtier_copy_run_t *d_runs, uint32_t run_cap,
                                     uint32_t *d_counters, uint32_t batch_base, bool non_replayable = false);

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
    cudaStream_t aux_stream[3];               // lazily created: side branches of the servicing pass (service.cu),
    cudaEvent_t aux_events[4];                //   [2] = the preprocess kernels of the streamed service entry point
    cudaEvent_t k1_events[UVMTIER_MAX_COPY_CHUNKS + 1];
    cudaStream_t d2h_stream;                  // lazily created: chunked copy-back of the streamed service entry point
    cudaEvent_t d2h_events[UVMTIER_MAX_COPY_CHUNKS + 1];
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
// internal, for the tiered loop (pmm.cu): calls whose sizes are decided on the device
NV_STATUS uvmtier_make_resident_counted(uvmtier_ctx_t *ctx, uint32_t n_requests, const uint32_t *n_requests_dev,
                                        const uvmtier_make_resident_req_t *d_requests, const uvmtier_page_mask_t *d_page_masks,
                                        const uvmtier_page_mask_t *d_prefetch_masks, const uvmtier_va_layout_t *lay,
                                        const uvmtier_block_state_t *state, uvmtier_make_resident_result_t *d_results,
                                        uvmtier_copy_run_t *d_runs, uint32_t run_cap, uint32_t *d_counters);
NV_STATUS uvmtier_copy_runs_enqueue(uvmtier_ctx_t *ctx, const uvmtier_copy_run_t *d_runs, const uint32_t *d_n_runs, uint32_t run_cap,
                                    uint64_t max_pages, uint32_t pair_mask);
void uvmtier_tiers_set_chunk_table(uvmtier_ctx_t *ctx, const uint32_t *chunk_of_slot);

static inline size_t uvmtier_align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
