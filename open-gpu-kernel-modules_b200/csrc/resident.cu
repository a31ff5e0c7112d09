// resident.cu -- C-ABI entry point of the batched uvm_va_block_make_resident (include/uvmtier.h section 1c).
// Reference: UVM/uvm_va_block.c:4877-5114 and callers uvm_migrate.c:225-289, uvm_va_block.c:12999-13130; see
// resident_kernels.cuh.
//
// Algorithmic bytes per request: 16 (request) + [64 page mask] + [64 prefetch mask] + the block's state read and
// written (7 masks x 64 B x 2, + 1 mask x 64 B x 2 per other GPU in the table) + 160 (result) + 8 per copy run.
#include "common.cuh"

#include "resident_kernels.cuh"

using namespace uvmtier_k;

// n_requests_dev != NULL: the number of requests is read on the device (n_requests is its upper bound and sizes the
// grid) -- the tiered loop decides how many va_blocks to evict without a round trip to the host.
NV_STATUS uvmtier_make_resident_counted(uvmtier_ctx_t *ctx, uint32_t n_requests, const uint32_t *n_requests_dev,
                                        const uvmtier_make_resident_req_t *d_requests, const uvmtier_page_mask_t *d_page_masks,
                                        const uvmtier_page_mask_t *d_prefetch_masks, const uvmtier_va_layout_t *lay,
                                        const uvmtier_block_state_t *state, uvmtier_make_resident_result_t *d_results,
                                        uvmtier_copy_run_t *d_runs, uint32_t run_cap, uint32_t *d_counters)
{
    if (!ctx || !lay || !state)
        return NV_ERR_INVALID_ARGUMENT;
    if (!n_requests)
        return NV_OK;
    if (!d_requests || !d_results || !d_counters || (run_cap && !d_runs) || !state->gpu_resident)
        return NV_ERR_INVALID_ARGUMENT;
    const size_t n_slots64 = (size_t)lay->n_spaces * lay->blocks_per_space;
    if (!n_slots64 || n_slots64 > kMaxSlots)
        return NV_ERR_INVALID_ARGUMENT;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));

    resident_kparams prm;
    prm.n_requests = n_requests;
    prm.n_slots = (uint32_t)n_slots64;
    prm.run_cap = run_cap;
    prm.gpu_id = lay->gpu_id;
    prm.n_requests_dev = n_requests_dev;
    if (state->n_peers > UVMTIER_MAX_PEER_GPUS)
        return NV_ERR_INVALID_ARGUMENT;
    peer_state_ptrs peers;
    memset(&peers, 0, sizeof(peers));
    peers.n_peers = state->n_peers;
    peers.direct = state->peer_direct;
    for (uint32_t k = 0; k < state->n_peers; k++) {
        if (!state->peer_resident[k] || (k && state->peer_gpu_id[k] <= state->peer_gpu_id[k - 1]) ||
            state->peer_gpu_id[k] == lay->gpu_id)
            return NV_ERR_INVALID_ARGUMENT;
        peers.resident[k] = (uint32_t *)state->peer_resident[k];
        peers.gpu_id[k] = state->peer_gpu_id[k];
    }

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

    const size_t warps_per_cta = kResThreads / 32;
    size_t grid = ((size_t)n_requests + warps_per_cta - 1) / warps_per_cta, cap = (size_t)ctx->sm_count * 16;
    if (grid > cap)
        grid = cap;
    UVMTIER_LAUNCH(k_make_resident, (int)grid, kResThreads, 0, ctx->stream, prm, d_requests, (const uint32_t *)d_page_masks,
                   (const uint32_t *)d_prefetch_masks, bs, peers, d_results, d_runs, d_counters);
    return NV_OK;
}

extern "C" NV_STATUS uvmtier_make_resident_device(uvmtier_ctx_t *ctx, uint32_t n_requests,
                                                  const uvmtier_make_resident_req_t *d_requests,
                                                  const uvmtier_page_mask_t *d_page_masks,
                                                  const uvmtier_page_mask_t *d_prefetch_masks,
                                                  const uvmtier_va_layout_t *lay, const uvmtier_block_state_t *state,
                                                  uvmtier_make_resident_result_t *d_results, uvmtier_copy_run_t *d_runs,
                                                  uint32_t run_cap, uint32_t *d_counters)
{
    return uvmtier_make_resident_counted(ctx, n_requests, NULL, d_requests, d_page_masks, d_prefetch_masks, lay, state, d_results,
                                         d_runs, run_cap, d_counters);
}
