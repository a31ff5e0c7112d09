// prefetch.cu -- K2: batched prefetch-hint kernels and their C-ABI entry points.
// Reference: UVM/uvm_perf_prefetch.c:425-445 (compute_ats), :327-413,447-502 (get_hint_va_block).
//
// Layout in HBM: masks are dense arrays of 64-byte uvm_page_mask_t, so the two va_blocks a warp
// works on are 128 contiguous bytes per load instruction (one full line, 4 sectors).  Each warp
// keeps everything in registers; no shared memory, no atomics.  Algorithmic bytes per va_block:
// 64 resident + 64 faulted [+64 thrashing] + 24 params + 4 info in, 64 mask + 4 residency + 4 info
// out = 228 B (292 B with thrashing); SURVEY 8d rounds this to 216/280.
#include "common.cuh"

#include "prefetch_kernels.cuh"

using namespace uvmtier_k;

namespace {

constexpr int kThreads = kPfThreads;

int pf_grid(const uvmtier_ctx_t *ctx, size_t n)
{
    const size_t blocks_per_cta = (kThreads / 32) * 2;
    size_t want = (n + blocks_per_cta - 1) / blocks_per_cta;
    size_t cap = (size_t)ctx->sm_count * 8;          // 8 resident CTAs of 256 threads per SM
    if (want > cap)
        want = cap;
    return (int)(want ? want : 1);
}

} // namespace

extern "C" {

NV_STATUS uvmtier_perf_prefetch_compute_ats_device(uvmtier_ctx_t *ctx, size_t n,
                                                   const uvmtier_page_mask_t *d_faulted,
                                                   const uvmtier_va_block_region_t *d_fregion,
                                                   const uvmtier_va_block_region_t *d_mregion,
                                                   const uvmtier_page_mask_t *d_residency,
                                                   uvmtier_page_mask_t *d_out)
{
    if (!ctx || (n && (!d_faulted || !d_fregion || !d_mregion || !d_out)))
        return NV_ERR_INVALID_ARGUMENT;
    if (!n)
        return NV_OK;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    pf_tunables tun = {ctx->cfg.uvm_perf_prefetch_enable, ctx->cfg.uvm_perf_prefetch_threshold,
                       ctx->cfg.uvm_perf_prefetch_min_faults};
    UVMTIER_LAUNCH(k_prefetch_compute_ats, pf_grid(ctx, n), kThreads, 0, ctx->stream, n, (const uint32_t *)d_faulted,
                   d_fregion, d_mregion, (const uint32_t *)d_residency, (uint32_t *)d_out, tun);
    return NV_OK;
}

NV_STATUS uvmtier_perf_prefetch_get_hint_va_block_device(uvmtier_ctx_t *ctx, size_t n,
                                                         const uvmtier_prefetch_block_t *d_blocks,
                                                         const uvmtier_page_mask_t *d_resident,
                                                         const uvmtier_page_mask_t *d_faulted,
                                                         const uvmtier_page_mask_t *d_thrashing,
                                                         const uvmtier_page_mask_t *d_cpu_read,
                                                         uvmtier_prefetch_info_t *d_info,
                                                         uvmtier_page_mask_t *d_out, uint32_t *d_out_residency)
{
    if (!ctx || (n && (!d_blocks || !d_resident || !d_faulted || !d_info || !d_out || !d_out_residency)))
        return NV_ERR_INVALID_ARGUMENT;
    if (!n)
        return NV_OK;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    pf_tunables tun = {ctx->cfg.uvm_perf_prefetch_enable, ctx->cfg.uvm_perf_prefetch_threshold,
                       ctx->cfg.uvm_perf_prefetch_min_faults};
    UVMTIER_LAUNCH(k_prefetch_get_hint, pf_grid(ctx, n), kThreads, 0, ctx->stream, n, d_blocks,
                   (const uint32_t *)d_resident, (const uint32_t *)d_faulted, (const uint32_t *)d_thrashing,
                   (const uint32_t *)d_cpu_read, d_info, (uint32_t *)d_out, d_out_residency, tun);
    return NV_OK;
}

// ---- host-buffer forms: stage through the context's device/pinned scratch ----

NV_STATUS uvmtier_perf_prefetch_compute_ats(uvmtier_ctx_t *ctx, size_t n, const uvmtier_page_mask_t *h_faulted,
                                            const uvmtier_va_block_region_t *h_fregion,
                                            const uvmtier_va_block_region_t *h_mregion,
                                            const uvmtier_page_mask_t *h_residency, uvmtier_page_mask_t *h_out)
{
    if (!ctx || (n && (!h_faulted || !h_fregion || !h_mregion || !h_out)))
        return NV_ERR_INVALID_ARGUMENT;
    if (!n)
        return NV_OK;
    const size_t mb = n * 64, rb = uvmtier_align_up(n * 4, 256);
    NV_STATUS st = uvmtier_stage_reserve(ctx, 3 * mb + 2 * rb, 0);
    if (st != NV_OK)
        return st;
    char *d = (char *)ctx->d_stage;
    uvmtier_page_mask_t *d_f = (uvmtier_page_mask_t *)d, *d_r = (uvmtier_page_mask_t *)(d + mb),
                        *d_o = (uvmtier_page_mask_t *)(d + 2 * mb);
    uvmtier_va_block_region_t *d_fr = (uvmtier_va_block_region_t *)(d + 3 * mb),
                              *d_mr = (uvmtier_va_block_region_t *)(d + 3 * mb + rb);
    UVMTIER_CUDA(cudaMemcpyAsync(d_f, h_faulted, mb, cudaMemcpyHostToDevice, ctx->stream));
    if (h_residency)
        UVMTIER_CUDA(cudaMemcpyAsync(d_r, h_residency, mb, cudaMemcpyHostToDevice, ctx->stream));
    UVMTIER_CUDA(cudaMemcpyAsync(d_fr, h_fregion, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    UVMTIER_CUDA(cudaMemcpyAsync(d_mr, h_mregion, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    st = uvmtier_perf_prefetch_compute_ats_device(ctx, n, d_f, d_fr, d_mr, h_residency ? d_r : NULL, d_o);
    if (st != NV_OK)
        return st;
    UVMTIER_CUDA(cudaMemcpyAsync(h_out, d_o, mb, cudaMemcpyDeviceToHost, ctx->stream));
    UVMTIER_CUDA(cudaStreamSynchronize(ctx->stream));
    return NV_OK;
}

NV_STATUS uvmtier_perf_prefetch_get_hint_va_block(uvmtier_ctx_t *ctx, size_t n, const uvmtier_prefetch_block_t *h_blocks,
                                                  const uvmtier_page_mask_t *h_resident,
                                                  const uvmtier_page_mask_t *h_faulted,
                                                  const uvmtier_page_mask_t *h_thrashing,
                                                  const uvmtier_page_mask_t *h_cpu_read, uvmtier_prefetch_info_t *h_info,
                                                  uvmtier_page_mask_t *h_out, uint32_t *h_out_residency)
{
    if (!ctx || (n && (!h_blocks || !h_resident || !h_faulted || !h_info || !h_out || !h_out_residency)))
        return NV_ERR_INVALID_ARGUMENT;
    if (!n)
        return NV_OK;
    const size_t mb = n * 64, bb = uvmtier_align_up(n * sizeof(uvmtier_prefetch_block_t), 256),
                 ib = uvmtier_align_up(n * 4, 256);
    NV_STATUS st = uvmtier_stage_reserve(ctx, 5 * mb + bb + 2 * ib, 0);
    if (st != NV_OK)
        return st;
    char *d = (char *)ctx->d_stage;
    uvmtier_page_mask_t *d_res = (uvmtier_page_mask_t *)d, *d_f = (uvmtier_page_mask_t *)(d + mb),
                        *d_thr = (uvmtier_page_mask_t *)(d + 2 * mb), *d_map = (uvmtier_page_mask_t *)(d + 3 * mb),
                        *d_o = (uvmtier_page_mask_t *)(d + 4 * mb);
    uvmtier_prefetch_block_t *d_b = (uvmtier_prefetch_block_t *)(d + 5 * mb);
    uvmtier_prefetch_info_t *d_i = (uvmtier_prefetch_info_t *)(d + 5 * mb + bb);
    uint32_t *d_hint = (uint32_t *)(d + 5 * mb + bb + ib);
    UVMTIER_CUDA(cudaMemcpyAsync(d_res, h_resident, mb, cudaMemcpyHostToDevice, ctx->stream));
    UVMTIER_CUDA(cudaMemcpyAsync(d_f, h_faulted, mb, cudaMemcpyHostToDevice, ctx->stream));
    if (h_thrashing)
        UVMTIER_CUDA(cudaMemcpyAsync(d_thr, h_thrashing, mb, cudaMemcpyHostToDevice, ctx->stream));
    if (h_cpu_read)
        UVMTIER_CUDA(cudaMemcpyAsync(d_map, h_cpu_read, mb, cudaMemcpyHostToDevice, ctx->stream));
    UVMTIER_CUDA(cudaMemcpyAsync(d_b, h_blocks, n * sizeof(*h_blocks), cudaMemcpyHostToDevice, ctx->stream));
    UVMTIER_CUDA(cudaMemcpyAsync(d_i, h_info, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    st = uvmtier_perf_prefetch_get_hint_va_block_device(ctx, n, d_b, d_res, d_f, h_thrashing ? d_thr : NULL,
                                                        h_cpu_read ? d_map : NULL, d_i, d_o, d_hint);
    if (st != NV_OK)
        return st;
    UVMTIER_CUDA(cudaMemcpyAsync(h_out, d_o, mb, cudaMemcpyDeviceToHost, ctx->stream));
    UVMTIER_CUDA(cudaMemcpyAsync(h_out_residency, d_hint, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    UVMTIER_CUDA(cudaMemcpyAsync(h_info, d_i, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    UVMTIER_CUDA(cudaStreamSynchronize(ctx->stream));
    return NV_OK;
}

} // extern "C"
