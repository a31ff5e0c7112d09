// pmm.cu -- host side of section (1d) of include/uvmtier.h: the eviction order of HBM root chunks
// (UVM/uvm_pmm_gpu.c:1395-1500) on the flat state table.  Kernels: pmm_kernels.cuh.
#include "common.cuh"

#include "pmm_kernels.cuh"

using namespace uvmtier_k;

// working space: [0..3] selection result, [4..7] counters, then a bitmap of n_slots bits
static size_t pmm_scratch_words(uint32_t n_slots) { return 8 + ((size_t)n_slots + 31) / 32; }

static bool pmm_ok(const uvmtier_ctx_t *ctx, const uvmtier_pmm_t *pmm)
{
    return ctx && pmm && pmm->d_stamp && pmm->d_scratch && pmm->n_slots;
}

extern "C" {

NV_STATUS uvmtier_pmm_create(uvmtier_ctx_t *ctx, const uvmtier_va_layout_t *lay, uint32_t capacity_chunks, uvmtier_pmm_t *out)
{
    if (!ctx || !lay || !out || !lay->n_spaces || !lay->blocks_per_space)
        return NV_ERR_INVALID_ARGUMENT;
    const size_t n = (size_t)lay->n_spaces * lay->blocks_per_space;
    if (n > kMaxSlots)
        return NV_ERR_INVALID_ARGUMENT;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    memset(out, 0, sizeof(*out));
    UVMTIER_CUDA(cudaMalloc(&out->d_stamp, n * sizeof(uint64_t)));
    UVMTIER_CUDA(cudaMalloc(&out->d_scratch, pmm_scratch_words((uint32_t)n) * sizeof(uint32_t)));
    UVMTIER_CUDA(cudaMalloc(&out->d_dram_stamp, n * sizeof(uint64_t)));
    UVMTIER_CUDA(cudaMemsetAsync(out->d_stamp, 0, n * sizeof(uint64_t), ctx->stream));
    UVMTIER_CUDA(cudaMemsetAsync(out->d_dram_stamp, 0, n * sizeof(uint64_t), ctx->stream));
    UVMTIER_CUDA(cudaStreamSynchronize(ctx->stream));
    out->n_slots = (uint32_t)n;
    out->capacity_chunks = capacity_chunks;
    out->clock = 0;
    return NV_OK;
}

NV_STATUS uvmtier_pmm_destroy(uvmtier_ctx_t *ctx, uvmtier_pmm_t *pmm)
{
    if (!ctx || !pmm)
        return NV_ERR_INVALID_ARGUMENT;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    UVMTIER_CUDA(cudaStreamSynchronize(ctx->stream));
    if (pmm->d_chunk_of_slot)
        uvmtier_tiers_set_chunk_table(ctx, NULL);
    if (pmm->d_stamp) cudaFree(pmm->d_stamp);
    if (pmm->d_scratch) cudaFree(pmm->d_scratch);
    if (pmm->d_chunk_of_slot) cudaFree(pmm->d_chunk_of_slot);
    if (pmm->d_free_bits) cudaFree(pmm->d_free_bits);
    if (pmm->d_dram_stamp) cudaFree(pmm->d_dram_stamp);
    memset(pmm, 0, sizeof(*pmm));
    return NV_OK;
}

// d_free_bits: one bit per chunk, then one word that counts allocations that found nothing free
static uint32_t pmm_free_words(const uvmtier_pmm_t *pmm) { return (pmm->capacity_chunks + 31) / 32; }

static pmm_chunk_table pmm_chunks(const uvmtier_pmm_t *pmm)
{
    pmm_chunk_table t;
    t.chunk_of_slot = pmm->d_chunk_of_slot;
    t.free_bits = pmm->d_free_bits;
    t.n_words = pmm->d_chunk_of_slot ? pmm_free_words(pmm) : 0;
    t.errors = pmm->d_free_bits ? pmm->d_free_bits + pmm_free_words(pmm) : NULL;
    return t;
}

NV_STATUS uvmtier_pmm_enable_chunk_table(uvmtier_ctx_t *ctx, uvmtier_pmm_t *pmm)
{
    if (!pmm_ok(ctx, pmm) || !pmm->capacity_chunks)
        return NV_ERR_INVALID_ARGUMENT;
    if (pmm->d_chunk_of_slot)
        return NV_OK;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    // only on a table without GPU pages: every stamp 0
    UVMTIER_CUDA(cudaMemsetAsync(pmm->d_scratch, 0, 4 * sizeof(uint32_t), ctx->stream));
    UVMTIER_LAUNCH(k_pmm_select, 1, kPmmSelectThreads, 0, ctx->stream, (const unsigned long long *)pmm->d_stamp, pmm->n_slots, 1u,
                   pmm->d_scratch);
    uint32_t sel[3];
    UVMTIER_CUDA(cudaMemcpyAsync(sel, pmm->d_scratch, sizeof(sel), cudaMemcpyDeviceToHost, ctx->stream));
    UVMTIER_CUDA(cudaStreamSynchronize(ctx->stream));
    if (sel[2] != 0)
        return NV_ERR_INVALID_STATE;
    const uint32_t words = pmm_free_words(pmm);
    UVMTIER_CUDA(cudaMalloc(&pmm->d_chunk_of_slot, (size_t)pmm->n_slots * sizeof(uint32_t)));
    UVMTIER_CUDA(cudaMalloc(&pmm->d_free_bits, ((size_t)words + 1) * sizeof(uint32_t)));
    UVMTIER_CUDA(cudaMemsetAsync(pmm->d_chunk_of_slot, 0xff, (size_t)pmm->n_slots * sizeof(uint32_t), ctx->stream));
    uint32_t *h = (uint32_t *)malloc(((size_t)words + 1) * sizeof(uint32_t));
    if (!h)
        return NV_ERR_NO_MEMORY;
    for (uint32_t w = 0; w < words; w++) {
        const uint32_t left = pmm->capacity_chunks - w * 32u;
        h[w] = left >= 32u ? 0xffffffffu : ((1u << left) - 1u);
    }
    h[words] = 0;
    const cudaError_t e = cudaMemcpy(pmm->d_free_bits, h, ((size_t)words + 1) * sizeof(uint32_t), cudaMemcpyHostToDevice);
    free(h);
    UVMTIER_CUDA(e);
    UVMTIER_CUDA(cudaStreamSynchronize(ctx->stream));
    uvmtier_tiers_set_chunk_table(ctx, pmm->d_chunk_of_slot);
    return NV_OK;
}

NV_STATUS uvmtier_pmm_note_service_device(uvmtier_ctx_t *ctx, uvmtier_pmm_t *pmm, uint32_t n_batches,
                                          const uvmtier_fault_batch_t *d_batches, const uvmtier_service_group_t *d_service)
{
    if (!pmm_ok(ctx, pmm))
        return NV_ERR_INVALID_ARGUMENT;
    if (!ctx->faults)
        return NV_ERR_INVALID_STATE;
    if (!n_batches)
        return NV_OK;
    if (!d_batches || !d_service)
        return NV_ERR_INVALID_ARGUMENT;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    const uint32_t B = ctx->faults->gp.max_batch_size;
    const uint64_t total = (uint64_t)n_batches * B;
    uint64_t grid = (total + kPmmThreads - 1) / kPmmThreads;
    if (grid > (uint64_t)ctx->sm_count * 8)
        grid = (uint64_t)ctx->sm_count * 8;
    UVMTIER_LAUNCH(k_pmm_note_service, (int)grid, kPmmThreads, 0, ctx->stream, n_batches, B, d_batches, d_service,
                   (unsigned long long *)pmm->d_stamp, pmm->n_slots, (unsigned long long)pmm->clock, (unsigned long long *)NULL,
                   pmm_chunks(pmm));
    pmm->clock += total;
    return NV_OK;
}

NV_STATUS uvmtier_pmm_note_make_resident_device(uvmtier_ctx_t *ctx, uvmtier_pmm_t *pmm, uint32_t n_requests,
                                                const uvmtier_make_resident_req_t *d_requests,
                                                const uvmtier_make_resident_result_t *d_results,
                                                const uvmtier_block_state_t *state)
{
    if (!pmm_ok(ctx, pmm) || !state || !state->gpu_resident)
        return NV_ERR_INVALID_ARGUMENT;
    if (!n_requests)
        return NV_OK;
    if (!d_requests || !d_results)
        return NV_ERR_INVALID_ARGUMENT;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    const uint32_t grid = (n_requests + kPmmThreads - 1) / kPmmThreads;
    // with a chunk table, a call that both brings va_blocks to the GPU and evicts others hands out chunks while it takes
    // them back: correct, but the newcomers may not see the chunks freed by the same call
    UVMTIER_LAUNCH(k_pmm_note_make_resident, (int)grid, kPmmThreads, 0, ctx->stream, n_requests, d_requests, d_results,
                   (const uint32_t *)state->gpu_resident, (unsigned long long *)pmm->d_stamp, pmm->n_slots,
                   (unsigned long long)pmm->clock, (unsigned long long *)NULL, (const uint32_t *)NULL, pmm_chunks(pmm),
                   (unsigned long long *)pmm->d_dram_stamp, (unsigned long long)pmm->round, (const uint32_t *)state->cpu_resident[0]);
    pmm->clock += n_requests;
    return NV_OK;
}

NV_STATUS uvmtier_pmm_count_needed_device(uvmtier_ctx_t *ctx, uvmtier_pmm_t *pmm, uint32_t n_batches,
                                          const uvmtier_fault_group_t *d_groups, const uvmtier_fault_batch_t *d_batches,
                                          const uvmtier_va_layout_t *lay, uint32_t *d_out)
{
    if (!pmm_ok(ctx, pmm) || !lay || !d_out || (size_t)lay->n_spaces * lay->blocks_per_space != pmm->n_slots)
        return NV_ERR_INVALID_ARGUMENT;
    if (!ctx->faults)
        return NV_ERR_INVALID_STATE;
    if (n_batches && (!d_groups || !d_batches))
        return NV_ERR_INVALID_ARGUMENT;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    service_kparams prm;
    memset(&prm, 0, sizeof(prm));
    prm.va_base = lay->va_base;
    prm.B = ctx->faults->gp.max_batch_size;
    prm.n_batches = n_batches;
    prm.blocks_per_space = lay->blocks_per_space;
    prm.n_spaces = lay->n_spaces;
    prm.gpu_id = lay->gpu_id;
    prm.n_slots = pmm->n_slots;
    prm.B_shift = svc_log2_or_none(prm.B);
    uint32_t *seen = pmm->d_scratch + 8;
    UVMTIER_CUDA(cudaMemsetAsync(seen, 0, ((size_t)pmm->n_slots + 31) / 32 * sizeof(uint32_t), ctx->stream));
    UVMTIER_CUDA(cudaMemsetAsync(d_out, 0, 2 * sizeof(uint32_t), ctx->stream));
    uint64_t work = (uint64_t)n_batches * prm.B;
    if (work < pmm->n_slots)
        work = pmm->n_slots;
    uint64_t grid = (work + kPmmThreads - 1) / kPmmThreads;
    if (grid > (uint64_t)ctx->sm_count * 8)
        grid = (uint64_t)ctx->sm_count * 8;
    UVMTIER_LAUNCH(k_pmm_count_needed, (int)grid, kPmmThreads, 0, ctx->stream, prm, d_groups, d_batches, ctx->faults->d_table,
                   (const unsigned long long *)pmm->d_stamp, seen, d_out);
    return NV_OK;
}

NV_STATUS uvmtier_pmm_pick_chunks_to_evict_device(uvmtier_ctx_t *ctx, uvmtier_pmm_t *pmm, uint32_t k, uint32_t dest_tier,
                                                  uvmtier_make_resident_req_t *d_requests, uint32_t *d_count)
{
    if (!pmm_ok(ctx, pmm) || !d_count || (k && !d_requests) || (dest_tier != UVMTIER_TIER_DRAM && dest_tier != UVMTIER_TIER_CXL))
        return NV_ERR_INVALID_ARGUMENT;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    UVMTIER_CUDA(cudaMemsetAsync(d_count, 0, sizeof(uint32_t), ctx->stream));
    if (!k)
        return NV_OK;
    uint32_t *sel = pmm->d_scratch;
    UVMTIER_LAUNCH(k_pmm_select, 1, kPmmSelectThreads, 0, ctx->stream, (const unsigned long long *)pmm->d_stamp, pmm->n_slots, k, sel);
    uint32_t grid = (pmm->n_slots + kPmmThreads - 1) / kPmmThreads;
    if (grid > (uint32_t)ctx->sm_count * 8)
        grid = (uint32_t)ctx->sm_count * 8;
    UVMTIER_LAUNCH(k_pmm_emit_evictions, (int)grid, kPmmThreads, 0, ctx->stream, (const unsigned long long *)pmm->d_stamp,
                   pmm->n_slots, sel, dest_tier, d_requests, d_count);
    return NV_OK;
}

NV_STATUS uvmtier_block_state_count_pages_device(uvmtier_ctx_t *ctx, const uvmtier_va_layout_t *lay,
                                                 const uvmtier_block_state_t *state, uint32_t *d_out)
{
    if (!ctx || !lay || !state || !state->gpu_resident || !d_out || !lay->n_spaces || !lay->blocks_per_space)
        return NV_ERR_INVALID_ARGUMENT;
    const size_t n = (size_t)lay->n_spaces * lay->blocks_per_space;
    if (n > kMaxSlots)
        return NV_ERR_INVALID_ARGUMENT;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    UVMTIER_CUDA(cudaMemsetAsync(d_out, 0, 3 * sizeof(uint32_t), ctx->stream));
    uint64_t grid = (n * 16 + kPmmThreads - 1) / kPmmThreads;
    if (grid > (uint64_t)ctx->sm_count * 8)
        grid = (uint64_t)ctx->sm_count * 8;
    UVMTIER_LAUNCH(k_state_count_pages, (int)grid, kPmmThreads, 0, ctx->stream, (uint32_t)n, (const uint32_t *)state->gpu_resident,
                   (const uint32_t *)state->cpu_resident[0], (const uint32_t *)state->cpu_resident[1], d_out);
    return NV_OK;
}

NV_STATUS uvmtier_service_fault_batches_tiered_device(uvmtier_ctx_t *ctx, uvmtier_pmm_t *pmm, uint32_t n_batches,
                                                      const uvmtier_ordered_fault_t *d_ordered,
                                                      const uvmtier_fault_group_t *d_groups, uvmtier_fault_batch_t *d_batches,
                                                      const uvmtier_va_layout_t *lay, const uvmtier_block_state_t *state,
                                                      uint16_t *d_entry_flags, uvmtier_service_group_t *d_service,
                                                      uint32_t chunk_batches, uint64_t dram_capacity_pages, int flags,
                                                      uvmtier_tiering_stats_t *out_stats)
{
    if (out_stats)
        memset(out_stats, 0, sizeof(*out_stats));
    if (!pmm_ok(ctx, pmm) || !lay || !state || !chunk_batches || (size_t)lay->n_spaces * lay->blocks_per_space != pmm->n_slots)
        return NV_ERR_INVALID_ARGUMENT;
    const bool execute_copies = (flags & UVMTIER_TIERED_EXECUTE_COPIES) != 0, demote = (flags & UVMTIER_TIERED_DEMOTE) != 0;
    if (!ctx->faults || (execute_copies && !ctx->tiers))
        return NV_ERR_INVALID_STATE;
    if (state->n_peers)
        return NV_ERR_NOT_SUPPORTED;
    if (!n_batches)
        return NV_OK;
    if (!d_ordered || !d_groups || !d_batches || !d_entry_flags || !d_service || n_batches > 65536u)   // 16-bit batch numbers
        return NV_ERR_INVALID_ARGUMENT;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    const uint32_t B = ctx->faults->gp.max_batch_size;
    const uint64_t chunk_entries = (uint64_t)chunk_batches * B;
    const uint32_t n_slots = pmm->n_slots;
    if (pmm->d_chunk_of_slot && chunk_entries > pmm->capacity_chunks) {
        uvmtier_set_error("tiered servicing with a chunk table: a round (chunk_batches * max_batch_size entries) may fault on more "
                          "va_blocks than HBM has chunks", cudaErrorInvalidValue, __FILE__, __LINE__);
        return NV_ERR_INVALID_ARGUMENT;
    }
    // a round evicts at most every va_block that holds a chunk, a va_block's GPU pages are at most 256 runs, and the DRAM
    // pages of a demoted va_block likewise
    const uint64_t evict_run_cap64 = (uint64_t)n_slots * 256u;
    const uint32_t evict_run_cap = (uint32_t)(evict_run_cap64 < 0x7fffffffull ? evict_run_cap64 : 0x7fffffffull);
    // copy runs of one servicing pass, worst case: every page a va_block receives is a run of its own, a va_block
    // receives at most its 512 pages, and a pass touches at most min(entries, slots) va_blocks -- so the list is never
    // cut and the copies never run on a truncated list
    const uint64_t svc_blocks = chunk_entries < n_slots ? chunk_entries : n_slots;
    const uint64_t svc_cap64 = svc_blocks * UVMTIER_PAGES_PER_VA_BLOCK;
    const uint32_t svc_run_cap = (uint32_t)(svc_cap64 < 0x7fffffffull ? svc_cap64 : 0x7fffffffull);
    const uint64_t all_pages = (uint64_t)n_slots * UVMTIER_PAGES_PER_VA_BLOCK;
    // working space: plan + counters, the selections' per-slot arrays, requests, results, the two run arrays
    size_t off = 0;
    auto take = [&](size_t bytes) { const size_t at = off; off = uvmtier_align_up(off + bytes, 256); return at; };
    const size_t o_plan = take(kPlanWords * sizeof(uint32_t)), o_acc = take(4 * sizeof(unsigned long long)),
                 o_touched = take(((size_t)n_slots + 31) / 32 * sizeof(uint32_t)), o_ew = take((size_t)n_slots * 4),
                 o_ea = take((size_t)n_slots * 4), o_dk = take((size_t)n_slots * 8), o_dw = take((size_t)n_slots * 4),
                 o_req = take((size_t)n_slots * sizeof(uvmtier_make_resident_req_t)),
                 o_res = take((size_t)n_slots * sizeof(uvmtier_make_resident_result_t)),
                 o_eruns = take((size_t)evict_run_cap * sizeof(uvmtier_copy_run_t)),
                 o_sruns = take((size_t)svc_run_cap * sizeof(uvmtier_copy_run_t));
    NV_STATUS st = uvmtier_stage_reserve(ctx, off, 512);
    if (st != NV_OK)
        return st;
    char *d = (char *)ctx->d_stage;
    uint32_t *plan = (uint32_t *)(d + o_plan);                            // layout: k_pmm_plan; [40] requests [42] eviction runs
    unsigned long long *d_acc = (unsigned long long *)(d + o_acc);        //         [44] demotion runs [46] service runs
    uint32_t *touched = (uint32_t *)(d + o_touched), *e_weight = (uint32_t *)(d + o_ew), *e_aux = (uint32_t *)(d + o_ea),
             *d_weight = (uint32_t *)(d + o_dw);
    unsigned long long *d_key = (unsigned long long *)(d + o_dk);
    uvmtier_make_resident_req_t *d_req = (uvmtier_make_resident_req_t *)(d + o_req);
    uvmtier_make_resident_result_t *d_res = (uvmtier_make_resident_result_t *)(d + o_res);
    uvmtier_copy_run_t *d_eruns = (uvmtier_copy_run_t *)(d + o_eruns), *d_sruns = (uvmtier_copy_run_t *)(d + o_sruns);
    uint32_t *h_cnt = (uint32_t *)ctx->h_stage;
    UVMTIER_CUDA(cudaMemsetAsync(d, 0, o_touched, ctx->stream));
    UVMTIER_CUDA(cudaMemsetAsync(d_entry_flags, 0, (size_t)n_batches * B * sizeof(uint16_t), ctx->stream));

    service_kparams prm;
    memset(&prm, 0, sizeof(prm));
    prm.va_base = lay->va_base;
    prm.B = B;
    prm.blocks_per_space = lay->blocks_per_space;
    prm.n_spaces = lay->n_spaces;
    prm.gpu_id = lay->gpu_id;
    prm.n_slots = n_slots;
    prm.B_shift = svc_log2_or_none(B);
    const pmm_chunk_table chunks = pmm_chunks(pmm);
    const int slot_grid = (int)((((uint64_t)n_slots + kPmmThreads - 1) / kPmmThreads) < (uint64_t)ctx->sm_count * 8
                                    ? ((uint64_t)n_slots + kPmmThreads - 1) / kPmmThreads : (uint64_t)ctx->sm_count * 8);
    const uint32_t pair_hbm_dram = UVMTIER_RUN_SRC_HBM * 3 + UVMTIER_RUN_DST_DRAM, pair_hbm_cxl = UVMTIER_RUN_SRC_HBM * 3 + UVMTIER_RUN_DST_CXL,
                   pair_dram_cxl = UVMTIER_RUN_SRC_DRAM * 3 + UVMTIER_RUN_DST_CXL, pair_dram_hbm = UVMTIER_RUN_SRC_DRAM * 3 + UVMTIER_RUN_DST_HBM,
                   pair_cxl_hbm = UVMTIER_RUN_SRC_CXL * 3 + UVMTIER_RUN_DST_HBM, pair_zero_hbm = UVMTIER_RUN_SRC_ZERO * 3 + UVMTIER_RUN_DST_HBM;
    uint32_t passes = 0;
    for (uint32_t b0 = 0; b0 < n_batches; b0 += chunk_batches) {
        const uint32_t nb = n_batches - b0 < chunk_batches ? n_batches - b0 : chunk_batches;
        const size_t e0 = (size_t)b0 * B;
        const uint64_t total = (uint64_t)nb * B;
        uint64_t egrid = (total + kPmmThreads - 1) / kPmmThreads;
        if (egrid > (uint64_t)ctx->sm_count * 8)
            egrid = (uint64_t)ctx->sm_count * 8;
        pmm->round++;
        // ---- what the round needs, what is there ----
        UVMTIER_CUDA(cudaMemsetAsync(touched, 0, ((size_t)n_slots + 31) / 32 * sizeof(uint32_t), ctx->stream));
        UVMTIER_CUDA(cudaMemsetAsync(plan, 0, 32 * sizeof(uint32_t), ctx->stream));          // totals, targets, selections
        UVMTIER_CUDA(cudaMemsetAsync(plan + 40, 0, 8 * sizeof(uint32_t), ctx->stream));      // request / run counters
        prm.n_batches = nb;
        UVMTIER_LAUNCH(k_pmm_mark_touched, (int)egrid, kPmmThreads, 0, ctx->stream, prm, d_groups + e0,
                       (const uvmtier_fault_batch_t *)(d_batches + b0), (const uvmtier_instance_ptr_entry_t *)ctx->faults->d_table, touched);
        UVMTIER_LAUNCH(k_pmm_slot_keys, slot_grid, kPmmThreads, 0, ctx->stream, n_slots, (const unsigned long long *)pmm->d_stamp,
                       (const unsigned long long *)pmm->d_dram_stamp, (const uint32_t *)touched, (const uint32_t *)state->gpu_resident,
                       (const uint32_t *)state->cpu_resident[0], (const uint32_t *)state->cpu_resident[1], e_weight, e_aux, d_key,
                       d_weight, plan);
        // ---- who leaves HBM, who leaves DRAM, where the victims go ----
        UVMTIER_LAUNCH(k_pmm_plan, 1, 32, 0, ctx->stream, plan, 0, pmm->capacity_chunks, (unsigned long long)dram_capacity_pages, (int)demote);
        UVMTIER_LAUNCH(k_pmm_select_weighted, 1, kPmmSelectThreads, 0, ctx->stream, (const unsigned long long *)pmm->d_stamp,
                       (const uint32_t *)e_weight, (const uint32_t *)e_aux, n_slots, (const uint32_t *)(plan + 8), plan + 16);
        UVMTIER_LAUNCH(k_pmm_plan, 1, 32, 0, ctx->stream, plan, 1, pmm->capacity_chunks, (unsigned long long)dram_capacity_pages, (int)demote);
        if (demote)
            UVMTIER_LAUNCH(k_pmm_select_weighted, 1, kPmmSelectThreads, 0, ctx->stream, (const unsigned long long *)d_key,
                           (const uint32_t *)d_weight, (const uint32_t *)NULL, n_slots, (const uint32_t *)(plan + 9), plan + 24);
        UVMTIER_LAUNCH(k_pmm_plan, 1, 32, 0, ctx->stream, plan, 2, pmm->capacity_chunks, (unsigned long long)dram_capacity_pages, (int)demote);
        if (demote) {
            // ---- DRAM -> CXL for the coldest DRAM va_blocks (page mask = the block's row of cpu_resident[0]) ----
            UVMTIER_LAUNCH(k_pmm_emit_selected, slot_grid, kPmmThreads, 0, ctx->stream, (const unsigned long long *)d_key, n_slots,
                           (const uint32_t *)(plan + 24), 0, (const uint32_t *)(plan + 10), d_req, plan + 40);
            st = uvmtier_make_resident_counted(ctx, n_slots, plan + 40, d_req, state->cpu_resident[0], NULL, lay, state, d_res, d_eruns,
                                               evict_run_cap, plan + 44);
            if (st != NV_OK)
                return st;
            UVMTIER_LAUNCH(k_pmm_note_make_resident, slot_grid, kPmmThreads, 0, ctx->stream, n_slots,
                           (const uvmtier_make_resident_req_t *)d_req, (const uvmtier_make_resident_result_t *)d_res,
                           (const uint32_t *)state->gpu_resident, (unsigned long long *)pmm->d_stamp, n_slots,
                           (unsigned long long)pmm->clock, d_acc + 2, (const uint32_t *)(plan + 40), chunks,
                           (unsigned long long *)pmm->d_dram_stamp, (unsigned long long)pmm->round,
                           (const uint32_t *)state->cpu_resident[0]);
            if (execute_copies) {
                st = uvmtier_copy_runs_enqueue(ctx, d_eruns, plan + 44, evict_run_cap, all_pages, 1u << pair_dram_cxl);
                if (st != NV_OK)
                    return st;
            }
        }
        // ---- evictions: make_resident to the chosen tier, the copies out, then the list moves and the chunks given back ----
        UVMTIER_LAUNCH(k_pmm_emit_selected, slot_grid, kPmmThreads, 0, ctx->stream, (const unsigned long long *)pmm->d_stamp, n_slots,
                       (const uint32_t *)(plan + 16), 1, (const uint32_t *)(plan + 10), d_req, plan + 41);
        st = uvmtier_make_resident_counted(ctx, n_slots, plan + 41, d_req, NULL, NULL, lay, state, d_res, d_eruns, evict_run_cap, plan + 42);
        if (st != NV_OK)
            return st;
        if (execute_copies) {
            st = uvmtier_copy_runs_enqueue(ctx, d_eruns, plan + 42, evict_run_cap, all_pages, (1u << pair_hbm_dram) | (1u << pair_hbm_cxl));
            if (st != NV_OK)
                return st;
        }
        UVMTIER_LAUNCH(k_pmm_note_make_resident, slot_grid, kPmmThreads, 0, ctx->stream, n_slots,
                       (const uvmtier_make_resident_req_t *)d_req, (const uvmtier_make_resident_result_t *)d_res,
                       (const uint32_t *)state->gpu_resident, (unsigned long long *)pmm->d_stamp, n_slots,
                       (unsigned long long)pmm->clock, d_acc, (const uint32_t *)(plan + 41), chunks,
                       (unsigned long long *)pmm->d_dram_stamp, (unsigned long long)pmm->round, (const uint32_t *)state->cpu_resident[0]);
        // ---- service the round, note its list moves (new va_blocks take chunks), bring its pages in ----
        st = uvmtier_service_pass_chunk(ctx, nb, d_ordered + e0, d_groups + e0, d_batches + b0, lay, state, d_entry_flags + e0,
                                        d_service + e0, d_sruns, svc_run_cap, plan + 46, b0);
        if (st != NV_OK)
            return st;
        UVMTIER_LAUNCH(k_pmm_note_service, (int)egrid, kPmmThreads, 0, ctx->stream, nb, B,
                       (const uvmtier_fault_batch_t *)(d_batches + b0), (const uvmtier_service_group_t *)(d_service + e0),
                       (unsigned long long *)pmm->d_stamp, n_slots, (unsigned long long)pmm->clock, d_acc + 1, chunks);
        pmm->clock += total;
        if (execute_copies) {
            st = uvmtier_copy_runs_enqueue(ctx, d_sruns, plan + 46, svc_run_cap, svc_cap64,
                                           (1u << pair_dram_hbm) | (1u << pair_cxl_hbm) | (1u << pair_zero_hbm));
            if (st != NV_OK)
                return st;
        }
        passes++;
    }
    // the tallies and the chunks in use now
    UVMTIER_CUDA(cudaMemsetAsync(plan, 0, 8 * sizeof(uint32_t), ctx->stream));
    UVMTIER_CUDA(cudaMemsetAsync(touched, 0, ((size_t)n_slots + 31) / 32 * sizeof(uint32_t), ctx->stream));
    UVMTIER_LAUNCH(k_pmm_slot_keys, slot_grid, kPmmThreads, 0, ctx->stream, n_slots, (const unsigned long long *)pmm->d_stamp,
                   (const unsigned long long *)pmm->d_dram_stamp, (const uint32_t *)touched, (const uint32_t *)state->gpu_resident,
                   (const uint32_t *)state->cpu_resident[0], (const uint32_t *)state->cpu_resident[1], e_weight, e_aux, d_key, d_weight,
                   plan);
    UVMTIER_CUDA(cudaMemcpyAsync(h_cnt, plan, kPlanWords * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
    UVMTIER_CUDA(cudaMemcpyAsync(h_cnt + kPlanWords, d_acc, 3 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, ctx->stream));
    uint32_t chunk_errors = 0;
    if (chunks.errors)
        UVMTIER_CUDA(cudaMemcpyAsync(&chunk_errors, chunks.errors, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
    UVMTIER_CUDA(cudaStreamSynchronize(ctx->stream));
    uvmtier_tiering_stats_t stats;
    memset(&stats, 0, sizeof(stats));
    stats.passes = passes;
    stats.eviction_waves = h_cnt[32];
    stats.evicted_to_dram = h_cnt[33];
    stats.evicted_to_cxl = h_cnt[34];
    stats.blocks_demoted = h_cnt[35];
    stats.overcommitted_passes = h_cnt[36];
    stats.peak_chunks_in_use = h_cnt[37];
    stats.chunks_in_use = h_cnt[1];
    stats.pages_evicted = ((const unsigned long long *)(h_cnt + kPlanWords))[0];
    stats.pages_promoted = ((const unsigned long long *)(h_cnt + kPlanWords))[1];
    stats.pages_demoted = ((const unsigned long long *)(h_cnt + kPlanWords))[2];
    if (out_stats)
        *out_stats = stats;
    if (chunk_errors)
        return NV_ERR_NO_MEMORY;
    return NV_OK;
}

} // extern "C"
