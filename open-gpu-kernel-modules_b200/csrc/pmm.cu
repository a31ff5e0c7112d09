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
    UVMTIER_CUDA(cudaMemsetAsync(out->d_stamp, 0, n * sizeof(uint64_t), ctx->stream));
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
    if (pmm->d_stamp) cudaFree(pmm->d_stamp);
    if (pmm->d_scratch) cudaFree(pmm->d_scratch);
    memset(pmm, 0, sizeof(*pmm));
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
                   (unsigned long long *)pmm->d_stamp, pmm->n_slots, (unsigned long long)pmm->clock);
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
    UVMTIER_LAUNCH(k_pmm_note_make_resident, (int)grid, kPmmThreads, 0, ctx->stream, n_requests, d_requests, d_results,
                   (const uint32_t *)state->gpu_resident, (unsigned long long *)pmm->d_stamp, pmm->n_slots,
                   (unsigned long long)pmm->clock);
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
                                                      uint32_t chunk_batches, uint64_t dram_capacity_pages, int execute_copies,
                                                      uvmtier_tiering_stats_t *out_stats)
{
    if (out_stats)
        memset(out_stats, 0, sizeof(*out_stats));
    if (!pmm_ok(ctx, pmm) || !lay || !state || !chunk_batches || (size_t)lay->n_spaces * lay->blocks_per_space != pmm->n_slots)
        return NV_ERR_INVALID_ARGUMENT;
    if (!ctx->faults || (execute_copies && !ctx->tiers))
        return NV_ERR_INVALID_STATE;
    if (!n_batches)
        return NV_OK;
    if (!d_ordered || !d_groups || !d_batches || !d_entry_flags || !d_service || n_batches > 65536u)   // 16-bit batch numbers
        return NV_ERR_INVALID_ARGUMENT;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    const uint32_t B = ctx->faults->gp.max_batch_size;
    const uint64_t chunk_entries = (uint64_t)chunk_batches * B;
    // victims per wave: never more than the va_blocks one chunk of batches can touch, and at most 65536 (the run
    // array of a wave is sized for them; a larger overflow takes more waves)
    const uint64_t kmax64 = chunk_entries < pmm->n_slots ? chunk_entries : pmm->n_slots;
    const uint32_t kmax = (uint32_t)(kmax64 < 65536 ? kmax64 : 65536);
    const uint32_t evict_run_cap = kmax * 256u;                           // a block's GPU pages: at most 256 runs
    // copy runs of one servicing pass, worst case: every page a va_block receives is a run of its own, a va_block
    // receives at most its 512 pages, and a pass touches at most min(entries, slots) va_blocks -- so the list is never
    // cut and the copies never run on a truncated list
    const uint64_t svc_blocks = chunk_entries < pmm->n_slots ? chunk_entries : pmm->n_slots;
    const uint64_t svc_cap64 = svc_blocks * UVMTIER_PAGES_PER_VA_BLOCK;
    const uint32_t svc_run_cap = (uint32_t)(svc_cap64 < 0x7fffffffull ? svc_cap64 : 0x7fffffffull);
    // working space: counters, tallies, requests, results, the two run arrays
    const size_t o_cnt = 0, o_acc = 256, o_req = 512, o_res = uvmtier_align_up(o_req + (size_t)kmax * sizeof(uvmtier_make_resident_req_t), 256),
                 o_eruns = uvmtier_align_up(o_res + (size_t)kmax * sizeof(uvmtier_make_resident_result_t), 256),
                 o_sruns = uvmtier_align_up(o_eruns + (size_t)evict_run_cap * sizeof(uvmtier_copy_run_t), 256),
                 bytes = o_sruns + (size_t)svc_run_cap * sizeof(uvmtier_copy_run_t);
    NV_STATUS st = uvmtier_stage_reserve(ctx, bytes, 256);
    if (st != NV_OK)
        return st;
    char *d = (char *)ctx->d_stage;
    uint32_t *d_cnt = (uint32_t *)(d + o_cnt);                            // [0] needed [1] in use [2..4] pages [5] picked
    unsigned long long *d_acc = (unsigned long long *)(d + o_acc);        //   [6] eviction runs [8] service runs
    uvmtier_make_resident_req_t *d_req = (uvmtier_make_resident_req_t *)(d + o_req);
    uvmtier_make_resident_result_t *d_res = (uvmtier_make_resident_result_t *)(d + o_res);
    uvmtier_copy_run_t *d_eruns = (uvmtier_copy_run_t *)(d + o_eruns), *d_sruns = (uvmtier_copy_run_t *)(d + o_sruns);
    uint32_t *h_cnt = (uint32_t *)ctx->h_stage;
    UVMTIER_CUDA(cudaMemsetAsync(d, 0, 512, ctx->stream));
    UVMTIER_CUDA(cudaMemsetAsync(d_entry_flags, 0, (size_t)n_batches * B * sizeof(uint16_t), ctx->stream));
    uvmtier_tiering_stats_t stats;
    memset(&stats, 0, sizeof(stats));
    for (uint32_t b0 = 0; b0 < n_batches; b0 += chunk_batches) {
        const uint32_t nb = n_batches - b0 < chunk_batches ? n_batches - b0 : chunk_batches;
        const size_t e0 = (size_t)b0 * B;
        // Make room: count what the chunk needs, evict the overflow, count again -- a victim that this very chunk of
        // batches faults on comes back and needs a chunk again, so the second count can ask for a few more.
        // Rounds repeat until nothing overflows.  A victim that this chunk of batches faults on needs a chunk again, so
        // the overflow shrinks geometrically rather than at once; a wave takes at most kmax victims.  The bound covers
        // emptying HBM wave by wave plus the recounts; a pass that still does not fit is serviced over capacity and
        // counted (stats.overcommitted_passes).
        uint32_t needed = 0, in_use = 0;
        bool fits = false;
        const int max_rounds = (int)(pmm->n_slots / (kmax ? kmax : 1u)) + 64;
        for (int round = 0; round < max_rounds; round++) {
            st = uvmtier_pmm_count_needed_device(ctx, pmm, nb, d_groups + e0, d_batches + b0, lay, d_cnt);
            if (st != NV_OK)
                return st;
            st = uvmtier_block_state_count_pages_device(ctx, lay, state, d_cnt + 2);
            if (st != NV_OK)
                return st;
            UVMTIER_CUDA(cudaMemcpyAsync(h_cnt, d_cnt, 10 * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
            UVMTIER_CUDA(cudaStreamSynchronize(ctx->stream));
            if (h_cnt[6] > evict_run_cap || h_cnt[8] > svc_run_cap)       // an earlier copy list was cut
                return NV_ERR_NO_MEMORY;
            needed = h_cnt[0];
            in_use = h_cnt[1];
            const uint64_t dram_pages = h_cnt[3];
            uint32_t k = (uint64_t)in_use + needed > pmm->capacity_chunks ? in_use + needed - pmm->capacity_chunks : 0;
            if (k > in_use)
                k = in_use;                                               // the chunk alone overflows HBM: evict all there is
            if (k > kmax)
                k = kmax;
            if (!k) {
                fits = true;
                break;
            }
            const uint32_t dest = dram_pages + (uint64_t)k * UVMTIER_PAGES_PER_VA_BLOCK <= dram_capacity_pages ? UVMTIER_TIER_DRAM
                                                                                                                : UVMTIER_TIER_CXL;
            st = uvmtier_pmm_pick_chunks_to_evict_device(ctx, pmm, k, dest, d_req, d_cnt + 5);
            if (st != NV_OK)
                return st;
            UVMTIER_CUDA(cudaMemsetAsync(d_cnt + 6, 0, 2 * sizeof(uint32_t), ctx->stream));
            st = uvmtier_make_resident_device(ctx, k, d_req, NULL, NULL, lay, state, d_res, d_eruns, evict_run_cap, d_cnt + 6);
            if (st != NV_OK)
                return st;
            UVMTIER_LAUNCH(k_pmm_note_make_resident, (int)((k + kPmmThreads - 1) / kPmmThreads), kPmmThreads, 0, ctx->stream, k,
                           (const uvmtier_make_resident_req_t *)d_req, (const uvmtier_make_resident_result_t *)d_res,
                           (const uint32_t *)state->gpu_resident, (unsigned long long *)pmm->d_stamp, pmm->n_slots,
                           (unsigned long long)pmm->clock, d_acc);
            pmm->clock += k;
            if (execute_copies) {
                st = uvmtier_copy_runs_device(ctx, d_eruns, d_cnt + 6, evict_run_cap, 0);
                if (st != NV_OK)
                    return st;
            }
            stats.eviction_waves++;
            if (dest == UVMTIER_TIER_DRAM) stats.evicted_to_dram += k; else stats.evicted_to_cxl += k;
            in_use -= k;                                                  // what the next count will find (if there is none)
        }
        if (!fits || (uint64_t)in_use + needed > pmm->capacity_chunks)
            stats.overcommitted_passes++;                                 // e.g. the chunk of batches alone needs more than HBM holds
        // service the chunk, note its list moves, bring its pages in
        UVMTIER_CUDA(cudaMemsetAsync(d_cnt + 8, 0, 2 * sizeof(uint32_t), ctx->stream));
        st = uvmtier_service_pass_chunk(ctx, nb, d_ordered + e0, d_groups + e0, d_batches + b0, lay, state, d_entry_flags + e0,
                                        d_service + e0, d_sruns, svc_run_cap, d_cnt + 8, b0);
        if (st != NV_OK)
            return st;
        const uint64_t total = (uint64_t)nb * B;
        uint64_t grid = (total + kPmmThreads - 1) / kPmmThreads;
        if (grid > (uint64_t)ctx->sm_count * 8)
            grid = (uint64_t)ctx->sm_count * 8;
        UVMTIER_LAUNCH(k_pmm_note_service, (int)grid, kPmmThreads, 0, ctx->stream, nb, B,
                       (const uvmtier_fault_batch_t *)(d_batches + b0), (const uvmtier_service_group_t *)(d_service + e0),
                       (unsigned long long *)pmm->d_stamp, pmm->n_slots, (unsigned long long)pmm->clock, d_acc + 1);
        pmm->clock += total;
        if (execute_copies) {
            st = uvmtier_copy_runs_device(ctx, d_sruns, d_cnt + 8, svc_run_cap, 0);
            if (st != NV_OK)
                return st;
        }
        stats.passes++;
        const uint32_t after = in_use + needed;
        if (after > stats.peak_chunks_in_use)
            stats.peak_chunks_in_use = after;
        stats.chunks_in_use = after;
    }
    // the tallies, the chunks in use now, and whether a run array overflowed in the last chunk (the others were
    // checked as the loop went)
    st = uvmtier_pmm_count_needed_device(ctx, pmm, 0, NULL, NULL, lay, d_cnt);
    if (st != NV_OK)
        return st;
    UVMTIER_CUDA(cudaMemcpyAsync(h_cnt, d_cnt, 10 * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
    UVMTIER_CUDA(cudaMemcpyAsync(h_cnt + 16, d_acc, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, ctx->stream));
    UVMTIER_CUDA(cudaStreamSynchronize(ctx->stream));
    if (h_cnt[6] > evict_run_cap || h_cnt[8] > svc_run_cap)
        return NV_ERR_NO_MEMORY;
    stats.chunks_in_use = h_cnt[1];
    stats.pages_evicted = ((const unsigned long long *)(h_cnt + 16))[0];
    stats.pages_promoted = ((const unsigned long long *)(h_cnt + 16))[1];
    if (out_stats)
        *out_stats = stats;
    return NV_OK;
}

} // extern "C"
