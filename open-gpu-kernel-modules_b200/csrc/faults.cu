// faults.cu -- C-ABI entry points of the fault-batch preprocessing leg (K0/K1).
// Reference: UVM/uvm_gpu_replayable_faults.c (fault_buffer_init_replayable_faults :139-213,
// fetch_fault_buffer_entries :844-986, preprocess_fault_batch :1134-1178).
//
// Layout in HBM.  The ring is the raw array of 32-byte NVC369 packets.  Batch b owns the slice
// [b*B, (b+1)*B) of every per-entry output array, so CTAs never share an output line.
// Algorithmic bytes per ring entry: 32 read + 4 disposition + 16 per coalesced entry
// + 16 per va_block group (+ 32 per batch) written.
#include "common.cuh"

#include <stdlib.h>

#define UVMTIER_AC_PREPROCESS
#include "access_counter_kernels.cuh"   // includes fault_kernels.cuh

using namespace uvmtier_k;


void uvmtier_fault_state_destroy(uvmtier_ctx_t *ctx)
{
    if (!ctx->faults)
        return;
    if (ctx->faults->d_table)
        cudaFree(ctx->faults->d_table);
    if (ctx->faults->d_big)
        cudaFree(ctx->faults->d_big);
    free(ctx->faults);
    ctx->faults = NULL;
}

extern "C" {

NV_STATUS uvmtier_fault_buffer_init(uvmtier_ctx_t *ctx, const uvmtier_fault_gpu_params_t *gp,
                                    const uvmtier_instance_ptr_entry_t *h_table, size_t n_rows)
{
    if (!ctx || !gp || (n_rows && !h_table))
        return NV_ERR_INVALID_ARGUMENT;
    if (gp->max_batch_size < 1 || gp->max_batch_size > kBigMaxBatch || gp->va_bits < 40 || gp->va_bits > 57 ||
        n_rows > kFaultMaxTableRows || gp->utlb_per_gpc_count == 0 || gp->gpc_count == 0 ||
        gp->utlb_per_gpc_count * gp->gpc_count > 4096)
        return NV_ERR_INVALID_ARGUMENT;

    uint32_t threads = 32, idx_bits = 5;
    while (threads < gp->max_batch_size) {
        threads <<= 1;
        idx_bits++;
    }
    const uint32_t rank_bits = 64 - kPageBits - 3 - idx_bits;
    for (size_t i = 0; i < n_rows; i++)
        if (h_table[i].sort_rank >= (1u << rank_bits))
            return NV_ERR_INVALID_ARGUMENT;

    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    uvmtier_fault_state_destroy(ctx);
    uvmtier_fault_state *fs = (uvmtier_fault_state *)calloc(1, sizeof(*fs));
    if (!fs)
        return NV_ERR_NO_MEMORY;
    fs->gp = *gp;
    fs->gp.coalesce = gp->coalesce != 0;
    fs->n_rows = (uint32_t)n_rows;
    fs->ac_granularity_2m = 1;
    fs->ac_cpu_chunk_2m = 1;
    static std::atomic<uint32_t> generation{0};
    fs->generation = ++generation;
    fs->threads = threads;
    fs->idx_bits = idx_bits;
    fs->smem_bytes = fault_smem_bytes(threads, gp->utlb_per_gpc_count * gp->gpc_count, (uint32_t)n_rows);
    if (gp->max_batch_size > kFaultMaxThreads) {
        // more entries per batch than a CTA has threads: the strided kernel, per-uTLB tables in shared memory
        fs->big_T = threads;
        fs->threads = kBigThreads;
        fs->smem_bytes = (((size_t)gp->utlb_per_gpc_count * gp->gpc_count * 8 + 15) & ~(size_t)15) +
                         (threads <= kBigSmemEntries ? (size_t)threads * kBigScanBytesPerEntry : 0);
    }
    ctx->faults = fs;
    UVMTIER_CUDA(cudaMalloc(&fs->d_table, (n_rows ? n_rows : 1) * sizeof(*h_table)));
    if (n_rows) {
        UVMTIER_CUDA(cudaMemcpyAsync(fs->d_table, h_table, n_rows * sizeof(*h_table), cudaMemcpyHostToDevice, ctx->stream));
        UVMTIER_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    if (fs->smem_bytes > 200 * 1024)
        return NV_ERR_INVALID_ARGUMENT;                     // more uTLBs than a CTA's shared memory can track
    if (fs->smem_bytes > 48 * 1024 && !fs->big_T)
        UVMTIER_CUDA(cudaFuncSetAttribute(k_fault_fetch_preprocess, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)fs->smem_bytes));
    if (fs->smem_bytes > 48 * 1024 && fs->big_T)
        UVMTIER_CUDA(cudaFuncSetAttribute(k_fault_fetch_preprocess_big, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)fs->smem_bytes));
    return NV_OK;
}

NV_STATUS uvmtier_fault_buffer_set_utlb_pending_output(uvmtier_ctx_t *ctx, uint16_t *d_utlb_pending)
{
    if (!ctx)
        return NV_ERR_INVALID_ARGUMENT;
    if (!ctx->faults)
        return NV_ERR_INVALID_STATE;
    ctx->faults->d_utlb_pending = d_utlb_pending;
    static std::atomic<uint32_t> bump{0x80000000u};
    ctx->faults->generation = ++bump;                       // a replayed graph of the streamed entry point holds the old pointer
    return NV_OK;
}

NV_STATUS uvmtier_fault_buffer_parse_entries_device(uvmtier_ctx_t *ctx, const uvmtier_fault_packet_t *d_packets, size_t n,
                                                    uvmtier_fault_buffer_entry_t *d_out)
{
    if (!ctx || (n && (!d_packets || !d_out)))
        return NV_ERR_INVALID_ARGUMENT;
    if (!ctx->faults)
        return NV_ERR_INVALID_STATE;
    if (!n)
        return NV_OK;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    size_t grid = (n + 255) / 256, cap = (size_t)ctx->sm_count * 8;
    if (grid > cap)
        grid = cap;
    UVMTIER_LAUNCH(k_fault_parse, (int)grid, 256, 0, ctx->stream, d_packets, n, ctx->faults->gp.utlb_per_gpc_count,
                   ctx->faults->gp.va_bits, d_out);
    return NV_OK;
}

NV_STATUS uvmtier_fault_buffer_parse_entries(uvmtier_ctx_t *ctx, const uvmtier_fault_packet_t *h_packets, size_t n,
                                             uvmtier_fault_buffer_entry_t *h_out)
{
    if (!ctx || (n && (!h_packets || !h_out)))
        return NV_ERR_INVALID_ARGUMENT;
    if (!ctx->faults)
        return NV_ERR_INVALID_STATE;
    if (!n)
        return NV_OK;
    const size_t pb = uvmtier_align_up(n * 32, 256);
    NV_STATUS st = uvmtier_stage_reserve(ctx, pb + n * sizeof(*h_out), 0);
    if (st != NV_OK)
        return st;
    char *d = (char *)ctx->d_stage;
    UVMTIER_CUDA(cudaMemcpyAsync(d, h_packets, n * 32, cudaMemcpyHostToDevice, ctx->stream));
    st = uvmtier_fault_buffer_parse_entries_device(ctx, (const uvmtier_fault_packet_t *)d, n,
                                                   (uvmtier_fault_buffer_entry_t *)(d + pb));
    if (st != NV_OK)
        return st;
    UVMTIER_CUDA(cudaMemcpyAsync(h_out, d + pb, n * sizeof(*h_out), cudaMemcpyDeviceToHost, ctx->stream));
    UVMTIER_CUDA(cudaStreamSynchronize(ctx->stream));
    return NV_OK;
}

NV_STATUS uvmtier_fetch_preprocess_fault_batches_device(uvmtier_ctx_t *ctx, uvmtier_fault_packet_t *d_ring,
                                                        uint32_t ring_entries, uint32_t get, uint32_t put,
                                                        int clear_valid, uint32_t *d_disposition,
                                                        uvmtier_ordered_fault_t *d_ordered, uvmtier_fault_group_t *d_groups,
                                                        uvmtier_fault_batch_t *d_batches,
                                                        uvmtier_fault_buffer_entry_t *d_entries, uint32_t *out_get)
{
    return uvmtier_fetch_preprocess_on_stream(ctx, ctx ? ctx->stream : NULL, d_ring, ring_entries, get, put, clear_valid,
                                              d_disposition, d_ordered, d_groups, d_batches, d_entries, out_get);
}

} // extern "C"

// Working space of the big-batch kernel for up to n_batches batches per launch (allocates: not while a stream is capturing;
// the streamed entry points reserve for the whole ring before they start).
NV_STATUS uvmtier_fault_big_reserve(uvmtier_ctx_t *ctx, uint32_t n_batches)
{
    uvmtier_fault_state *fs = ctx->faults;
    if (!fs->big_T)
        return NV_OK;
    const size_t need = (size_t)n_batches * fs->big_T * kBigBytesPerEntry;
    if (need <= fs->big_bytes)
        return NV_OK;
    UVMTIER_CUDA(cudaDeviceSynchronize());                  // a launch on another of the context's streams may still read it
    if (fs->d_big)
        UVMTIER_CUDA(cudaFree(fs->d_big));
    fs->d_big = NULL;
    fs->big_bytes = 0;
    UVMTIER_CUDA(cudaMalloc(&fs->d_big, need));
    fs->big_bytes = need;
    return NV_OK;
}

NV_STATUS uvmtier_fetch_preprocess_on_stream(uvmtier_ctx_t *ctx, cudaStream_t stream, uvmtier_fault_packet_t *d_ring,
                                             uint32_t ring_entries, uint32_t get, uint32_t put, int clear_valid,
                                             uint32_t *d_disposition, uvmtier_ordered_fault_t *d_ordered,
                                             uvmtier_fault_group_t *d_groups, uvmtier_fault_batch_t *d_batches,
                                             uvmtier_fault_buffer_entry_t *d_entries, uint32_t *out_get, uint32_t utlb_batch_base)
{
    if (!ctx || !d_ring || !ring_entries || get >= ring_entries || put >= ring_entries)
        return NV_ERR_INVALID_ARGUMENT;
    if (!ctx->faults)
        return NV_ERR_INVALID_STATE;
    const uvmtier_fault_state *fs = ctx->faults;
    const uint32_t n = put >= get ? put - get : ring_entries - get + put;
    if (out_get)
        *out_get = put;
    if (!n)
        return NV_OK;
    if (!d_disposition || !d_ordered || !d_groups || !d_batches)
        return NV_ERR_INVALID_ARGUMENT;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    fault_kparams prm;
    prm.B = fs->gp.max_batch_size;
    prm.utlb_per_gpc = fs->gp.utlb_per_gpc_count;
    prm.utlb_count = fs->gp.utlb_per_gpc_count * fs->gp.gpc_count;
    prm.va_bits = fs->gp.va_bits;
    prm.coalesce = fs->gp.coalesce && ctx->cfg.uvm_perf_fault_coalesce;
    prm.ring_entries = ring_entries;
    prm.get = get;
    prm.n_entries = n;
    prm.tbl_n = fs->n_rows;
    prm.idx_bits = fs->idx_bits;
    prm.clear_valid = clear_valid != 0;
    const uint32_t n_batches = (n + prm.B - 1) / prm.B;
    // the optional per-uTLB counts of batch b land at d_utlb_pending[(utlb_batch_base + b) * utlb_count + u]
    uint16_t *up = fs->d_utlb_pending ? fs->d_utlb_pending + (size_t)utlb_batch_base * prm.utlb_count : NULL;
    if (fs->big_T) {
        const NV_STATUS rs = uvmtier_fault_big_reserve(ctx, n_batches);
        if (rs != NV_OK)
            return rs;
        UVMTIER_LAUNCH(k_fault_fetch_preprocess_big, n_batches, kBigThreads, fs->smem_bytes, stream, prm, fs->big_T, fs->d_big, d_ring,
                       (const uvmtier_instance_ptr_entry_t *)fs->d_table, d_disposition, d_ordered, d_groups, d_batches, d_entries, up);
        return NV_OK;
    }
    UVMTIER_LAUNCH(k_fault_fetch_preprocess, n_batches, fs->threads, fs->smem_bytes, stream, prm, d_ring, fs->d_table,
                   d_disposition, d_ordered, d_groups, d_batches, d_entries, up);
    return NV_OK;
}

extern "C" {

NV_STATUS uvmtier_fetch_preprocess_fault_batches_streamed(uvmtier_ctx_t *ctx, const uvmtier_fault_packet_t *h_ring,
                                                          uvmtier_fault_packet_t *d_ring, uint32_t ring_entries, uint32_t get,
                                                          uint32_t put, int clear_valid, uint32_t chunk_batches,
                                                          uint32_t *d_disposition, uvmtier_ordered_fault_t *d_ordered,
                                                          uvmtier_fault_group_t *d_groups, uvmtier_fault_batch_t *d_batches,
                                                          uvmtier_fault_buffer_entry_t *d_entries, uint32_t *out_get)
{
    if (!ctx || !h_ring || !d_ring || !ring_entries || get >= ring_entries || put >= ring_entries)
        return NV_ERR_INVALID_ARGUMENT;
    if (!ctx->faults)
        return NV_ERR_INVALID_STATE;
    const uint32_t n = put >= get ? put - get : ring_entries - get + put;
    if (out_get)
        *out_get = put;
    if (!n)
        return NV_OK;
    if (!d_disposition || !d_ordered || !d_groups || !d_batches)
        return NV_ERR_INVALID_ARGUMENT;
    uvmtier_chunk_plan plan;
    NV_STATUS st = uvmtier_streamed_begin(ctx, ring_entries, get, put, chunk_batches, &plan);
    if (st != NV_OK)
        return st;
    for (uint32_t c = 0; c < plan.n_chunks; c++) {
        uint32_t e0, e1, first, last;
        st = uvmtier_streamed_copy_chunk(ctx, plan, c, h_ring, d_ring, ring_entries, get, &e0, &e1, &first, &last);
        if (st != NV_OK)
            return st;
        st = uvmtier_fetch_preprocess_on_stream(ctx, ctx->stream, d_ring, ring_entries, first, last, clear_valid,
                                                d_disposition + e0, d_ordered + e0, d_groups + e0, d_batches + c * plan.per,
                                                d_entries ? d_entries + e0 : NULL, NULL, c * plan.per);
        if (st != NV_OK)
            return st;
    }
    return NV_OK;
}

NV_STATUS uvmtier_fetch_preprocess_fault_batches(uvmtier_ctx_t *ctx, uvmtier_fault_packet_t *h_ring, uint32_t ring_entries,
                                                 uint32_t get, uint32_t put, int clear_valid, uint32_t *h_disposition,
                                                 uvmtier_ordered_fault_t *h_ordered, uvmtier_fault_group_t *h_groups,
                                                 uvmtier_fault_batch_t *h_batches, uvmtier_fault_buffer_entry_t *h_entries,
                                                 uint32_t *out_get)
{
    if (!ctx || !h_ring || !ring_entries || get >= ring_entries || put >= ring_entries)
        return NV_ERR_INVALID_ARGUMENT;
    if (!ctx->faults)
        return NV_ERR_INVALID_STATE;
    const uint32_t n = put >= get ? put - get : ring_entries - get + put;
    if (!n) {
        if (out_get)
            *out_get = put;
        return NV_OK;
    }
    const uint32_t B = ctx->faults->gp.max_batch_size, n_batches = (n + B - 1) / B;
    const size_t rb = uvmtier_align_up((size_t)ring_entries * 32, 256), db = uvmtier_align_up((size_t)n * 4, 256),
                 ob = uvmtier_align_up((size_t)n * 16, 256), gb = ob, bb = uvmtier_align_up((size_t)n_batches * 32, 256),
                 eb = h_entries ? uvmtier_align_up((size_t)n * sizeof(*h_entries), 256) : 0;
    NV_STATUS st = uvmtier_stage_reserve(ctx, rb + db + ob + gb + bb + eb, 0);
    if (st != NV_OK)
        return st;
    char *d = (char *)ctx->d_stage;
    uvmtier_fault_packet_t *d_ring = (uvmtier_fault_packet_t *)d;
    uint32_t *d_disp = (uint32_t *)(d + rb);
    uvmtier_ordered_fault_t *d_ord = (uvmtier_ordered_fault_t *)(d + rb + db);
    uvmtier_fault_group_t *d_grp = (uvmtier_fault_group_t *)(d + rb + db + ob);
    uvmtier_fault_batch_t *d_bat = (uvmtier_fault_batch_t *)(d + rb + db + ob + gb);
    uvmtier_fault_buffer_entry_t *d_ent = h_entries ? (uvmtier_fault_buffer_entry_t *)(d + rb + db + ob + gb + bb) : NULL;
    UVMTIER_CUDA(cudaMemcpyAsync(d_ring, h_ring, (size_t)ring_entries * 32, cudaMemcpyHostToDevice, ctx->stream));
    // slots past num_coalesced / num_block_groups are not written by the kernel: give them a defined value
    UVMTIER_CUDA(cudaMemsetAsync(d_ord, 0xff, ob + gb, ctx->stream));
    if (d_ent)
        UVMTIER_CUDA(cudaMemsetAsync(d_ent, 0, eb, ctx->stream));
    st = uvmtier_fetch_preprocess_fault_batches_device(ctx, d_ring, ring_entries, get, put, clear_valid, d_disp, d_ord,
                                                       d_grp, d_bat, d_ent, out_get);
    if (st != NV_OK)
        return st;
    if (clear_valid)
        UVMTIER_CUDA(cudaMemcpyAsync(h_ring, d_ring, (size_t)ring_entries * 32, cudaMemcpyDeviceToHost, ctx->stream));
    UVMTIER_CUDA(cudaMemcpyAsync(h_disposition, d_disp, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    UVMTIER_CUDA(cudaMemcpyAsync(h_ordered, d_ord, (size_t)n * 16, cudaMemcpyDeviceToHost, ctx->stream));
    UVMTIER_CUDA(cudaMemcpyAsync(h_groups, d_grp, (size_t)n * 16, cudaMemcpyDeviceToHost, ctx->stream));
    UVMTIER_CUDA(cudaMemcpyAsync(h_batches, d_bat, (size_t)n_batches * 32, cudaMemcpyDeviceToHost, ctx->stream));
    if (h_entries)
        UVMTIER_CUDA(cudaMemcpyAsync(h_entries, d_ent, (size_t)n * sizeof(*h_entries), cudaMemcpyDeviceToHost, ctx->stream));
    UVMTIER_CUDA(cudaStreamSynchronize(ctx->stream));
    return NV_OK;
}

} // extern "C"

// Non-replayable faults (include/uvmtier.h section 1g): parse + translate, one run per fault, then the servicing pass
// without the "already authorised" shortcut.
extern "C" NV_STATUS uvmtier_service_non_replayable_faults_device(uvmtier_ctx_t *ctx, const uvmtier_fault_packet_t *d_shadow_buffer,
                                                                  uint32_t n_faults, const uvmtier_va_layout_t *lay,
                                                                  const uvmtier_block_state_t *state,
                                                                  uvmtier_fault_buffer_entry_t *d_entries, uint16_t *d_entry_flags,
                                                                  uvmtier_service_group_t *d_service, uvmtier_copy_run_t *d_runs,
                                                                  uint32_t run_cap, uint32_t *d_counters)
{
    if (!ctx || !lay || !state)
        return NV_ERR_INVALID_ARGUMENT;
    if (!ctx->faults)
        return NV_ERR_INVALID_STATE;
    if (!n_faults)
        return NV_OK;
    if (!d_shadow_buffer)
        return NV_ERR_INVALID_ARGUMENT;
    const uint32_t B = ctx->faults->gp.max_batch_size;
    const uint32_t nb = (n_faults + B - 1) / B;
    const size_t total = (size_t)nb * B;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    // working space: the sorted view the servicing pass takes (here in buffer order) and its run and batch arrays
    const size_t ob = uvmtier_align_up(total * sizeof(uvmtier_ordered_fault_t), 256), gb = uvmtier_align_up(total * sizeof(uvmtier_fault_group_t), 256);
    NV_STATUS st = uvmtier_stage_reserve(ctx, ob + gb + (size_t)nb * sizeof(uvmtier_fault_batch_t), 0);
    if (st != NV_OK)
        return st;
    uvmtier_ordered_fault_t *d_ord = (uvmtier_ordered_fault_t *)ctx->d_stage;
    uvmtier_fault_group_t *d_grp = (uvmtier_fault_group_t *)((char *)ctx->d_stage + ob);
    uvmtier_fault_batch_t *d_bat = (uvmtier_fault_batch_t *)((char *)ctx->d_stage + ob + gb);
    if (!d_entry_flags || !d_service || !d_counters || (run_cap && !d_runs) || nb > 65536u)
        return NV_ERR_INVALID_ARGUMENT;
    UVMTIER_CUDA(cudaMemsetAsync(d_counters, 0, 2 * sizeof(uint32_t), ctx->stream));
    UVMTIER_CUDA(cudaMemsetAsync(d_entry_flags, 0, total * sizeof(uint16_t), ctx->stream));
    UVMTIER_CUDA(cudaMemsetAsync(d_service, 0, total * sizeof(uvmtier_service_group_t), ctx->stream));   // dropped entries: all zero
    size_t grid = (n_faults + 255) / 256, cap = (size_t)ctx->sm_count * 8;
    if (grid > cap)
        grid = cap;
    UVMTIER_LAUNCH(k_nonreplayable_prepare, (int)grid, 256, 0, ctx->stream, d_shadow_buffer, n_faults, B,
                   ctx->faults->gp.utlb_per_gpc_count, ctx->faults->gp.va_bits, ctx->faults->d_table, ctx->faults->n_rows, d_ord, d_grp,
                   d_bat, d_entry_flags, d_entries);
    return uvmtier_service_pass_chunk(ctx, nb, d_ord, d_grp, d_bat, lay, state, d_entry_flags, d_service, d_runs, run_cap, d_counters, 0, true);
}

// ------------------------------------------------------------------------------------------------
// Access-counter notifications, front end (include/uvmtier.h section 1e).
extern "C" NV_STATUS uvmtier_access_counters_init(uvmtier_ctx_t *ctx, const uvmtier_access_counter_config_t *cfg)
{
    if (!ctx)
        return NV_ERR_INVALID_ARGUMENT;
    if (!ctx->faults)
        return NV_ERR_INVALID_STATE;                       // the table and the batch size come from uvmtier_fault_buffer_init
    const uint64_t gran = cfg && cfg->granularity_bytes ? cfg->granularity_bytes : (2ull << 20);
    const uint64_t chunk = cfg && cfg->cpu_chunk_bytes ? cfg->cpu_chunk_bytes : (2ull << 20);
    if ((gran != (64ull << 10) && gran != (2ull << 20) && gran != (16ull << 20) && gran != (16ull << 30)) ||   // :198-218
        (chunk != 4096 && chunk != (64ull << 10) && chunk != (2ull << 20)))
        return NV_ERR_INVALID_ARGUMENT;
    ctx->faults->ac_granularity_2m = gran == (2ull << 20);
    ctx->faults->ac_cpu_chunk_2m = chunk == (2ull << 20);
    return NV_OK;
}

extern "C" NV_STATUS uvmtier_access_counters_fetch_preprocess_device(uvmtier_ctx_t *ctx, uvmtier_access_counter_packet_t *d_ring,
                                                                     uint32_t ring_entries, uint32_t get, uint32_t put,
                                                                     int clear_valid, uvmtier_ordered_notification_t *d_ordered,
                                                                     uvmtier_fault_group_t *d_groups, uvmtier_fault_batch_t *d_batches,
                                                                     uvmtier_access_counter_entry_t *d_entries, uint32_t *out_get)
{
    if (!ctx || !d_ring || !ring_entries || get >= ring_entries || put >= ring_entries)
        return NV_ERR_INVALID_ARGUMENT;
    if (!ctx->faults)
        return NV_ERR_INVALID_STATE;
    const uvmtier_fault_state *fs = ctx->faults;
    const uint32_t n = put >= get ? put - get : ring_entries - get + put;
    if (out_get)
        *out_get = put;
    if (!n)
        return NV_OK;
    if (!d_ordered || !d_groups || !d_batches)
        return NV_ERR_INVALID_ARGUMENT;
    if (fs->big_T)
        return NV_ERR_NOT_SUPPORTED;                        // notification batches are one thread per entry: max_batch_size <= 1024
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    ac_kparams prm;
    prm.B = fs->gp.max_batch_size;
    prm.va_bits = fs->gp.va_bits;
    prm.ring_entries = ring_entries;
    prm.get = get;
    prm.n_entries = n;
    prm.tbl_n = fs->n_rows;
    prm.idx_bits = fs->idx_bits;
    prm.clear_valid = clear_valid != 0;
    UVMTIER_LAUNCH(k_access_counters_fetch_preprocess, (n + prm.B - 1) / prm.B, fs->threads, 0, ctx->stream, prm, d_ring,
                   fs->d_table, d_ordered, d_groups, d_batches, d_entries);
    return NV_OK;
}

// ------------------------------------------------------------------------------------------------
// Chunk plan and host->device streaming shared by the streamed entry points (here and in service.cu).
NV_STATUS uvmtier_streamed_begin(uvmtier_ctx_t *ctx, uint32_t ring_entries, uint32_t get, uint32_t put, uint32_t chunk_batches,
                                 uvmtier_chunk_plan *plan)
{
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    plan->n = put >= get ? put - get : ring_entries - get + put;
    plan->B = ctx->faults->gp.max_batch_size;
    plan->n_batches = (plan->n + plan->B - 1) / plan->B;
    uint32_t per = chunk_batches ? chunk_batches : (plan->n_batches + 7) / 8;
    if ((plan->n_batches + per - 1) / per > UVMTIER_MAX_COPY_CHUNKS)
        per = (plan->n_batches + UVMTIER_MAX_COPY_CHUNKS - 1) / UVMTIER_MAX_COPY_CHUNKS;
    plan->per = per;
    plan->n_chunks = (plan->n_batches + per - 1) / per;
    {
        const NV_STATUS rs = uvmtier_fault_big_reserve(ctx, per);         // before any capture starts
        if (rs != NV_OK)
            return rs;
    }
    if (!ctx->copy_stream)
        UVMTIER_CUDA(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
    for (uint32_t c = 0; c <= plan->n_chunks; c++)
        if (!ctx->copy_events[c])
            UVMTIER_CUDA(cudaEventCreateWithFlags(&ctx->copy_events[c], cudaEventDisableTiming));
    // the copies may overwrite d_ring only after everything already queued on the context's stream (an earlier
    // pass reading it) is done
    UVMTIER_CUDA(cudaEventRecord(ctx->copy_events[plan->n_chunks], ctx->stream));
    UVMTIER_CUDA(cudaStreamWaitEvent(ctx->copy_stream, ctx->copy_events[plan->n_chunks], 0));
    return NV_OK;
}

NV_STATUS uvmtier_streamed_copy_chunk(uvmtier_ctx_t *ctx, const uvmtier_chunk_plan &plan, uint32_t c,
                                      const uvmtier_fault_packet_t *h_ring, uvmtier_fault_packet_t *d_ring,
                                      uint32_t ring_entries, uint32_t get, uint32_t *e0_out, uint32_t *e1_out,
                                      uint32_t *first_out, uint32_t *last_out, cudaStream_t waiter)
{
    const uint32_t B = plan.B, per = plan.per, n = plan.n;
    const uint32_t e0 = c * per * B, e1 = (c + 1) * per * B < n ? (c + 1) * per * B : n;       // entries, from `get`
    const uint32_t first = (uint32_t)(((uint64_t)get + e0) % ring_entries), cnt = e1 - e0;
    const uint32_t head = cnt < ring_entries - first ? cnt : ring_entries - first;          // up to the wrap (:975)
    UVMTIER_CUDA(cudaMemcpyAsync(d_ring + first, h_ring + first, (size_t)head * 32, cudaMemcpyHostToDevice, ctx->copy_stream));
    if (head < cnt)
        UVMTIER_CUDA(cudaMemcpyAsync(d_ring, h_ring, (size_t)(cnt - head) * 32, cudaMemcpyHostToDevice, ctx->copy_stream));
    UVMTIER_CUDA(cudaEventRecord(ctx->copy_events[c], ctx->copy_stream));
    UVMTIER_CUDA(cudaStreamWaitEvent(waiter ? waiter : ctx->stream, ctx->copy_events[c], 0));
    *e0_out = e0;
    *e1_out = e1;
    *first_out = first;
    *last_out = (uint32_t)(((uint64_t)get + e1) % ring_entries);
    return NV_OK;
}
