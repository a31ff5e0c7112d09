// simt_kernels.cpp -- C entry points that run the product's kernels under the CPU SIMT emulator.
// TEST INFRASTRUCTURE ONLY ("not gpu" tests); see simt.h.
#include "simt.h"

#include <vector>

#include "prefetch_kernels.cuh"

using namespace uvmtier_k;

extern "C" {

void simt_prefetch_compute_ats(size_t n, const uint32_t *faulted, const uvmtier_va_block_region_t *fr,
                               const uvmtier_va_block_region_t *mr, const uint32_t *residency, uint32_t *out,
                               uint32_t enable, uint32_t threshold, uint32_t min_faults, unsigned grid)
{
    pf_tunables tun = {enable, threshold, min_faults};
    simt::launch(dim3(grid), dim3(kPfThreads), [&] { k_prefetch_compute_ats(n, faulted, fr, mr, residency, out, tun); });
}

void simt_prefetch_get_hint(size_t n, const uvmtier_prefetch_block_t *blocks, const uint32_t *resident,
                            const uint32_t *faulted, const uint32_t *thrashing, const uint32_t *cpu_read,
                            uvmtier_prefetch_info_t *info, uint32_t *out, uint32_t *out_residency, uint32_t enable,
                            uint32_t threshold, uint32_t min_faults, unsigned grid)
{
    pf_tunables tun = {enable, threshold, min_faults};
    simt::launch(dim3(grid), dim3(kPfThreads),
                 [&] { k_prefetch_get_hint(n, blocks, resident, faulted, thrashing, cpu_read, info, out, out_residency, tun); });
}

} // extern "C"

// ---------------------------------------------------------------- fault leg
#include "fault_kernels.cuh"

extern "C" {

void simt_fault_parse(const uvmtier_fault_packet_t *packets, size_t n, const uvmtier_fault_gpu_params_t *gp,
                      uvmtier_fault_buffer_entry_t *out)
{
    simt::launch(dim3(2), dim3(256), [&] { k_fault_parse(packets, n, gp->utlb_per_gpc_count, gp->va_bits, out); });
}

// Same launch geometry as uvmtier_fetch_preprocess_fault_batches_device.
int simt_fault_fetch_preprocess(uvmtier_fault_packet_t *ring, uint32_t ring_entries, uint32_t get, uint32_t put,
                                const uvmtier_fault_gpu_params_t *gp, const uvmtier_instance_ptr_entry_t *tbl,
                                uint32_t tbl_n, int clear_valid, uint32_t *disp, uvmtier_ordered_fault_t *ordered,
                                uvmtier_fault_group_t *groups, uvmtier_fault_batch_t *batches,
                                uvmtier_fault_buffer_entry_t *entries, uint16_t *utlb_pending, int force_cta_kernel)
{
    uint32_t threads = 32, idx_bits = 5;
    while (threads < gp->max_batch_size) { threads <<= 1; idx_bits++; }
    const uint32_t n = put >= get ? put - get : ring_entries - get + put;
    if (!n)
        return 0;
    fault_kparams prm;
    prm.B = gp->max_batch_size;
    prm.utlb_per_gpc = gp->utlb_per_gpc_count;
    prm.utlb_count = gp->utlb_per_gpc_count * gp->gpc_count;
    prm.va_bits = gp->va_bits;
    prm.coalesce = gp->coalesce != 0;
    prm.ring_entries = ring_entries;
    prm.get = get;
    prm.n_entries = n;
    prm.tbl_n = tbl_n;
    prm.idx_bits = idx_bits;
    prm.clear_valid = clear_valid != 0;
    const uint32_t n_batches = (n + prm.B - 1) / prm.B;
    (void)force_cta_kernel;
    if (prm.B > kFaultMaxThreads) {
        // batches of more than 1024 entries: the strided kernel (64 emulated threads are enough to exercise the striding)
        std::vector<unsigned char> work((size_t)n_batches * threads * kBigBytesPerEntry);
        simt::launch(dim3(n_batches), dim3(64), [&] {
            k_fault_fetch_preprocess_big(prm, threads, work.data(), ring, tbl, disp, ordered, groups, batches, entries, utlb_pending);
        });
        return (int)n_batches;
    }
    if (fault_smem_bytes(threads, prm.utlb_count, tbl_n) > sizeof(simt::g_dynamic_smem))
        return -1;
    simt::launch(dim3(n_batches), dim3(threads),
                 [&] { k_fault_fetch_preprocess(prm, ring, tbl, disp, ordered, groups, batches, entries, utlb_pending); });
    return (int)n_batches;
}

} // extern "C"
