// simt_kernels.cpp -- C entry points that run the product's kernels under the CPU SIMT emulator.
// TEST INFRASTRUCTURE ONLY ("not gpu" tests); see simt.h.
#include "simt.h"

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
