// simt_service.cpp -- runs the servicing-pass kernels (service_kernels.cuh) under the CPU SIMT emulator with
// the same launch sequence as uvmtier_service_fault_batches_device.  TEST INFRASTRUCTURE ONLY.
#include "simt.h"

#include <vector>

#include "service_kernels.cuh"
#include "service_tpb.cuh"

using namespace uvmtier_k;

extern "C" int simt_service_fault_batches(uint32_t n_batches, uint32_t B, const uvmtier_ordered_fault_t *ordered,
                                          const uvmtier_fault_group_t *groups, uvmtier_fault_batch_t *batches,
                                          const uvmtier_instance_ptr_entry_t *tbl, const uvmtier_va_layout_t *lay,
                                          const uvmtier_block_state_t *state, uint32_t pf_enable, uint32_t pf_threshold,
                                          uint32_t pf_min_faults, uint16_t *entry_flags, uvmtier_service_group_t *service,
                                          uvmtier_copy_run_t *runs, uint32_t run_cap, uint32_t *counters, int warp_per_block)
{
    const uint32_t n_slots = lay->n_spaces * lay->blocks_per_space;
    const size_t total = (size_t)n_batches * B;
    service_kparams prm;
    prm.va_base = lay->va_base; prm.B = B; prm.n_batches = n_batches; prm.blocks_per_space = lay->blocks_per_space;
    prm.n_spaces = lay->n_spaces; prm.gpu_id = lay->gpu_id; prm.n_slots = n_slots; prm.pf_enable = pf_enable;
    prm.pf_threshold = pf_threshold; prm.pf_min_faults = pf_min_faults; prm.run_cap = run_cap;
    block_state_ptrs bs;
    bs.gpu_resident = (uint32_t *)state->gpu_resident;
    bs.cpu_resident[0] = (uint32_t *)state->cpu_resident[0]; bs.cpu_resident[1] = (uint32_t *)state->cpu_resident[1];
    bs.gpu_pte[0] = (uint32_t *)state->gpu_pte[0]; bs.gpu_pte[1] = (uint32_t *)state->gpu_pte[1]; bs.gpu_pte[2] = (uint32_t *)state->gpu_pte[2];
    bs.prefetch_info = state->prefetch_info; bs.logical_prot = state->logical_prot; bs.preferred_location = state->preferred_location;

    std::vector<uint32_t> count(n_slots, 0), offset(n_slots, 0), cursor(n_slots, 0), sums(1024, 0), unmapped(n_batches, 0), list(total + 1, 0);
    counters[0] = counters[1] = 0;
    memset(entry_flags, 0, total * sizeof(uint16_t));
    const uint32_t n_sums = (n_slots + kScanItems - 1) / kScanItems;
    simt::launch(dim3(2), dim3(kSvcThreads), [&] { k_service_count(prm, groups, batches, tbl, count.data(), unmapped.data(), list.data(), 0, offset.data(), cursor.data()); });
    simt::launch(dim3(n_sums), dim3(256), [&] { k_scan_sums(count.data(), n_slots, sums.data()); });
    simt::launch(dim3(1), dim3(1024), [&] { k_scan_top(sums.data(), n_sums); });
    simt::launch(dim3(n_sums), dim3(256), [&] { k_scan_final(count.data(), n_slots, sums.data(), offset.data()); });
    simt::launch(dim3(2), dim3(kSvcThreads), [&] { k_service_count(prm, groups, batches, tbl, count.data(), unmapped.data(), list.data(), 1, offset.data(), cursor.data()); });
    if (warp_per_block) {
        simt::launch(dim3(3), dim3(kSvcThreads), [&] { k_service_blocks(prm, ordered, groups, batches, bs, count.data(), offset.data(), list.data(), entry_flags, service, runs, counters); });
    }
    else {
        simt::launch(dim3(5), dim3(32), [&] { k_service_sort_lists(n_slots, count.data(), offset.data(), list.data()); });
        simt::launch(dim3((n_slots + kTpbThreads - 1) / kTpbThreads), dim3(kTpbThreads),
                     [&] { k_service_blocks_tpb(prm, ordered, groups, batches, bs, count.data(), offset.data(), list.data(), entry_flags, service, runs, counters); });
    }
    simt::launch(dim3((n_batches + 127) / 128), dim3(128), [&] { k_service_unmapped(prm, ordered, groups, batches, tbl, unmapped.data(), entry_flags, service); });
    return 0;
}
