// simt_pmm.cpp -- runs the eviction-order kernels (pmm_kernels.cuh) under the CPU SIMT emulator.
// TEST INFRASTRUCTURE ONLY.
#include "simt.h"

#include "pmm_kernels.cuh"

using namespace uvmtier_k;

extern "C" int simt_pmm_note_service(uint32_t n_batches, uint32_t B, const uvmtier_fault_batch_t *batches,
                                     const uvmtier_service_group_t *service, unsigned long long *stamp, uint32_t n_slots,
                                     unsigned long long clock)
{
    simt::launch(dim3(3), dim3(kPmmThreads), [&] { k_pmm_note_service(n_batches, B, batches, service, stamp, n_slots, clock); });
    return 0;
}

extern "C" int simt_pmm_note_make_resident(uint32_t n_requests, const uvmtier_make_resident_req_t *reqs,
                                           const uvmtier_make_resident_result_t *results, const uint32_t *gpu_resident,
                                           unsigned long long *stamp, uint32_t n_slots, unsigned long long clock)
{
    simt::launch(dim3(2), dim3(kPmmThreads), [&] { k_pmm_note_make_resident(n_requests, reqs, results, gpu_resident, stamp, n_slots, clock); });
    return 0;
}

extern "C" int simt_pmm_count_needed(uint32_t n_batches, uint32_t B, const uvmtier_fault_group_t *groups,
                                     const uvmtier_fault_batch_t *batches, const uvmtier_instance_ptr_entry_t *tbl,
                                     const uvmtier_va_layout_t *lay, const unsigned long long *stamp, uint32_t *seen, uint32_t *out)
{
    service_kparams prm;
    memset(&prm, 0, sizeof(prm));
    prm.va_base = lay->va_base; prm.B = B; prm.n_batches = n_batches; prm.blocks_per_space = lay->blocks_per_space;
    prm.n_spaces = lay->n_spaces; prm.gpu_id = lay->gpu_id; prm.n_slots = lay->n_spaces * lay->blocks_per_space;
    prm.B_shift = svc_log2_or_none(B);
    simt::launch(dim3(3), dim3(kPmmThreads), [&] { k_pmm_count_needed(prm, groups, batches, tbl, stamp, seen, out); });
    return 0;
}

// sel: 4 words of working space; requests: room for k; returns the number of requests written
extern "C" uint32_t simt_pmm_pick(const unsigned long long *stamp, uint32_t n_slots, uint32_t k, uint32_t dest_tier,
                                  uvmtier_make_resident_req_t *requests)
{
    uint32_t sel[4] = {0, 0, 0, 0}, count = 0;
    simt::launch(dim3(1), dim3(kPmmSelectThreads), [&] { k_pmm_select(stamp, n_slots, k, sel); });
    simt::launch(dim3(2), dim3(kPmmThreads), [&] { k_pmm_emit_evictions(stamp, n_slots, sel, dest_tier, requests, &count); });
    return count;
}
