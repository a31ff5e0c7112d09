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

// The tiered loop's planning kernels: touched -> per-slot keys and totals, the three plan steps, the two weighted selections.
// plan: kPlanWords words (zeroed by the caller); returns nothing, the caller reads plan / the arrays.
extern "C" int simt_pmm_plan_round(uint32_t n_slots, const unsigned long long *stamp, const unsigned long long *dram_stamp,
                                   const uint32_t *touched, const uint32_t *gpu_resident, const uint32_t *cpu0, const uint32_t *cpu1,
                                   uint32_t capacity_chunks, unsigned long long dram_capacity_pages, int demote, uint32_t *e_weight,
                                   uint32_t *e_aux, unsigned long long *d_key, uint32_t *d_weight, uint32_t *plan,
                                   uvmtier_make_resident_req_t *evict_req, uvmtier_make_resident_req_t *demote_req)
{
    simt::launch(dim3(3), dim3(kPmmThreads), [&] { k_pmm_slot_keys(n_slots, stamp, dram_stamp, touched, gpu_resident, cpu0, cpu1, e_weight, e_aux, d_key, d_weight, plan); });
    simt::launch(dim3(1), dim3(32), [&] { k_pmm_plan(plan, 0, capacity_chunks, dram_capacity_pages, demote); });
    simt::launch(dim3(1), dim3(kPmmSelectThreads), [&] { k_pmm_select_weighted(stamp, e_weight, e_aux, n_slots, plan + 8, plan + 16); });
    simt::launch(dim3(1), dim3(32), [&] { k_pmm_plan(plan, 1, capacity_chunks, dram_capacity_pages, demote); });
    if (demote)
        simt::launch(dim3(1), dim3(kPmmSelectThreads), [&] { k_pmm_select_weighted(d_key, d_weight, (const uint32_t *)NULL, n_slots, plan + 9, plan + 24); });
    simt::launch(dim3(1), dim3(32), [&] { k_pmm_plan(plan, 2, capacity_chunks, dram_capacity_pages, demote); });
    if (demote)
        simt::launch(dim3(2), dim3(kPmmThreads), [&] { k_pmm_emit_selected(d_key, n_slots, plan + 24, 0, plan + 10, demote_req, plan + 40); });
    simt::launch(dim3(2), dim3(kPmmThreads), [&] { k_pmm_emit_selected(stamp, n_slots, plan + 16, 1, plan + 10, evict_req, plan + 41); });
    return 0;
}
