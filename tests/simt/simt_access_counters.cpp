// simt_access_counters.cpp -- runs the access-counter kernels (access_counter_kernels.cuh) under the CPU SIMT emulator.
// TEST INFRASTRUCTURE ONLY.
#include "simt.h"

#include <vector>

#define UVMTIER_AC_PREPROCESS
#define UVMTIER_AC_SERVICE
#include "access_counter_kernels.cuh"
#include "service_lists.cuh"

using namespace uvmtier_k;

extern "C" int simt_access_counters_fetch_preprocess(uvmtier_access_counter_packet_t *ring, uint32_t ring_entries, uint32_t get,
                                                     uint32_t put, uint32_t B, uint32_t va_bits,
                                                     const uvmtier_instance_ptr_entry_t *tbl, uint32_t tbl_n, int clear_valid,
                                                     uvmtier_ordered_notification_t *ordered, uvmtier_fault_group_t *groups,
                                                     uvmtier_fault_batch_t *batches, uvmtier_access_counter_entry_t *entries)
{
    const uint32_t n = put >= get ? put - get : ring_entries - get + put;
    if (!n)
        return 0;
    uint32_t threads = 32, idx_bits = 5;
    while (threads < B) { threads <<= 1; idx_bits++; }
    ac_kparams prm;
    prm.B = B; prm.va_bits = va_bits; prm.ring_entries = ring_entries; prm.get = get; prm.n_entries = n;
    prm.tbl_n = tbl_n; prm.idx_bits = idx_bits; prm.clear_valid = clear_valid != 0;
    simt::launch(dim3((n + B - 1) / B), dim3(threads),
                 [&] { k_access_counters_fetch_preprocess(prm, ring, tbl, ordered, groups, batches, entries); });
    return 0;
}

extern "C" int simt_access_counters_service(uint32_t n_batches, uint32_t B, const uvmtier_ordered_notification_t *ordered,
                                            const uvmtier_fault_group_t *groups, uvmtier_fault_batch_t *batches,
                                            const uvmtier_instance_ptr_entry_t *tbl, const uvmtier_va_layout_t *lay,
                                            const uvmtier_block_state_t *state, uint32_t pf_enable, uint32_t pf_threshold,
                                            uint32_t pf_min_faults, int granularity_2m, int cpu_chunk_2m,
                                            uvmtier_service_group_t *service, uvmtier_copy_run_t *runs, uint32_t run_cap,
                                            uint32_t *counters, uint32_t heavy_min)
{
    const uint32_t n_slots = lay->n_spaces * lay->blocks_per_space;
    const size_t total = (size_t)n_batches * B;
    service_kparams prm;
    memset(&prm, 0, sizeof(prm));
    prm.va_base = lay->va_base; prm.B = B; prm.n_batches = n_batches; prm.blocks_per_space = lay->blocks_per_space;
    prm.n_spaces = lay->n_spaces; prm.gpu_id = lay->gpu_id; prm.n_slots = n_slots; prm.pf_enable = pf_enable;
    prm.pf_threshold = pf_threshold; prm.pf_min_faults = pf_min_faults; prm.run_cap = run_cap; prm.B_shift = svc_log2_or_none(B);
    block_state_ptrs bs;
    bs.gpu_resident = (uint32_t *)state->gpu_resident;
    bs.cpu_resident[0] = (uint32_t *)state->cpu_resident[0]; bs.cpu_resident[1] = (uint32_t *)state->cpu_resident[1];
    bs.gpu_pte[0] = (uint32_t *)state->gpu_pte[0]; bs.gpu_pte[1] = (uint32_t *)state->gpu_pte[1]; bs.gpu_pte[2] = (uint32_t *)state->gpu_pte[2];
    bs.prefetch_info = state->prefetch_info; bs.logical_prot = state->logical_prot; bs.preferred_location = state->preferred_location;
    bs.gpu_evicted = (uint32_t *)state->gpu_evicted;
    bs.read_duplicated = (uint32_t *)state->read_duplicated;

    std::vector<unsigned long long> cw64(n_slots + 1, 0);
    const uint2 *cw = (const uint2 *)cw64.data();
    std::vector<uint32_t> offset(n_slots + 1, 0), cursor(4, 0), unmapped(n_batches, 0), list(total + 1, 0), big(n_slots + 2, 0),
        heavy(n_slots + 2, 0), hist(kLightBins, 0), light_rank(n_slots + 1, 0), light(n_slots + 1, 0);
    std::vector<uint2> slotpos(total + 1);
    counters[0] = counters[1] = 0;
    simt::launch(dim3(2), dim3(kSvcThreads), [&] { k_service_count(prm, groups, batches, tbl, cw64.data(), unmapped.data(), slotpos.data()); });
    simt::launch(dim3((n_slots + 255) / 256), dim3(256), [&] { k_service_alloc(n_slots, cw, heavy_min, offset.data(), cursor.data(), hist.data(), light_rank.data()); });
    simt::launch(dim3(2), dim3(kSvcThreads), [&] { k_service_fill((uint32_t)total, slotpos.data(), offset.data(), list.data()); });
    simt::launch(dim3(3), dim3(kSortThreads), [&] { k_service_sort_lists(n_slots, cw, offset.data(), list.data(), heavy_min, big.data(), heavy.data(), hist.data(), light_rank.data(), light.data()); });
    simt::launch(dim3(2), dim3(kSortBigThreads), [&] { k_service_sort_big(cw, offset.data(), list.data(), big.data()); });
    ac_service_cfg cfg;
    cfg.granularity_2m = granularity_2m != 0;
    cfg.cpu_chunk_2m = cpu_chunk_2m != 0;
    simt::launch(dim3(2), dim3(kSvcThreads), [&] { k_access_counters_blocks(prm, cfg, ordered, groups, bs, cw, offset.data(), list.data(), light.data(), cursor.data() + 1, service, runs, counters); });
    simt::launch(dim3(2), dim3(kSvcThreads), [&] { k_access_counters_blocks(prm, cfg, ordered, groups, bs, cw, offset.data(), list.data(), heavy.data() + 1, heavy.data(), service, runs, counters); });
    return 0;
}
