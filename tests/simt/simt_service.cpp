// simt_service.cpp -- runs the servicing-pass kernels (service_kernels.cuh, service_grp.cuh) under the CPU SIMT
// emulator with the same launch sequence as service_pass() in service.cu.  TEST INFRASTRUCTURE ONLY.
//
// lanes_per_block: unused (kept for the callers' signature).  heavy_min: slots whose runs hold at least this many
// entries are serviced by the warp-per-block kernel (0xffffffff = none, 0 = all), the others by a group of four lanes each.
#include "simt.h"

#include <vector>

#include "service_kernels.cuh"
#include "service_lists.cuh"
#include "service_grp.cuh"

using namespace uvmtier_k;

static int run_service(uint32_t n_batches, uint32_t B, const uvmtier_ordered_fault_t *ordered,
                       const uvmtier_fault_group_t *groups, uvmtier_fault_batch_t *batches,
                       const uvmtier_instance_ptr_entry_t *tbl, const uvmtier_va_layout_t *lay,
                       const uvmtier_block_state_t *state, uint32_t pf_enable, uint32_t pf_threshold,
                       uint32_t pf_min_faults, uint16_t *entry_flags, uvmtier_service_group_t *service,
                       uvmtier_copy_run_t *runs, uint32_t run_cap, uint32_t *counters, int lanes_per_block,
                       uint32_t heavy_min, uint32_t non_replayable, bool zero_outputs, const svc_thrash_params *thrash = NULL);

// with the thrashing hint inside the loop; tb = the ten table pointers of tests/thrashlib.py's HintState (counters last)
extern "C" int simt_service_fault_batches_thrashing(uint32_t n_batches, uint32_t B, const uvmtier_ordered_fault_t *ordered,
                                                    const uvmtier_fault_group_t *groups, uvmtier_fault_batch_t *batches,
                                                    const uvmtier_instance_ptr_entry_t *tbl, const uvmtier_va_layout_t *lay,
                                                    const uvmtier_block_state_t *state, uint32_t pf_enable, uint32_t pf_threshold,
                                                    uint32_t pf_min_faults, uint16_t *entry_flags, uvmtier_service_group_t *service,
                                                    uvmtier_copy_run_t *runs, uint32_t run_cap, uint32_t *counters, uint32_t heavy_min,
                                                    void *const *tb, const uvmtier_thrashing_hint_config_t *cfg,
                                                    const unsigned long long *batch_time)
{
    svc_thrash_params tp;
    tp.th.blocks = (uvmtier_thrashing_block_t *)tb[0];
    tp.th.page_word = (unsigned long long *)tb[1];
    tp.th.page_procs = (uint32_t *)tb[2];
    tp.th.thrashing_pages = (uint32_t *)tb[3];
    tp.th.page_throttle = (unsigned long long *)tb[4];
    tp.th.page_hint = (uint32_t *)tb[5];
    tp.th.page_throttled = (uint32_t *)tb[6];
    tp.th.page_deadline = (unsigned long long *)tb[7];
    tp.th.pinned_pages = (uint32_t *)tb[8];
    tp.cfg = cfg;
    tp.batch_time = batch_time;
    tp.tally = (uint32_t *)tb[9];
    return run_service(n_batches, B, ordered, groups, batches, tbl, lay, state, pf_enable, pf_threshold, pf_min_faults, entry_flags,
                       service, runs, run_cap, counters, 4, heavy_min, 0, true, &tp);
}

extern "C" int simt_service_fault_batches(uint32_t n_batches, uint32_t B, const uvmtier_ordered_fault_t *ordered,
                                          const uvmtier_fault_group_t *groups, uvmtier_fault_batch_t *batches,
                                          const uvmtier_instance_ptr_entry_t *tbl, const uvmtier_va_layout_t *lay,
                                          const uvmtier_block_state_t *state, uint32_t pf_enable, uint32_t pf_threshold,
                                          uint32_t pf_min_faults, uint16_t *entry_flags, uvmtier_service_group_t *service,
                                          uvmtier_copy_run_t *runs, uint32_t run_cap, uint32_t *counters, int lanes_per_block,
                                          uint32_t heavy_min)
{
    return run_service(n_batches, B, ordered, groups, batches, tbl, lay, state, pf_enable, pf_threshold, pf_min_faults, entry_flags,
                       service, runs, run_cap, counters, lanes_per_block, heavy_min, 0, true);
}

static int run_service(uint32_t n_batches, uint32_t B, const uvmtier_ordered_fault_t *ordered,
                       const uvmtier_fault_group_t *groups, uvmtier_fault_batch_t *batches,
                       const uvmtier_instance_ptr_entry_t *tbl, const uvmtier_va_layout_t *lay,
                       const uvmtier_block_state_t *state, uint32_t pf_enable, uint32_t pf_threshold,
                       uint32_t pf_min_faults, uint16_t *entry_flags, uvmtier_service_group_t *service,
                       uvmtier_copy_run_t *runs, uint32_t run_cap, uint32_t *counters, int lanes_per_block,
                       uint32_t heavy_min, uint32_t non_replayable, bool zero_outputs, const svc_thrash_params *thrash)
{
    svc_thrash_params tp;
    memset(&tp, 0, sizeof(tp));
    if (thrash)
        tp = *thrash;
    const uvmtier_thrashing_block_t *thr_blocks = tp.th.blocks;
    const uint32_t n_slots = lay->n_spaces * lay->blocks_per_space;
    const size_t total = (size_t)n_batches * B;
    service_kparams prm;
    memset(&prm, 0, sizeof(prm));
    prm.non_replayable = non_replayable;
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

    // (runs, entries) per slot packed in 64 bits, like the device array
    std::vector<unsigned long long> cw64(n_slots + 1, 0);
    const uint2 *cw = (const uint2 *)cw64.data();
    std::vector<uint32_t> offset(n_slots + 1, 0), cursor(4, 0), unmapped(n_batches, 0), list(total + 1, 0), big(n_slots + 2, 0),
        heavy(n_slots + 2, 0);
    std::vector<uint2> slotpos(total + 1);
    if (zero_outputs) {
        counters[0] = counters[1] = 0;
        memset(entry_flags, 0, total * sizeof(uint16_t));
    }
    simt::launch(dim3(2), dim3(kSvcThreads), [&] { k_service_count(prm, groups, batches, tbl, cw64.data(), unmapped.data(), slotpos.data()); });
    std::vector<uint32_t> hist(kLightBins, 0), light_rank(n_slots + 1, 0), light(n_slots + 1, 0);
    simt::launch(dim3((n_slots + 255) / 256), dim3(256), [&] { k_service_alloc(n_slots, cw, heavy_min, offset.data(), cursor.data(), hist.data(), light_rank.data(), thr_blocks); });
    simt::launch(dim3(2), dim3(kSvcThreads), [&] { k_service_fill((uint32_t)total, slotpos.data(), offset.data(), list.data()); });
    simt::launch(dim3(3), dim3(kSortThreads), [&] { k_service_sort_lists(n_slots, cw, offset.data(), list.data(), heavy_min, big.data(), heavy.data(), hist.data(), light_rank.data(), light.data(), thr_blocks); });
    simt::launch(dim3(2), dim3(kSortBigThreads), [&] { k_service_sort_big(cw, offset.data(), list.data(), big.data()); });
    (void)lanes_per_block;
    const uint32_t ggrid = 2;   // fewer CTAs than work items: the warps keep pulling from the counter
    simt::launch(dim3(ggrid), dim3(kGrpThreads), [&] { k_service_blocks_grp<kGrpLanes, 1>(prm, ordered, groups, batches, bs, cw, offset.data(), list.data(), light.data(), cursor.data() + 1, entry_flags, service, runs, counters, cursor.data() + 2); });
    simt::launch(dim3(3), dim3(kSvcThreads), [&] { k_service_blocks(prm, ordered, groups, batches, bs, cw, offset.data(), list.data(), heavy.data(), entry_flags, service, runs, counters, tp); });
    simt::launch(dim3((n_batches + 127) / 128), dim3(128), [&] { k_service_unmapped(prm, ordered, groups, batches, tbl, unmapped.data(), entry_flags, service); });
    return 0;
}


// ---------------------------------------------------------------- non-replayable faults
#include "fault_kernels.cuh"

extern "C" int simt_service_non_replayable_faults(const uvmtier_fault_packet_t *shadow, uint32_t n, const uvmtier_fault_gpu_params_t *gp,
                                                  const uvmtier_instance_ptr_entry_t *tbl, uint32_t tbl_n, const uvmtier_va_layout_t *lay,
                                                  const uvmtier_block_state_t *state, uint32_t pf_enable, uint32_t pf_threshold,
                                                  uint32_t pf_min_faults, uvmtier_fault_buffer_entry_t *entries, uint16_t *entry_flags,
                                                  uvmtier_service_group_t *service, uvmtier_copy_run_t *runs, uint32_t run_cap,
                                                  uint32_t *counters, uint32_t heavy_min)
{
    const uint32_t B = gp->max_batch_size, nb = (n + B - 1) / B;
    const size_t total = (size_t)nb * B;
    std::vector<uvmtier_ordered_fault_t> ordered(total);
    std::vector<uvmtier_fault_group_t> groups(total);
    std::vector<uvmtier_fault_batch_t> batches(nb);
    counters[0] = counters[1] = 0;
    memset(entry_flags, 0, total * sizeof(uint16_t));
    memset(service, 0, total * sizeof(*service));
    simt::launch(dim3(2), dim3(256), [&] { k_nonreplayable_prepare(shadow, n, B, gp->utlb_per_gpc_count, gp->va_bits, tbl, tbl_n, ordered.data(), groups.data(), batches.data(), entry_flags, entries); });
    return run_service(nb, B, ordered.data(), groups.data(), batches.data(), tbl, lay, state, pf_enable, pf_threshold, pf_min_faults,
                       entry_flags, service, runs, run_cap, counters, 4, heavy_min, 1, false);
}
