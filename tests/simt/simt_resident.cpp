// simt_resident.cpp -- runs k_make_resident (resident_kernels.cuh) under the CPU SIMT emulator.
// TEST INFRASTRUCTURE ONLY.
#include "simt.h"

#include "resident_kernels.cuh"

using namespace uvmtier_k;

extern "C" int simt_make_resident(uint32_t n_requests, const uvmtier_make_resident_req_t *reqs, const uint32_t *page_masks,
                                  const uint32_t *prefetch_masks, uint32_t n_slots, const uvmtier_block_state_t *state,
                                  uvmtier_make_resident_result_t *results, uvmtier_copy_run_t *runs, uint32_t run_cap,
                                  uint32_t *counters, uint32_t gpu_id)
{
    resident_kparams prm;
    prm.n_requests = n_requests; prm.n_slots = n_slots; prm.run_cap = run_cap; prm.gpu_id = gpu_id; prm.n_requests_dev = NULL;
    peer_state_ptrs peers;
    memset(&peers, 0, sizeof(peers));
    peers.n_peers = state->n_peers; peers.direct = state->peer_direct;
    for (uint32_t k = 0; k < state->n_peers; k++) {
        peers.resident[k] = (uint32_t *)state->peer_resident[k];
        peers.gpu_id[k] = state->peer_gpu_id[k];
    }
    block_state_ptrs bs;
    bs.gpu_resident = (uint32_t *)state->gpu_resident;
    bs.cpu_resident[0] = (uint32_t *)state->cpu_resident[0]; bs.cpu_resident[1] = (uint32_t *)state->cpu_resident[1];
    bs.gpu_pte[0] = (uint32_t *)state->gpu_pte[0]; bs.gpu_pte[1] = (uint32_t *)state->gpu_pte[1]; bs.gpu_pte[2] = (uint32_t *)state->gpu_pte[2];
    bs.prefetch_info = state->prefetch_info; bs.logical_prot = state->logical_prot; bs.preferred_location = state->preferred_location;
    bs.gpu_evicted = (uint32_t *)state->gpu_evicted;
    bs.read_duplicated = (uint32_t *)state->read_duplicated;
    simt::launch(dim3(3), dim3(kResThreads), [&] { k_make_resident(prm, reqs, page_masks, prefetch_masks, bs, peers, results, runs, counters); });
    return 0;
}
