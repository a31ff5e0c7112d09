/*
 * oracle_pmm.c -- TEST INFRASTRUCTURE ONLY: CPU restatement of the reference's root-chunk eviction order
 * (kernel-open/nvidia-uvm/uvm_pmm_gpu.c): the two lists root_chunks.va_block_used / va_block_unused, whose every
 * move goes to the tail (root_chunk_update_eviction_list :1430-1448, chunk_update_lists_locked :627-651), and
 * pick_root_chunk_to_evict (:1460-1500: head of unused, else head of used), restated as one stamp per va_block:
 *   0 = no root chunk, else (bit 63 if on va_block_used) | sequence number of the last move.
 *
 * Pinning: oracle_pmm_replay is compared with the reference's own functions compiled verbatim
 * (oracle/_ref/libref_pmm.so, ref_pmm_glue.c) on random event sequences (tests/test_oracle_pmm.py).
 */
#include <stdint.h>
#include <stdlib.h>

#include "ref_api.h"

#define PMM_USED (1ull << 63)
enum { EV_ALLOC = 0, EV_USED = 1, EV_UNUSED = 2, EV_PICK = 3 };

/* Same contract as ref_pmm_replay (ref_pmm_glue.c). */
uint32_t oracle_pmm_replay(uint32_t n_slots, uint32_t n_events, const uint32_t *events, uint32_t *victims)
{
    uint64_t *stamp = (uint64_t *)calloc(n_slots, sizeof(uint64_t)), clock = 0;
    uint32_t i, s, nv = 0;
    for (i = 0; i < n_events; i++) {
        const uint32_t kind = events[2 * i], slot = events[2 * i + 1];
        if (kind == EV_ALLOC)
            stamp[slot] = PMM_USED | ++clock;                   /* chunk_update_lists_locked :639-643 */
        else if (kind == EV_USED) {
            if (stamp[slot]) stamp[slot] = PMM_USED | ++clock;  /* uvm_pmm_gpu_mark_root_chunk_used :1450-1453 */
        }
        else if (kind == EV_UNUSED) {
            if (stamp[slot]) stamp[slot] = ++clock;             /* uvm_pmm_gpu_mark_root_chunk_unused :1455-1458 */
        }
        else {
            uint32_t best = 0xFFFFFFFFu;
            for (s = 0; s < n_slots; s++)                       /* head of unused, else head of used :1484-1490 */
                if (stamp[s] && (best == 0xFFFFFFFFu || stamp[s] < stamp[best]))
                    best = s;
            victims[nv++] = best;
            if (best != 0xFFFFFFFFu)
                stamp[best] = 0;
        }
    }
    free(stamp);
    return nv;
}

/* The list moves of one servicing pass: record (b, g) at b * B + g; serviced runs in (batch, run) order. */
void oracle_pmm_note_service(uint64_t *stamp, uint32_t n_slots, uint64_t clock, uint32_t n_batches, uint32_t B,
                             const ref_fault_batch_t *batch, const ref_service_group_t *svc_by_position)
{
    uint32_t b, g;
    for (b = 0; b < n_batches; b++)
        for (g = 0; g < batch[b].num_block_groups; g++) {
            const ref_service_group_t *r = &svc_by_position[(size_t)b * B + g];
            if (r->slot < n_slots && r->page_fault_count)
                stamp[r->slot] = PMM_USED | (clock + (uint64_t)b * B + g + 1);
        }
}

/* The list moves of one make_resident call; st = the state table after the call. */
void oracle_pmm_note_make_resident(uint64_t *stamp, uint32_t n_slots, uint64_t clock, uint32_t n_requests,
                                   const ref_make_resident_req_t *reqs, const ref_make_resident_result_t *results,
                                   const ref_block_state_t *st)
{
    uint32_t q, w;
    for (q = 0; q < n_requests; q++) {
        const ref_make_resident_req_t *rq = &reqs[q];
        uint64_t any = 0;
        if (results[q].slot == 0xFFFFFFFFu || rq->slot >= n_slots)
            continue;
        for (w = 0; w < 8; w++) any |= st->gpu_resident[8 * (size_t)rq->slot + w];
        if (rq->dest_tier == 0) {
            if (any) stamp[rq->slot] = PMM_USED | (clock + q + 1);             /* block_mark_memory_used */
        }
        else if (!any) {
            if (rq->flags & 4) stamp[rq->slot] = 0;                              /* evicted and freed */
            else if (stamp[rq->slot] && results[q].migrated[2]) stamp[rq->slot] = clock + q + 1;
        }
    }
}

static int cmp_u64(const void *a, const void *b)
{
    const uint64_t x = *(const uint64_t *)a, y = *(const uint64_t *)b;
    return x < y ? -1 : x > y;
}

/* pick_root_chunk_to_evict k times: the victims' slots in eviction order; returns how many. */
uint32_t oracle_pmm_pick(const uint64_t *stamp, uint32_t n_slots, uint32_t k, uint32_t *victims)
{
    uint64_t *key = (uint64_t *)malloc((size_t)n_slots * 2 * sizeof(uint64_t));
    uint32_t s, n = 0;
    for (s = 0; s < n_slots; s++)
        if (stamp[s]) { key[2 * n] = stamp[s]; key[2 * n + 1] = s; n++; }
    qsort(key, n, 2 * sizeof(uint64_t), cmp_u64);
    if (k > n) k = n;
    for (s = 0; s < k; s++) victims[s] = (uint32_t)key[2 * s + 1];
    free(key);
    return k;
}
