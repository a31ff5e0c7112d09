/*
 * oracle_resident.c -- TEST INFRASTRUCTURE ONLY: CPU restatement of uvm_va_block_make_resident on the flat
 * per-slot state table (UVM = kernel-open/nvidia-uvm, uvm_va_block.c):
 *   M1  uvm_va_block_make_resident :5086-5114 = _copy :4877-4976 + _finish :5042-5084
 *   M1b block_populate_pages :3099-3182 (destination pages exist for everything that is copied or first touched)
 *   M2  block_copy_resident_pages :4660-4847, _mask :4460-4536, _from :4379-4452
 *   M4  block_make_resident_update_state :4990-5040, block_copy_set_first_touch_residency :4581-4619,
 *       break_read_duplication_in_region :4538-4579, block_make_resident_clear_evicted :4978-4988
 *   M6  uvm_va_block_evict_chunks :12999-13130 (request flag 4: page_mask = gpu resident pages, cause EVICTION)
 *   M1r uvm_va_block_make_resident_read_duplicate :5164-5332, block_prep_read_duplicate_mapping :5120-5162
 *       (request flag 8; state table with a read_duplicated mask)
 *
 * Processors of the table: one GPU ("this GPU", the destination and the only mapper), the CPU with two NUMA nodes
 * (0 = DRAM, 1 = CXL), and up to four other GPUs of the box that hold pages of the same va_blocks (residency only).
 * Harness policy as in oracle_service.c: residency is exclusive unless the caller read-duplicates (not with other
 * GPUs in the table), CPU destination pages are allocated on the node the request names (block_populate_pages_cpu's
 * first choice), this GPU copies from the CPU directly and from another GPU directly when the table says so
 * (peer_direct), else through CPU staging pages on node 0.
 *
 * Pinning: the reference holds no vectors for make_resident masks, so this file is checked against the
 * reference's own code: oracle/ref_resident_glue.c compiles uvm_va_block_make_resident, _copy, _finish,
 * block_copy_resident_pages{,_mask,_from}, block_make_resident_update_state, block_copy_set_first_touch_residency,
 * break_read_duplication_in_region and block_make_resident_clear_evicted VERBATIM from uvm_va_block.c against the
 * shims (oracle/_ref/libref_resident.so; page allocation, PTE writes and the copy push are stand-ins that record
 * masks), and tests/test_oracle_resident.py::test_oracle_matches_the_reference_make_resident requires identical
 * results, copy runs and block state on random requests.  M6 (flag 4) builds the request the way
 * uvm_va_block_evict_chunks does (:13080-13110) and then runs the pinned make_resident; the chunk walk above it is
 * out of scope.
 */
#include "oracle_common.h"
#include "ref_api.h"

uint32_t oracle_emit_runs(const or_mask_t *pages, const or_mask_t *prefetch, uint32_t slot, unsigned src, unsigned dst,
                          ref_copy_run_t *runs, uint32_t at, uint32_t cap);

static or_mask_t *node_mask(or_mask_t *cpu, int nid) { return &cpu[nid]; }

/* uvm_gpu_id_t of the GPU the table belongs to: only the ORDER of the GPU sources of a CPU destination depends on it
 * (closest first = ascending id, uvm_va_space.c:1879-1924), and only when the table names other GPUs. */
static uint32_t g_this_gpu_id = 1;
void oracle_make_resident_set_gpu_id(uint32_t gpu_id) { g_this_gpu_id = gpu_id; }

/* Processes the requests in order.  Returns the number of copy runs needed (only run_cap are written). */
uint32_t oracle_make_resident(uint32_t n_requests, const ref_make_resident_req_t *reqs, const uint64_t *page_masks,
                              const uint64_t *prefetch_masks, uint32_t n_slots, const ref_block_state_t *st,
                              ref_make_resident_result_t *results, ref_copy_run_t *runs, uint32_t run_cap)
{
    uint32_t q, nrun = 0;
    for (q = 0; q < n_requests; q++) {
        const ref_make_resident_req_t *rq = &reqs[q];
        ref_make_resident_result_t *res = &results[q];
        const uint32_t slot = rq->slot;
        const bool dst_gpu = rq->dest_tier == 0;
        const int nid = rq->dest_tier == 2 ? 1 : 0;                  /* make_resident.dest_nid */
        const bool rdup = (rq->flags & 8) != 0;                      /* uvm_va_block_make_resident_read_duplicate */
        or_mask_t gpu, cpu[2], pte[3], evicted, rd, pm, pf, copy, from[3], migrated, touch, now, unmap, *dst_res;
        or_mask_t peer[4], from_peer[4], staged;
        int i;
        uint32_t k = 0, n_unmapped = 0, np = st->n_peers, j;

        memset(res, 0, sizeof(*res));
        if (slot >= n_slots || rq->region_first > rq->region_outer || rq->region_outer > OR_PAGES || rq->dest_tier > 2 ||
            (rdup && (!st->read_duplicated || rq->cause == 4 || (rq->flags & 4) || np))) {   /* :5304 asserts !EVICTION */
            res->slot = 0xFFFFFFFFu;
            continue;
        }
        res->slot = slot;
        res->run_offset = nrun;
        or_load(&gpu, st->gpu_resident + 8 * (size_t)slot);
        or_load(&cpu[0], st->cpu_resident[0] + 8 * (size_t)slot);
        or_load(&cpu[1], st->cpu_resident[1] + 8 * (size_t)slot);
        for (i = 0; i < 3; i++) or_load(&pte[i], st->gpu_pte[i] + 8 * (size_t)slot);
        or_zero(&evicted);
        if (st->gpu_evicted) or_load(&evicted, st->gpu_evicted + 8 * (size_t)slot);
        or_zero(&rd);
        if (st->read_duplicated) or_load(&rd, st->read_duplicated + 8 * (size_t)slot);
        for (j = 0; j < 4; j++) { or_zero(&peer[j]); or_zero(&from_peer[j]); }
        for (j = 0; j < np; j++) or_load(&peer[j], st->peer_resident[j] + 8 * (size_t)slot);
        or_zero(&staged);

        /* page_mask clipped to the region (uvm_page_mask_region_clear_outside); NULL = whole region */
        or_zero(&pm);
        or_fill_range(&pm, rq->region_first, rq->region_outer);
        if (rq->flags & 1) {
            or_mask_t m;
            or_load(&m, page_masks + 8 * (size_t)rq->mask_row);
            or_and(&pm, &pm, &m);
        }
        if (rq->flags & 4)
            or_and(&pm, &pm, &gpu);                                   /* pages_to_evict :13078 */
        or_zero(&pf);
        if (rq->flags & 2)
            or_load(&pf, prefetch_masks + 8 * (size_t)rq->mask_row);

        dst_res = dst_gpu ? &gpu : node_mask(cpu, nid);
        if (rdup) {
            /* pages entering read duplication (:5222-5243): for every processor holding pages of
             * page_mask & ~read_duplicated, unmap everybody else and revoke its write access
             * (block_prep_read_duplicate_mapping :5120-5162).  The GPU is the only mapper of the table:
             * it loses its mappings of such CPU pages, and WRITE / ATOMIC on such pages of its own. */
            or_mask_t fresh, on_cpu, on_gpu, hit;
            or_andnot(&fresh, &pm, &rd);
            or_or(&on_cpu, &cpu[0], &cpu[1]);
            or_and(&on_cpu, &on_cpu, &fresh);
            or_and(&hit, &on_cpu, &pte[0]);
            n_unmapped += or_weight(&hit);
            for (i = 0; i < 3; i++) or_andnot(&pte[i], &pte[i], &on_cpu);
            or_and(&on_gpu, &gpu, &fresh);
            or_andnot(&pte[1], &pte[1], &on_gpu);
            or_andnot(&pte[2], &pte[2], &on_gpu);
        }
        else {
            /* make_resident_copy :4919-4926: unmap every mapped processor from the pages not resident on the
             * destination; :4931-4939: and every processor but the destination from the read-duplicated pages */
            or_mask_t gone;
            or_andnot(&gone, &pm, dst_res);
            if (!dst_gpu) {
                or_mask_t dup;
                or_and(&dup, &pm, &rd);
                or_or(&gone, &gone, &dup);
            }
            or_and(&unmap, &gone, &pte[0]);
            n_unmapped += or_weight(&unmap);
            for (i = 0; i < 3; i++) or_andnot(&pte[i], &pte[i], &gone);
        }

        /* block_copy_resident_pages: copy_page_mask = page_mask & ~resident(dst, dest_nid)  :4687-4692 */
        or_andnot(&copy, &pm, dst_res);
        for (i = 0; i < 3; i++) or_zero(&from[i]);
        if (dst_gpu) {
            /* first everything from the processors with direct copy access, closest first (:4707-4745, :4486-4497):
             * the other GPUs in ascending id, then the CPU, node after node */
            or_mask_t rest = copy, hop;
            for (j = 0; j < np; j++) {
                if (!((st->peer_direct >> j) & 1u)) continue;
                or_and(&from_peer[j], &rest, &peer[j]);
                or_andnot(&rest, &rest, &from_peer[j]);
                k += oracle_emit_runs(&from_peer[j], &pf, slot, 3u + j, rq->dest_tier, runs, nrun + k, run_cap);
            }
            or_and(&from[0], &rest, &cpu[0]);
            or_andnot(&rest, &rest, &from[0]);
            k += oracle_emit_runs(&from[0], &pf, slot, 0, rq->dest_tier, runs, nrun + k, run_cap);
            or_and(&from[1], &rest, &cpu[1]);
            or_andnot(&rest, &rest, &from[1]);
            k += oracle_emit_runs(&from[1], &pf, slot, 1, rq->dest_tier, runs, nrun + k, run_cap);
            /* then what only GPUs without direct access hold: staged through CPU pages (pages_staged :3140-3170, allocated
             * on node 0), copied there from those GPUs (:4754-4785), then from the CPU to the destination (:4803-4829) */
            for (j = 0; j < np; j++) {
                if ((st->peer_direct >> j) & 1u) continue;
                or_and(&hop, &rest, &peer[j]);
                or_andnot(&rest, &rest, &hop);
                or_or(&from_peer[j], &from_peer[j], &hop);
                or_or(&staged, &staged, &hop);
                k += oracle_emit_runs(&hop, &pf, slot, 3u + j, 1 /* DRAM */, runs, nrun + k, run_cap);
            }
            k += oracle_emit_runs(&staged, &pf, slot, 0, rq->dest_tier, runs, nrun + k, run_cap);
        }
        else {
            /* destination CPU: everything resident elsewhere is a source, closest first -- the CPU itself (the
             * other node; the destination node holds none of copy_page_mask), then the GPUs in ascending id (:4754-4785) */
            or_mask_t rest = copy;
            bool self_done = false;
            or_and(&from[nid ^ 1], &rest, &cpu[nid ^ 1]);
            or_andnot(&rest, &rest, &from[nid ^ 1]);
            k += oracle_emit_runs(&from[nid ^ 1], &pf, slot, (unsigned)(nid ^ 1), rq->dest_tier, runs, nrun + k, run_cap);
            for (j = 0; j <= np; j++) {
                if (!self_done && (j == np || st->peer_gpu_id[j] > g_this_gpu_id)) {
                    or_and(&from[2], &rest, &gpu);
                    or_andnot(&rest, &rest, &from[2]);
                    k += oracle_emit_runs(&from[2], &pf, slot, 2, rq->dest_tier, runs, nrun + k, run_cap);
                    self_done = true;
                }
                if (j < np) {
                    or_and(&from_peer[j], &rest, &peer[j]);
                    or_andnot(&rest, &rest, &from_peer[j]);
                    k += oracle_emit_runs(&from_peer[j], &pf, slot, 3u + j, rq->dest_tier, runs, nrun + k, run_cap);
                }
            }
        }
        or_or(&migrated, &from[0], &from[1]);
        or_or(&migrated, &migrated, &from[2]);
        for (j = 0; j < np; j++) or_or(&migrated, &migrated, &from_peer[j]);
        or_andnot(&touch, &copy, &migrated);                          /* resident nowhere: populated + zeroed */
        k += oracle_emit_runs(&touch, &pf, slot, 7, rq->dest_tier, runs, nrun + k, run_cap);
        res->run_count = (uint16_t)k;
        nrun += k;
        for (i = 0; i < 3; i++) res->migrated[i] = (uint16_t)or_weight(&from[i]);
        res->migrated[0] = (uint16_t)(res->migrated[0] + or_weight(&staged));      /* the second hop of the staged pages */
        for (j = 0; j < np; j++) res->migrated_peer[j] = (uint16_t)or_weight(&from_peer[j]);
        res->first_touch = (uint16_t)or_weight(&touch);
        res->unmapped = (uint16_t)n_unmapped;

        or_or(&now, &migrated, &touch);
        if (rdup) {
            /* :5259-5302: first touch and migrated pages become resident at the destination, the sources KEEP
             * theirs -- except the other CPU node (one CPU page table: the pages switch node, :5288-5297) --
             * and every migrated page is read-duplicated from now on */
            or_or(dst_res, dst_res, &now);
            if (!dst_gpu)
                or_andnot(&cpu[nid ^ 1], &cpu[nid ^ 1], &migrated);
            or_or(&rd, &rd, &migrated);
            if (dst_gpu)
                or_andnot(&evicted, &evicted, &migrated);             /* :5306-5307 */
            or_store(res->pages_changed_residency, &touch);           /* only first touch reports (:4614-4617) */
        }
        else {
            /* _finish: update_state (dst |= migrated; evicted), first touch (dst |= touch), then every other
             * residency loses region & page_mask and read duplication ends there  (:5061-5074, :4538-4579) */
            or_or(dst_res, dst_res, &now);
            or_andnot(&rd, &rd, &pm);
            for (j = 0; j < np; j++) or_andnot(&peer[j], &peer[j], &pm);
            if (dst_gpu) {
                or_andnot(&cpu[0], &cpu[0], &pm);
                or_andnot(&cpu[1], &cpu[1], &pm);
                or_andnot(&evicted, &evicted, &migrated);             /* :5038-5039 */
            }
            else {
                or_andnot(&cpu[nid ^ 1], &cpu[nid ^ 1], &pm);         /* uvm_va_block_cpu_clear_resident_all_chunks :672-700 */
                or_andnot(&gpu, &gpu, &pm);
                if (rq->cause == 4 && or_weight(&from[2]))
                    or_or(&evicted, &evicted, &migrated);             /* :5023-5037: every involved GPU gets copy_mask */
            }
            or_store(res->pages_changed_residency, &now);
        }
        or_store(res->pages_migrated, &migrated);

        or_store(st->gpu_resident + 8 * (size_t)slot, &gpu);
        or_store(st->cpu_resident[0] + 8 * (size_t)slot, &cpu[0]);
        or_store(st->cpu_resident[1] + 8 * (size_t)slot, &cpu[1]);
        for (i = 0; i < 3; i++) or_store(st->gpu_pte[i] + 8 * (size_t)slot, &pte[i]);
        if (st->gpu_evicted) or_store(st->gpu_evicted + 8 * (size_t)slot, &evicted);
        if (st->read_duplicated) or_store(st->read_duplicated + 8 * (size_t)slot, &rd);
        for (j = 0; j < np; j++) or_store(st->peer_resident[j] + 8 * (size_t)slot, &peer[j]);
    }
    return nrun;
}
