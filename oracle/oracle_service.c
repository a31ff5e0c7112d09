/*
 * oracle_service.c -- TEST INFRASTRUCTURE ONLY: CPU restatement of the per-va_block servicing of a
 * sorted fault batch (UVM = kernel-open/nvidia-uvm):
 *   F4  service_fault_batch / _dispatch / _block_locked      uvm_gpu_replayable_faults.c:1375-1597,1946-2035,2232-2377
 *       check_fault_access_permissions                        :1293-1360
 *   P4  uvm_va_block_get_prefetch_hint                        uvm_va_block.c:11828-11881 (calls oracle_prefetch.c)
 *   M1-M4 make_resident mask algebra                          uvm_va_block.c:4660-4847 (copy lists), :4990-5084 (state)
 *       mapping after service                                 uvm_va_block.c:11019-11066, 12063-12080
 *
 * Harness policy, stated once here and in DESIGN.md: one faulting GPU per state table,
 * no read duplication, no thrashing, no accessed_by, every range migratable, no range groups; the
 * destination of every serviced fault is the faulting GPU (block_select_processor_residency falls through
 * to `return processor_id`, uvm_va_block.c:11560-11712, when the page is not resident on it); CPU memory
 * has two NUMA nodes (0 = DRAM, 1 = CXL); both tiers and HBM back a va_block contiguously.
 *
 * Pinning: no known-answer vectors exist in the reference for this stage.  The ENTRY LOOP (duplicates,
 * permissions, authorised-already, new_residency / region, batch counters, fatal / invalid-prefetch marks) is
 * pinned against the reference's own service_fault_batch_block_locked, check_fault_access_permissions,
 * fault_entry_duplicate_flags, mark_fault_* and service_fault_batch_fatal_notify compiled verbatim
 * (oracle/_ref/libref_service.so, ref_service_glue.c; tests/test_oracle_service.py -k reference) -- which found
 * and fixed one divergence (duplicates of faults outside every va_range are counted too, :1692).  The BACK HALF
 * (oracle_service_block_apply) is three steps: the prefetch hint = oracle_prefetch.c, pinned by the reference's
 * known-answer vectors and by libref_prefetch.so; residency = oracle_make_resident, pinned by the reference's
 * uvm_va_block_make_resident compiled verbatim (libref_resident.so, see oracle_resident.c); and the mapping of the
 * serviced pages on the faulting GPU (uvm_va_block.c:11019-11066, 12063-12080), which is the one step that is a
 * restatement only: its callees (block_map_gpu_to, PTE batches) write page tables and are out of scope.
 * Batches are applied in trace order, exactly like uvm_parent_gpu_service_replayable_faults (:2906-2960).
 */
#include "oracle_common.h"
#include "oracle_thrashing.h"
#include "ref_api.h"

enum { ACC_PREFETCH = 0, ACC_READ, ACC_WRITE, ACC_ATOMIC_WEAK, ACC_ATOMIC_STRONG, ACC_COUNT };
enum { PROT_NONE = 0, PROT_RO = 1, PROT_RW = 2, PROT_RWA = 3 };
enum { REASON_INVALID_ADDRESS = 1, REASON_INVALID_PERMISSIONS = 2 };   /* UvmEventFatalReason, uvm_types.h:514-529 */
enum { CANCEL_ALL = 0, CANCEL_WRITE_AND_ATOMIC = 1 };                  /* uvm_hal_types.h:262-271 */

uint32_t oracle_make_resident(uint32_t n_requests, const ref_make_resident_req_t *reqs, const uint64_t *page_masks,
                              const uint64_t *prefetch_masks, uint32_t n_slots, const ref_block_state_t *st,
                              ref_make_resident_result_t *results, ref_copy_run_t *runs, uint32_t run_cap);
void oracle_prefetch_get_hint_va_block(size_t n, const ref_prefetch_block_t *blk, const uint64_t *resident,
                                       const uint64_t *faulted, const uint64_t *thrashing,
                                       const uint64_t *cpu_read_mapped, ref_prefetch_info_t *info,
                                       uint64_t *out_mask, uint32_t *out_residency);

/* uvm_fault_access_type_to_prot, uvm_hal_types.h:522-537 */
static unsigned prot_of(unsigned acc)
{
    if (acc == ACC_ATOMIC_STRONG) return PROT_RWA;
    if (acc == ACC_ATOMIC_WEAK || acc == ACC_WRITE) return PROT_RW;
    return PROT_RO;
}

/* uvm_va_block_check_logical_permissions for a GPU fault on a migratable managed range
 * (uvm_va_block.c:12361-12416 with uvm_enable_builtin_tests set, as in the reference's own test runs) */
static bool logical_ok(unsigned logical_prot, unsigned acc) { return prot_of(acc) <= logical_prot; }

/* The back half's two callees.  Default: the restatements in this directory.  oracle_service_set_back_half() swaps in
 * functions of the same contract -- in practice ref_prefetch_get_hint_va_block (oracle/_ref/libref_prefetch.so: the
 * reference's uvm_perf_prefetch.c compiled unmodified) and ref_make_resident (libref_resident.so: the reference's
 * uvm_va_block_make_resident compiled verbatim), so that the whole CPU arm runs the reference's own code. */
typedef void (*prefetch_hint_fn)(size_t, const ref_prefetch_block_t *, const uint64_t *, const uint64_t *, const uint64_t *,
                                 const uint64_t *, ref_prefetch_info_t *, uint64_t *, uint32_t *);
/* ref_make_resident also takes the id of the GPU the table belongs to (the restatement reads it from nowhere: one GPU) */
typedef uint32_t (*ref_make_resident_fn)(uint32_t, const ref_make_resident_req_t *, const uint64_t *, const uint64_t *, uint32_t,
                                         uint32_t, const ref_block_state_t *, ref_make_resident_result_t *, ref_copy_run_t *, uint32_t);
static prefetch_hint_fn g_prefetch_hint = oracle_prefetch_get_hint_va_block;
static ref_make_resident_fn g_ref_make_resident;
void oracle_service_set_back_half(void *prefetch_hint, void *ref_make_resident)
{
    g_prefetch_hint = prefetch_hint ? (prefetch_hint_fn)prefetch_hint : oracle_prefetch_get_hint_va_block;
    g_ref_make_resident = (ref_make_resident_fn)ref_make_resident;
}

/* Thrashing hints inside the fault loop (uvm_gpu_replayable_faults.c:1524-1545): off by default (the harness policy of the
 * bench: "no thrashing"); oracle_service_set_thrashing() attaches the thrashing restatement's tables and one time stamp
 * per batch (NV_GETTIME() when the batch is serviced).  A fault that reaches the hint:
 *   THROTTLE -> mark_fault_throttled (:1537): flagged, not serviced
 *   PIN      -> the page joins thrashing_pin_mask; uvm_va_block_select_residency takes the pin's residency
 *               (uvm_va_block.c:11624-11628).  On this GPU: serviced as ever.  Elsewhere: flagged and left to the caller --
 *               the servicing pass moves pages to the faulting GPU only (stated in include/uvmtier.h).
 * and the prefetcher leaves the va_block's thrashing pages alone (uvm_perf_prefetch.c:378-381, 407). */
static const thrash_cfg_t *g_th_cfg;
static const thrash_tables_t *g_th;
static const uint64_t *g_batch_time;
void oracle_service_set_thrashing(const thrash_cfg_t *cfg, const thrash_tables_t *th, const uint64_t *batch_time_ns)
{
    g_th_cfg = cfg; g_th = cfg ? th : NULL; g_batch_time = batch_time_ns;
}

static void mask_ptr_load(or_mask_t *m, uint64_t *base, uint32_t slot) { or_load(m, base + 8 * (size_t)slot); }
static void mask_ptr_store(uint64_t *base, uint32_t slot, const or_mask_t *m) { or_store(base + 8 * (size_t)slot, m); }

/* Emits the copy runs of `pages` from source `src`: consecutive pages with the same cause form one run
 * (block_copy_resident_pages_between, uvm_va_block.c:4143-4335, both sides block-contiguous). */
uint32_t oracle_emit_runs(const or_mask_t *pages, const or_mask_t *prefetch, uint32_t slot, unsigned src, unsigned dst,
                          ref_copy_run_t *runs, uint32_t at, uint32_t cap)
{
    unsigned p = 0;
    uint32_t n = 0;
    while (p < OR_PAGES) {
        unsigned q, cause;
        if (!or_test(pages, p)) { p++; continue; }
        cause = or_test(prefetch, p);
        for (q = p + 1; q < OR_PAGES && or_test(pages, q) && (unsigned)or_test(prefetch, q) == cause; q++) {}
        if (at + n < cap) {
            ref_copy_run_t *r = &runs[at + n];
            r->slot = slot; r->first_page = (uint16_t)p; r->num_pages = q - p; r->src = src; r->prefetch = cause; r->dst = dst;
        }
        n++;
        p = q;
    }
    return n;
}

/*
 * The back half of servicing one va_block run: uvm_va_block_service_locked (uvm_va_block.c:12307-12359) with what
 * service_fault_batch_block_locked built -- the new_residency mask of the faulting GPU and the faulted region.
 * Prefetch hint, make_resident (copy list, residency), mapping.  Fills the rest of *rec, appends copy runs at
 * runs[*nrun].  Also called by oracle/ref_service_glue.c, where the reference's own entry loop builds the mask.
 */
void oracle_service_block_apply(const ref_va_layout_t *lay, const ref_block_state_t *st, uint32_t slot, uint64_t block_start,
                                const uint64_t *new_residency, unsigned region_first, unsigned region_outer,
                                ref_service_group_t *rec, ref_copy_run_t *runs, uint32_t *nrun_io, uint32_t run_cap)
{
    or_mask_t pf_mask, cpu0, cpu1, any_cpu, newres, gpu_res, pte[3];
    ref_prefetch_block_t pb;
    uint32_t hint, nrun = *nrun_io;
    unsigned rf = region_first, ro = region_outer, p, i;
    const unsigned logical = st->logical_prot[slot];

    or_load(&newres, new_residency);
    mask_ptr_load(&gpu_res, st->gpu_resident, slot);
    for (i = 0; i < 3; i++) mask_ptr_load(&pte[i], st->gpu_pte[i], slot);

    /* get_prefetch_hint :11828-11881: exactly one destination processor */
    memset(&pb, 0, sizeof(pb));
    pb.start = block_start; pb.big_page_size = 65536; pb.num_pages = 512;
    pb.faulted_first = (uint16_t)rf; pb.faulted_outer = (uint16_t)ro;
    pb.new_residency = (uint16_t)lay->gpu_id; pb.preferred_location = st->preferred_location[slot];
    mask_ptr_load(&cpu0, st->cpu_resident[0], slot);
    mask_ptr_load(&cpu1, st->cpu_resident[1], slot);
    or_or(&any_cpu, &cpu0, &cpu1);
    pb.flags = REF_PF_HAS_RESIDENT_MASK | REF_PF_GPU_VA_SPACE_REGISTERED |
               ((or_empty(&gpu_res) && or_empty(&any_cpu)) ? REF_PF_BLOCK_RESIDENT_EMPTY : 0) | (g_th ? REF_PF_HAS_THRASHING : 0);
    g_prefetch_hint(1, &pb, gpu_res.w, newres.w, g_th ? (const uint64_t *)(g_th->thrashing_pages + (size_t)slot * 16) : NULL, NULL,
                    &st->prefetch_info[slot], pf_mask.w, &hint);
    rec->hint_residency = (uint16_t)hint;
    if (hint != 257) {
        or_or(&newres, &newres, &pf_mask);
        for (p = 0; p < OR_WORDS; p++) if (newres.w[p]) { rf = 64 * p + (unsigned)__builtin_ctzll(newres.w[p]); break; }
        for (p = OR_WORDS; p-- > 0;) if (newres.w[p]) { ro = 64 * p + 64 - (unsigned)__builtin_clzll(newres.w[p]); break; }
        rec->prefetch_count = (uint16_t)or_weight(&pf_mask);
    }
    else {
        or_zero(&pf_mask);
    }
    rec->region_first = (uint16_t)rf; rec->region_outer = (uint16_t)ro;
    or_store(rec->new_residency, &newres);

    /* uvm_va_block_make_resident(dest = the faulting GPU, region, page_mask = new_residency, prefetch mask, cause =
     * REPLAYABLE_FAULT): the restatement in oracle_resident.c, which is compared with the reference's own
     * uvm_va_block_make_resident (oracle/_ref/libref_resident.so).  Copy list, residency, eviction marks, unmapping. */
    {
        ref_make_resident_req_t rq;
        ref_make_resident_result_t res;
        memset(&rq, 0, sizeof(rq));
        rq.slot = slot; rq.region_first = (uint16_t)rf; rq.region_outer = (uint16_t)ro; rq.dest_tier = 0; rq.cause = 0;
        rq.flags = 1 | 2; rq.mask_row = 0;
        if (g_ref_make_resident)
            g_ref_make_resident(1, &rq, newres.w, pf_mask.w, 0xFFFFFFFFu, lay->gpu_id, st, &res,
                                runs ? runs + (nrun < run_cap ? nrun : run_cap) : NULL, nrun < run_cap ? run_cap - nrun : 0);
        else
            oracle_make_resident(1, &rq, newres.w, pf_mask.w, 0xFFFFFFFFu, st, &res, runs ? runs + (nrun < run_cap ? nrun : run_cap) : NULL,
                                 nrun < run_cap ? run_cap - nrun : 0);
        rec->migrated[0] = res.migrated[0];
        rec->migrated[1] = res.migrated[1];
        rec->first_touch = res.first_touch;
        rec->run_count = res.run_count;
        nrun += res.run_count;
        for (i = 0; i < 3; i++) mask_ptr_load(&pte[i], st->gpu_pte[i], slot);      /* after the unmapping */
    }

    /* service_finish :12063-12302: map the faulting GPU.  compute_new_permission (:11019-11066): the page is now
     * resident on the faulting GPU, which has native atomics on its own memory, and no other faultable processor
     * holds an atomic mapping of a page that just moved, so a read-write VMA maps READ_WRITE_ATOMIC whatever the
     * access type; a read-only VMA maps READ_ONLY. */
    {
        const unsigned new_prot = logical == PROT_RWA ? PROT_RWA : PROT_RO;
        for (i = 0; i < new_prot; i++) {
            or_or(&pte[i], &pte[i], &newres);
            mask_ptr_store(st->gpu_pte[i], slot, &pte[i]);
        }
    }
    *nrun_io = nrun;
}

/*
 * Services every batch in order.  ordered/groups/batch: the product-layout outputs of the preprocess
 * stage, batch b at offset b * max_batch_size.  Outputs:
 *   entry_flags[b*B + s]  REF_SF_* of sorted entry s
 *   svc[]                 one record per serviced-or-examined run, in (batch, group) order; returns count
 *   runs[]                copy runs, in record order (run_offset/run_count index into it)
 *   batch[].num_duplicate_faults / num_invalid_prefetch_faults are accumulated
 * *n_runs receives the number of runs needed (may exceed run_cap: then only run_cap were written).
 */
uint32_t oracle_service_fault_batches(uint32_t n_batches, uint32_t max_batch_size, const ref_ordered_fault_t *ordered,
                                      const ref_fault_group_t *groups, ref_fault_batch_t *batch,
                                      const ref_instance_ptr_entry_t *tbl, const ref_va_layout_t *lay,
                                      const ref_block_state_t *st, unsigned pf_enable, unsigned pf_threshold,
                                      unsigned pf_min_faults, uint16_t *entry_flags, ref_service_group_t *svc,
                                      ref_copy_run_t *runs, uint32_t run_cap, uint32_t *n_runs)
{
    uint32_t b, nsvc = 0, nrun = 0;
    extern int oracle_prefetch_init(unsigned, unsigned, unsigned);
    oracle_prefetch_init(pf_enable, pf_threshold, pf_min_faults);

    for (b = 0; b < n_batches; b++) {
        const size_t base = (size_t)b * max_batch_size;
        uint32_t g;
        if (batch[b].status != 0)
            continue;                                   /* batch is refetched, nothing serviced (:2946) */
        for (g = 0; g < batch[b].num_block_groups; g++) {
            const ref_fault_group_t *grp = &groups[base + g];
            const uint32_t first = grp->first_fault_index, count = grp->block_faults;
            const uint32_t va_space = tbl[grp->ctx].va_space;
            ref_service_group_t *rec;
            uint64_t off;
            uint32_t slot = 0xFFFFFFFFu, i, page_faults = 0;
            unsigned first_page = OR_PAGES, last_page = 0, logical;
            or_mask_t newres, gpu_res, pte[3];
            bool prev_fatal = false, prev_invalid_prefetch = false;

            for (i = 0; i < count; i++)
                entry_flags[base + first + i] = 0;
            if (grp->flags & 1)
                continue;                               /* fatal fault type: skipped by service_fault_batch (:2294-2303) */

            rec = &svc[nsvc++];
            memset(rec, 0, sizeof(*rec));
            rec->batch = (uint16_t)b; rec->group = g; rec->hint_residency = 257; rec->run_offset = nrun;
            off = grp->block_start - lay->va_base;
            if (grp->block_start >= lay->va_base && (off >> 21) < lay->blocks_per_space && va_space < lay->n_spaces)
                slot = va_space * lay->blocks_per_space + (uint32_t)(off >> 21);
            rec->slot = slot;

            if (slot == 0xFFFFFFFFu) {
                /* no va_range: service_fault_batch_fatal_notify one entry at a time (:1656-1698, :2022-2031) */
                for (i = 0; i < count; i++) {
                    const ref_ordered_fault_t *e = &ordered[base + first + i];
                    const unsigned acc = e->fault_access_type & 7;
                    uint16_t fl = 0;
                    bool dup = false;
                    if (first + i > 0) {
                        const ref_ordered_fault_t *pe = &ordered[base + first + i - 1];
                        dup = tbl[pe->ctx].va_space == va_space && pe->fault_address == e->fault_address;
                    }
                    if (dup && (entry_flags[base + first + i - 1] & REF_SF_INVALID_PREFETCH)) {
                        fl |= REF_SF_INVALID_PREFETCH;                   /* fault_entry_duplicate_flags :1257-1272 */
                        batch[b].num_invalid_prefetch_faults += e->num_instances;
                    }
                    if (dup)
                        fl |= REF_SF_DUPLICATE;
                    /* update_batch_and_notify_fault, called by service_fault_batch_fatal_notify (:1692, :1200-1203) */
                    batch[b].num_duplicate_faults += dup ? e->num_instances : (uint32_t)e->num_instances - 1;
                    if (acc == ACC_PREFETCH) {
                        fl |= REF_SF_INVALID_PREFETCH;
                        batch[b].num_invalid_prefetch_faults += e->num_instances;
                    }
                    else {
                        fl |= REF_SF_FATAL | (REASON_INVALID_ADDRESS << REF_SF_REASON_SHIFT) | (CANCEL_ALL << REF_SF_CANCEL_SHIFT);
                    }
                    entry_flags[base + first + i] = fl;
                }
                continue;
            }

            logical = st->logical_prot[slot];
            mask_ptr_load(&gpu_res, st->gpu_resident, slot);
            for (i = 0; i < 3; i++) mask_ptr_load(&pte[i], st->gpu_pte[i], slot);
            or_zero(&newres);

            /* ---- service_fault_batch_block_locked :1433-1580 ---- */
            for (i = 0; i < count; i++) {
                const ref_ordered_fault_t *e = &ordered[base + first + i];
                const unsigned acc = e->fault_access_type & 7;
                const unsigned page = (unsigned)((e->fault_address - grp->block_start) >> 12);
                unsigned service_acc = acc;
                uint16_t fl = 0;
                bool dup = false, entry_fatal = false, entry_invalid_prefetch = false;

                if (i > 0)
                    dup = ordered[base + first + i - 1].fault_address == e->fault_address;   /* same va_space inside a run */
                batch[b].num_duplicate_faults += dup ? e->num_instances : (uint32_t)e->num_instances - 1;  /* :1200-1203 */
                if (dup) {
                    fl |= REF_SF_DUPLICATE;
                    if (prev_invalid_prefetch) {                                             /* :1264-1267 */
                        entry_invalid_prefetch = true;
                        batch[b].num_invalid_prefetch_faults += e->num_instances;
                    }
                    if (!prev_fatal)
                        goto next;                                                           /* :1481-1483 */
                }

                /* check_fault_access_permissions :1293-1360 */
                if (!logical_ok(logical, acc)) {
                    if (acc == ACC_PREFETCH) {
                        entry_invalid_prefetch = true;
                        batch[b].num_invalid_prefetch_faults += e->num_instances;
                        goto next;
                    }
                    else {
                        unsigned cancel = CANCEL_ALL;
                        service_acc = ACC_COUNT;
                        if (acc > ACC_READ) {
                            cancel = CANCEL_WRITE_AND_ATOMIC;
                            if (e->access_type_mask & (1u << ACC_READ)) {
                                if (logical_ok(logical, ACC_READ))
                                    service_acc = ACC_READ;
                                else
                                    cancel = CANCEL_ALL;
                            }
                        }
                        entry_fatal = true;
                        fl |= REF_SF_FATAL | (REASON_INVALID_PERMISSIONS << REF_SF_REASON_SHIFT) | (cancel << REF_SF_CANCEL_SHIFT);
                        if (service_acc == ACC_COUNT)
                            goto next;
                    }
                }
                /* already authorised -> nothing to do (:1518-1522; uvm_va_block.c:2576-2615) */
                if (or_test(&pte[prot_of(service_acc) - 1], page)) {
                    fl |= REF_SF_AUTHORIZED;
                    goto next;
                }
                if (g_th) {                                                                  /* :1524-1545 */
                    thrash_query_t q;
                    thrash_hint_t h;
                    or_mask_t c0, c1;
                    mask_ptr_load(&c0, st->cpu_resident[0], slot);
                    mask_ptr_load(&c1, st->cpu_resident[1], slot);
                    memset(&q, 0, sizeof(q));
                    q.time_stamp = g_batch_time[b]; q.slot = slot; q.page = (uint16_t)page; q.requester = (uint8_t)lay->gpu_id;
                    q.resident_processors = (or_test(&gpu_res, page) ? 1u << lay->gpu_id : 0u) | ((or_test(&c0, page) || or_test(&c1, page)) ? 1u : 0u);
                    q.preferred_location = st->preferred_location[slot] == 257 ? 0xFFFF : st->preferred_location[slot];
                    oracle_thrashing_get_hints(1, &q, g_th_cfg, g_th, &h);
                    if (h.type == 2) {
                        fl |= REF_SF_THROTTLED;
                        goto next;
                    }
                    if (h.type == 1) {
                        if (h.pin_residency != lay->gpu_id) {
                            fl |= REF_SF_PIN_REMOTE;
                            goto next;
                        }
                        fl |= REF_SF_PINNED;
                    }
                }
                /* new residency = the faulting GPU (harness policy, or the pin); :1560-1579 */
                or_set(&newres, page);
                page_faults++;
                fl |= REF_SF_SERVICED | (service_acc << REF_SF_ACCESS_SHIFT);
                if (page < first_page) first_page = page;
                if (page > last_page) last_page = page;
            next:
                if (entry_invalid_prefetch) fl |= REF_SF_INVALID_PREFETCH;
                entry_flags[base + first + i] = fl;
                prev_fatal = entry_fatal;
                prev_invalid_prefetch = entry_invalid_prefetch;
            }

            rec->page_fault_count = (uint16_t)page_faults;
            if (!page_faults)
                continue;

            /* ---- uvm_va_block_service_locked, uvm_va_block.c:12307-12359 ---- */
            oracle_service_block_apply(lay, st, slot, grp->block_start, newres.w, first_page, last_page + 1, rec, runs, &nrun, run_cap);
        }
    }
    if (n_runs)
        *n_runs = nrun;
    return nsvc;
}
