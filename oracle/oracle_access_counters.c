/*
 * oracle_access_counters.c -- TEST INFRASTRUCTURE ONLY: CPU restatement of the access-counter notification path
 * (UVM = kernel-open/nvidia-uvm):
 *   parse      uvm_hal_turing_access_counter_buffer_parse_entry   uvm_turing_access_counter_buffer.c:154-198
 *              (get_address :105-120, get_access_counter_inst_aperture :88-103; packet layout clc365.h:33-80)
 *   fetch      fetch_access_counter_buffer_entries                uvm_gpu_access_counters.c:862-940 (mode BATCH_READY)
 *   preprocess preprocess_notifications :974-992, translate_notifications_instance_ptrs :942-972,
 *              comparators :798-860
 *   service    service_notifications_batch :1513-1615 / _in_block :1356-1427, expand_notification_block :1287-1354,
 *              service_va_block_locked :1079-1249 -> the back half of oracle_service.c
 * Harness policy as oracle_service.c; in addition the size of the CPU chunks behind a va_block is a parameter
 * (uvm_va_block_get_physical_size, uvm_va_block.c:2353-2373).  The kernel sort() is not stable: ties keep
 * notification_cache order (the canonical choice of the fault path).  Untranslated notifications (va_space NULL)
 * sort first and are skipped.
 *
 * Pinning: both halves are compared with the reference's own functions compiled verbatim
 * (tests/test_oracle_access_counters.py -k reference): the servicing half (expand_notification_block,
 * service_va_block_locked) by oracle/_ref/libref_access_counters.so -- records, copy runs and state identical; the
 * front half (uvm_hal_turing_access_counter_buffer_parse_entry, fetch_access_counter_buffer_entries,
 * translate_notifications_instance_ptrs, preprocess_notifications) by oracle/_ref/libref_access_counters_front.so --
 * parsed entries, counters, cleared VALID bits, runs and the sorted order identical up to the order of
 * notifications that compare equal, which the reference leaves undefined (unstable sort) and nothing downstream
 * depends on.
 */
#include <stdlib.h>

#include "oracle_common.h"
#include "ref_api.h"

typedef struct {
    uint64_t address, instance_ptr;
    uint32_t instance_aperture, mmu_engine_id, ve_id, counter_value, sub_granularity, bank, tag, valid;
} ref_ac_entry_t;                   /* 48 bytes, = uvmtier_access_counter_entry_t */
typedef struct {
    uint64_t address;
    uint16_t index, ctx;
    uint32_t sub_granularity;
} ref_ordered_notification_t;       /* 16 bytes */

void oracle_service_block_apply(const ref_va_layout_t *lay, const ref_block_state_t *st, uint32_t slot, uint64_t block_start,
                                const uint64_t *new_residency, unsigned region_first, unsigned region_outer,
                                ref_service_group_t *rec, ref_copy_run_t *runs, uint32_t *nrun_io, uint32_t run_cap);
int oracle_prefetch_init(unsigned, unsigned, unsigned);

static void ac_parse(const uint32_t *w, unsigned va_bits, ref_ac_entry_t *e)
{
    uint64_t addr = (uint64_t)w[2] | ((uint64_t)w[3] << 32);                     /* ADDR_LO + (ADDR_HI << 32) :111 */
    const unsigned sh = 64 - va_bits;
    e->address = (uint64_t)((int64_t)(addr << sh) >> sh);                        /* canonical form, uvm_gpu.c:428-459 */
    e->instance_ptr = ((uint64_t)(w[0] >> 12) | ((uint64_t)w[1] << 20)) << 12;   /* INST_LO 31:12, INST_HI; 4K page number :175-178 */
    e->instance_aperture = ((w[0] >> 10) & 3) == 0 ? 10 : 9;                     /* VID_MEM -> UVM_APERTURE_VID, else SYS :88-103 */
    e->mmu_engine_id = (w[7] >> 20) & 0x1ff;
    e->ve_id = (e->mmu_engine_id - 384) & 0xff;                                  /* get_ve_id, graphics engine */
    e->counter_value = w[6] & 0xffff;
    e->sub_granularity = w[4];
    e->bank = (w[0] >> 2) & 0xf;
    e->tag = w[5] & 0xfffff;
    e->valid = w[7] >> 31;
}

static int cmp_u64pair(const void *a, const void *b)
{
    const uint64_t *x = (const uint64_t *)a, *y = (const uint64_t *)b;
    if (x[0] != y[0]) return x[0] < y[0] ? -1 : 1;
    return x[1] < y[1] ? -1 : x[1] > y[1];
}

/* All batches between GET and PUT; outputs sliced by batch like the product's.  Returns the number of batches. */
uint32_t oracle_access_counters_fetch_preprocess(uint32_t *ring, uint32_t ring_entries, uint32_t get, uint32_t put, uint32_t B,
                                                 unsigned va_bits, const ref_instance_ptr_entry_t *tbl, uint32_t n_rows,
                                                 int clear_valid, ref_ordered_notification_t *ordered, ref_fault_group_t *groups,
                                                 ref_fault_batch_t *batch, ref_ac_entry_t *entries)
{
    const uint32_t n_total = put >= get ? put - get : ring_entries - get + put, nb = (n_total + B - 1) / B;
    uint32_t b, i;
    uint64_t *key = (uint64_t *)malloc((size_t)B * 2 * sizeof(uint64_t));
    ref_ac_entry_t *cache = (ref_ac_entry_t *)malloc((size_t)B * sizeof(*cache));
    uint16_t *ctx = (uint16_t *)malloc((size_t)B * sizeof(uint16_t));
    for (b = 0; b < nb; b++) {
        const uint32_t base = b * B, in_batch = n_total - base < B ? n_total - base : B;
        uint32_t n = 0, ng = 0;
        bool single = true;
        memset(&batch[b], 0, sizeof(batch[b]));
        for (i = 0; i < in_batch && entries; i++) {                              /* parity view: every entry of the batch */
            ac_parse(ring + 8 * (size_t)((get + base + i) % ring_entries), va_bits, &entries[base + i]);
            entries[base + i].valid = 0;
        }
        for (i = 0; i < in_batch; i++) {                                         /* :888-920 */
            uint32_t *w = ring + 8 * (size_t)((get + base + i) % ring_entries);
            ref_ac_entry_t e;
            ac_parse(w, va_bits, &e);
            if (!e.valid)
                break;                                                           /* stop at the first entry that is not ready */
            if (clear_valid)
                w[7] &= 0x7fffffffu;
            if (entries)
                entries[base + i].valid = 1;
            cache[n++] = e;
        }
        for (i = 0; i < n; i++) {                                                /* translate :942-972, single :925-930 */
            uint32_t k, rank = 0;
            ctx[i] = 0xFFFF;
            for (k = 0; k < n_rows; k++)
                if (tbl[k].instance_ptr == cache[i].instance_ptr && tbl[k].aperture == cache[i].instance_aperture &&
                    (tbl[k].ve_id == REF_VE_ID_ANY || tbl[k].ve_id == cache[i].ve_id)) {
                    ctx[i] = (uint16_t)k;
                    rank = tbl[k].sort_rank + 1u;
                    break;
                }
            if (cache[i].instance_ptr != cache[0].instance_ptr || cache[i].ve_id != cache[0].ve_id ||
                cache[i].instance_aperture != cache[0].instance_aperture)
                single = false;
            /* (va_space, gpu) rank, then address (page number), then cache order :834-851 */
            key[2 * i] = ((uint64_t)rank << 45) | ((cache[i].address >> 12) & ((1ull << 45) - 1));
            key[2 * i + 1] = i;
        }
        qsort(key, n, 2 * sizeof(uint64_t), cmp_u64pair);
        for (i = 0; i < n; i++) {
            const uint32_t idx = (uint32_t)key[2 * i + 1];
            ref_ordered_notification_t *o = &ordered[base + i];
            o->address = cache[idx].address; o->index = (uint16_t)idx; o->ctx = ctx[idx]; o->sub_granularity = cache[idx].sub_granularity;
            if (o->ctx == 0xFFFF)
                continue;                                                        /* va_space NULL: ignored */
            /* a run: same va_space and gpu, address inside the same va_block (:1381-1384) */
            if (i == 0 || (key[2 * i] >> 9) != (key[2 * (i - 1)] >> 9)) {
                ref_fault_group_t *g = &groups[base + ng++];
                g->block_start = o->address & ~((2ull << 20) - 1);
                g->ctx = o->ctx; g->first_fault_index = (uint16_t)i; g->block_faults = 0; g->flags = 0;
            }
            groups[base + ng - 1].block_faults++;
        }
        batch[b].num_cached_faults = n;
        batch[b].num_coalesced_faults = n;
        batch[b].is_single_instance_ptr = single;
        batch[b].num_block_groups = ng;
    }
    free(key); free(cache); free(ctx);
    return nb;
}

/*
 * service_notifications_batch for every batch in order.  svc[]: one record per run whose va_block exists, in
 * (batch, run) order; returns their number.  granularity_2m / cpu_chunk_2m: see include/uvmtier.h section 1e.
 */
uint32_t oracle_access_counters_service(uint32_t n_batches, uint32_t B, const ref_ordered_notification_t *ordered,
                                        const ref_fault_group_t *groups, const ref_fault_batch_t *batch,
                                        const ref_instance_ptr_entry_t *tbl, const ref_va_layout_t *lay, const ref_block_state_t *st,
                                        unsigned pf_enable, unsigned pf_threshold, unsigned pf_min_faults, int granularity_2m,
                                        int cpu_chunk_2m, ref_service_group_t *svc, ref_copy_run_t *runs, uint32_t run_cap,
                                        uint32_t *n_runs)
{
    uint32_t b, g, i, nsvc = 0, nrun = 0;
    oracle_prefetch_init(pf_enable, pf_threshold, pf_min_faults);
    for (b = 0; b < n_batches; b++) {
        const uint32_t base = b * B;
        for (g = 0; g < batch[b].num_block_groups; g++) {
            const ref_fault_group_t *grp = &groups[base + g];
            const uint32_t va_space = tbl[grp->ctx].va_space;
            const uint64_t off = grp->block_start - lay->va_base;
            or_mask_t accessed, gpu, cpu0, cpu1, cpu, pte0, newres;
            ref_service_group_t *rec;
            uint32_t slot, w;
            unsigned first_page = OR_PAGES, last_page = 0, pages, p;
            if (!(grp->block_start >= lay->va_base && (off >> 21) < lay->blocks_per_space && va_space < lay->n_spaces))
                continue;                                          /* no va_block at this address */
            slot = va_space * lay->blocks_per_space + (uint32_t)(off >> 21);
            rec = &svc[nsvc++];
            memset(rec, 0, sizeof(*rec));
            rec->slot = slot; rec->batch = (uint16_t)b; rec->group = g; rec->hint_residency = 257; rec->run_offset = nrun;
            or_load(&gpu, st->gpu_resident + 8 * (size_t)slot);
            or_load(&cpu0, st->cpu_resident[0] + 8 * (size_t)slot);
            or_load(&cpu1, st->cpu_resident[1] + 8 * (size_t)slot);
            or_load(&pte0, st->gpu_pte[0] + 8 * (size_t)slot);
            or_or(&cpu, &cpu0, &cpu1);
            or_zero(&accessed);
            for (i = 0; granularity_2m && i < grp->block_faults; i++) {          /* expand_notification_block :1287-1354 */
                const ref_ordered_notification_t *o = &ordered[base + grp->first_fault_index + i];
                const unsigned page = (unsigned)((o->address - grp->block_start) >> 12) & 511;
                if (or_test(&gpu, page) || !or_test(&cpu, page))                 /* closest resident copy :1307-1315 */
                    continue;
                if (!cpu_chunk_2m) {
                    or_set(&accessed, page);                                     /* :1325-1326 */
                }
                else {
                    unsigned r;
                    for (r = 0; r < 32; r++)                                     /* regions of 16 pages :1335-1346 */
                        if ((o->sub_granularity >> r) & 1)
                            or_fill_range(&accessed, 16 * r, 16 * r + 16);
                    or_and(&accessed, &accessed, &cpu);                          /* :1351 */
                }
            }
            /* service_va_block_locked: the GPU must still map something in the block (:1103-1104); pages already on
             * the GPU are stale (:1122-1123); every remaining page goes to the GPU (harness policy) */
            if (or_empty(&pte0))
                or_zero(&accessed);
            or_andnot(&newres, &accessed, &gpu);
            pages = or_weight(&newres);
            rec->page_fault_count = (uint16_t)pages;
            if (!pages)
                continue;
            for (w = 0; w < OR_WORDS; w++) if (newres.w[w]) { first_page = 64 * w + (unsigned)__builtin_ctzll(newres.w[w]); break; }
            for (p = OR_WORDS; p-- > 0;) if (newres.w[p]) { last_page = 64 * p + 63 - (unsigned)__builtin_clzll(newres.w[p]); break; }
            oracle_service_block_apply(lay, st, slot, grp->block_start, newres.w, first_page, last_page + 1, rec, runs, &nrun, run_cap);
        }
    }
    if (n_runs)
        *n_runs = nrun;
    return nsvc;
}
