/*
 * oracle_nonreplayable.c -- TEST INFRASTRUCTURE ONLY: CPU restatement of non-replayable fault servicing
 * (UVM = kernel-open/nvidia-uvm, uvm_gpu_non_replayable_faults.c):
 *   fetch_non_replayable_fault_buffer_entries  :179-236  parse every packet of the shadow-buffer copy; one instance,
 *                                                        access_type_mask = the one bit, is_fatal from the fault type
 *   service_fault_once                         :606-712  channel lookup (dropped when it is gone, :623-632), va_block
 *                                                        lookup, managed or not
 *   service_managed_fault_in_block_locked      :345-445  logical permissions -> fatal; else new_residency = the page,
 *                                                        region = the page, uvm_va_block_service_locked
 *   service_non_managed_fault                  :543-604  no ATS in the harness: fatal, NV_ERR_INVALID_ADDRESS
 * Faults are serviced one after the other in buffer order.  Harness policy as in oracle_service.c (the destination is
 * the faulting GPU); the back half is oracle_service_block_apply (prefetch hint, make_resident, mapping).
 *
 * Pinning: service_managed_fault_in_block_locked is compiled verbatim in oracle/_ref/libref_nonreplayable.so
 * (ref_nonreplayable_glue.c) and compared with this file by tests/test_oracle_nonreplayable.py; the parser is the one
 * the replayable path pins (libref_faults.so).
 */
#include "oracle_common.h"
#include "ref_api.h"

void oracle_faults_parse(const uint32_t *ring, uint32_t n_entries, const ref_fault_gpu_params_t *gp, ref_fault_entry_t *entries);
void oracle_service_block_apply(const ref_va_layout_t *lay, const ref_block_state_t *st, uint32_t slot, uint64_t block_start,
                                const uint64_t *new_residency, unsigned region_first, unsigned region_outer,
                                ref_service_group_t *rec, ref_copy_run_t *runs, uint32_t *nrun_io, uint32_t run_cap);
int oracle_prefetch_init(unsigned, unsigned, unsigned);

enum { NR_ACC_PREFETCH = 0, NR_ACC_READ, NR_ACC_WRITE, NR_ACC_ATOMIC_WEAK, NR_ACC_ATOMIC_STRONG };
static unsigned nr_prot_of(unsigned acc)     /* uvm_fault_access_type_to_prot, uvm_hal_types.h:522-537 */
{
    if (acc == NR_ACC_ATOMIC_STRONG) return 3;
    if (acc == NR_ACC_ATOMIC_WEAK || acc == NR_ACC_WRITE) return 2;
    return 1;
}

/*
 * packets: n x 8 words (the shadow-buffer copy).  Outputs, one per fault i:
 *   entries[i]       parsed entry (num_instances 1, access_type_mask the one bit)
 *   entry_flags[i]   REF_SF_SERVICED | access << 5, or REF_SF_FATAL | reason << 8 (1 address, 2 permissions, 3 fault
 *                    type), or 0 (channel gone: dropped)
 *   svc[i]           record of fault i (all zero for dropped and fatal-type faults); batch = i / B, group = i % B with
 *                    B = gp->max_batch_size, the slicing of the product's arrays
 *   runs / *n_runs   copy runs in fault order
 * Returns the number of faults that reached a va_block or were found outside every va_range.
 */
uint32_t oracle_service_non_replayable_faults(uint32_t n, const uint32_t *packets, const ref_fault_gpu_params_t *gp,
                                              const ref_instance_ptr_entry_t *tbl, size_t tbl_n, const ref_va_layout_t *lay,
                                              const ref_block_state_t *st, unsigned pf_enable, unsigned pf_threshold,
                                              unsigned pf_min_faults, ref_fault_entry_t *entries, uint16_t *entry_flags,
                                              ref_service_group_t *svc, ref_copy_run_t *runs, uint32_t run_cap, uint32_t *n_runs)
{
    const uint32_t B = gp->max_batch_size;
    uint32_t i, nrun = 0, examined = 0;
    oracle_prefetch_init(pf_enable, pf_threshold, pf_min_faults);
    oracle_faults_parse(packets, n, gp, entries);                                   /* :206-234 */
    for (i = 0; i < n; i++) {
        ref_fault_entry_t *e = &entries[i];
        ref_service_group_t *rec = &svc[i];
        size_t k;
        int ctx = -1;
        uint64_t block_start, off;
        uint32_t slot = 0xFFFFFFFFu;
        unsigned page;
        or_mask_t newres;

        e->num_instances = 1;                                                       /* :215 */
        e->access_type_mask = (uint8_t)(1u << e->fault_access_type);                /* :216 */
        entry_flags[i] = 0;
        memset(rec, 0, sizeof(*rec));
        for (k = 0; k < tbl_n; k++)                                                 /* uvm_parent_gpu_fault_entry_to_va_space, :623-632 */
            if (tbl[k].instance_ptr == e->instance_ptr_address && tbl[k].aperture == e->instance_ptr_aperture &&
                (tbl[k].ve_id == 0xFF || tbl[k].ve_id == e->ve_id)) { ctx = (int)k; break; }
        if (ctx < 0)
            continue;                                                               /* the channel is gone: dropped */
        if (e->flags & REF_FE_FATAL) {                                              /* :221-224, reported at :693-694 */
            entry_flags[i] = (uint16_t)(REF_SF_FATAL | (3u << REF_SF_REASON_SHIFT));
            continue;
        }
        examined++;
        rec->batch = (uint16_t)(i / B); rec->group = i % B; rec->hint_residency = 257; rec->run_offset = nrun;
        block_start = e->fault_address & ~((2ull << 20) - 1);
        off = block_start - lay->va_base;
        if (block_start >= lay->va_base && (off >> 21) < lay->blocks_per_space && tbl[ctx].va_space < lay->n_spaces)
            slot = tbl[ctx].va_space * lay->blocks_per_space + (uint32_t)(off >> 21);
        rec->slot = slot;
        if (slot == 0xFFFFFFFFu) {                                                  /* service_non_managed_fault :543-604 */
            entry_flags[i] = (uint16_t)(REF_SF_FATAL | (1u << REF_SF_REASON_SHIFT));
            rec->run_offset = 0;
            continue;
        }
        /* service_managed_fault_in_block_locked :345-445 */
        if (nr_prot_of(e->fault_access_type) > st->logical_prot[slot]) {           /* :372-387 */
            entry_flags[i] = (uint16_t)(REF_SF_FATAL | (2u << REF_SF_REASON_SHIFT));
            rec->run_offset = 0;
            continue;
        }
        page = (unsigned)((e->fault_address - block_start) >> 12);
        or_zero(&newres);
        or_set(&newres, page);                                                      /* :418-419 */
        rec->page_fault_count = 1;
        entry_flags[i] = (uint16_t)(REF_SF_SERVICED | ((unsigned)e->fault_access_type << REF_SF_ACCESS_SHIFT));
        oracle_service_block_apply(lay, st, slot, block_start, newres.w, page, page + 1, rec, runs, &nrun, run_cap);   /* :431 */
        if (!rec->run_count)
            rec->run_offset = 0;
    }
    if (n_runs)
        *n_runs = nrun;
    return examined;
}
