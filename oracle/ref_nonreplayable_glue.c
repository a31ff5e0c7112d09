/*
 * ref_nonreplayable_glue.c -- TEST INFRASTRUCTURE ONLY (oracle/_ref build -> libref_nonreplayable.so).
 *
 * Compiles the reference's non-replayable block servicing VERBATIM -- the line-range extract
 * kernel-open/nvidia-uvm/uvm_gpu_non_replayable_faults.c:345-445 (service_managed_fault_in_block_locked: perf event,
 * logical permission check -> fatal, select_residency, new_residency mask and region of the one page,
 * uvm_va_block_service_locked) -- on top of the service leg's shims (ref_service_glue.c: va_block, service context,
 * policy, thrashing and residency stand-ins; uvm_va_block_service_locked is routed to oracle_service_block_apply,
 * whose two callees can in turn be the reference's own prefetch file and make_resident).  Packets are parsed by the
 * reference's parser (ref_faults_glue.c).  The loop around it -- channel lookup, va_block lookup, the fatal paths of
 * service_fault_once :606-712 and service_non_managed_fault :543-604 -- is restated in the harness function below.
 */
#include "ref_service_glue.c"

static uvm_non_replayable_fault_buffer_t g_non_replayable;
/* the extract takes &gpu->parent->fault_buffer.non_replayable: the stand-in parent struct has no such member */
#define non_replayable replayable.non_replayable_ptr[0]
#define uvm_assert_rwsem_locked(l) ((void)0)
static const uvm_va_policy_t *uvm_va_policy_get(uvm_va_block_t *b, NvU64 addr) { (void)addr; return &b->managed_range->policy; }
static bool uvm_range_group_address_migratable(uvm_va_space_t *s, NvU64 addr) { (void)s; (void)addr; return true; }
static void uvm_processor_mask_set(uvm_processor_mask_t *m, uvm_processor_id_t id) { __set_bit(id.val, m->bitmap); }

#include "x_nonreplayable.h"    /* uvm_gpu_non_replayable_faults.c:345-445, verbatim */
#undef non_replayable

uint32_t ref_service_non_replayable_faults(uint32_t n, const uint32_t *packets, const ref_fault_gpu_params_t *gp,
                                           const ref_instance_ptr_entry_t *tbl, size_t tbl_n, const ref_va_layout_t *lay,
                                           const ref_block_state_t *st, unsigned pf_enable, unsigned pf_threshold,
                                           unsigned pf_min_faults, ref_fault_entry_t *entries, uint16_t *entry_flags,
                                           ref_service_group_t *svc, ref_copy_run_t *runs, uint32_t run_cap, uint32_t *n_runs)
{
    extern int oracle_prefetch_init(unsigned, unsigned, unsigned);
    const uint32_t B = gp->max_batch_size;
    uvm_parent_gpu_t parent;
    uvm_gpu_t gpu;
    uvm_va_space_t spaces[64];
    uvm_fault_buffer_entry_t fe;
    uint32_t i, nrun = 0, examined = 0;

    oracle_prefetch_init(pf_enable, pf_threshold, pf_min_faults);
    ref_faults_parse(packets, n, gp, entries);
    memset(&parent, 0, sizeof(parent));
    memset(spaces, 0, sizeof(spaces));
    parent.fault_buffer.replayable.block_service_context_ptr = &g_block_service_context;
    parent.fault_buffer.replayable.non_replayable_ptr = &g_non_replayable;
    g_block_service_context.block_context = &g_va_block_context;
    gpu.id.val = lay->gpu_id; gpu.parent = &parent;
    g_state = st; g_lay = lay;
    g_runs = runs; g_nrun = &nrun; g_run_cap = run_cap;

    for (i = 0; i < n; i++) {
        ref_fault_entry_t *e = &entries[i];
        ref_service_group_t *rec = &svc[i];
        size_t k;
        int ctx = -1;
        uint64_t block_start, off;
        uint32_t slot = 0xFFFFFFFFu;
        uvm_va_block_t blk;
        NV_STATUS status;

        e->num_instances = 1;
        e->access_type_mask = (uint8_t)(1u << e->fault_access_type);
        entry_flags[i] = 0;
        memset(rec, 0, sizeof(*rec));
        for (k = 0; k < tbl_n; k++)
            if (tbl[k].instance_ptr == e->instance_ptr_address && tbl[k].aperture == e->instance_ptr_aperture &&
                (tbl[k].ve_id == 0xFF || tbl[k].ve_id == e->ve_id)) { ctx = (int)k; break; }
        if (ctx < 0)
            continue;
        if (e->flags & REF_FE_FATAL) {
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
        if (slot == 0xFFFFFFFFu) {
            entry_flags[i] = (uint16_t)(REF_SF_FATAL | ((unsigned)uvm_tools_status_to_fatal_fault_reason(NV_ERR_INVALID_ADDRESS) << REF_SF_REASON_SHIFT));
            continue;
        }
        /* the verbatim function on this fault */
        memset(&fe, 0, sizeof(fe));
        fe.fault_address = e->fault_address;
        fe.fault_access_type = (uvm_fault_access_type_t)e->fault_access_type;
        fe.access_type_mask = e->access_type_mask;
        fe.num_instances = 1;
        spaces[tbl[ctx].va_space].index = tbl[ctx].va_space;
        fe.va_space = &spaces[tbl[ctx].va_space];
        fe.gpu = &gpu;
        memset(&blk, 0, sizeof(blk));
        blk.start = block_start; blk.end = block_start + (2ull << 20) - 1; blk.slot = slot;
        blk.va_space = fe.va_space;
        blk.managed_range = (void *)&blk.range_storage;
        blk.range_storage.policy.preferred_location.val = st->preferred_location[slot];
        g_block_service_context.num_retries = 0;
        g_rec = rec;
        status = service_managed_fault_in_block_locked(&blk, NULL, &fe, &g_block_service_context, true);
        if (status != NV_OK)
            return 0xFFFFFFFFu;
        if (fe.is_fatal)
            entry_flags[i] = (uint16_t)(REF_SF_FATAL | ((unsigned)fe.fatal_reason << REF_SF_REASON_SHIFT));
        else {
            entry_flags[i] = (uint16_t)(REF_SF_SERVICED | ((unsigned)fe.fault_access_type << REF_SF_ACCESS_SHIFT));
            rec->page_fault_count = 1;
        }
    }
    if (n_runs)
        *n_runs = nrun;
    return examined;
}
