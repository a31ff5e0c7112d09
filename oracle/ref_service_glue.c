/*
 * ref_service_glue.c -- TEST INFRASTRUCTURE ONLY (oracle/_ref build -> libref_service.so).
 *
 * Compiles the reference's per-va_block fault servicing loop VERBATIM -- line-range extracts of
 * kernel-open/nvidia-uvm/uvm_gpu_replayable_faults.c:
 *   :1214-1272  mark_fault_invalid_prefetch / _throttled / _fatal, fault_entry_duplicate_flags
 *   :1293-1360  check_fault_access_permissions
 *   :1375-1597  service_fault_batch_block_locked
 *   :1656-1693  service_fault_batch_fatal, service_fault_batch_fatal_notify
 * (plus, through ref_faults_glue.c, :1180-1212 check_fault_entry_duplicate / update_batch_and_notify_fault) --
 * against stand-ins for what the loop calls outside that file:
 *   uvm_va_block_check_logical_permissions   -> prot_of(access) <= logical_prot[slot]  (uvm_va_block.c:12361-12416
 *                                               with uvm_enable_builtin_tests, as in the reference's own test runs)
 *   uvm_va_block_page_is_gpu_authorized      -> the slot's gpu_pte bits                 (uvm_va_block.c:2576-2615)
 *   uvm_perf_thrashing_get_hint              -> NONE (harness policy), or -- ref_service_set_thrashing() -- the thrashing
 *                                               restatement's hint for (va_block, page, this GPU, the batch's time stamp)
 *   uvm_va_block_select_residency            -> the pin's residency for a PIN hint (uvm_va_block.c:11624-11628), else the
 *                                               faulting GPU                          (harness policy)
 *   uvm_va_block_service_locked              -> oracle_service_block_apply(): the CPU restatement of the prefetch
 *                                               hint + make_resident + mapping back half (oracle_service.c)
 * The batch-level driver below restates service_fault_batch (:2232-2377) / _dispatch (:1946-2035) on the flat
 * layout: fatal entries are skipped, an address in no va_range goes through service_fault_batch_fatal_notify
 * (verbatim), everything else through service_fault_batch_block_locked (verbatim).
 *
 * What this pins: the entry loop of F4 (duplicates, permissions, authorised-already, new_residency / region
 * construction, batch counters, per-entry fatal / invalid-prefetch marks).  The back half stays a restatement.
 */
#include "ref_faults_glue.c"    /* shim types + extracts :695-1212 */

#include "x_va_block_types.h"   /* uvm_va_block_types.h:36-80: uvm_page_mask_t, uvm_va_block_region_t, ... */

#define UVM_ID_MAX_PROCESSORS_S 258
typedef struct { DECLARE_BITMAP(bitmap, UVM_ID_MAX_PROCESSORS_S); } uvm_processor_mask_t;
static void uvm_processor_mask_zero(uvm_processor_mask_t *m) { bitmap_zero(m->bitmap, UVM_ID_MAX_PROCESSORS_S); }
static bool uvm_processor_mask_test_and_set(uvm_processor_mask_t *m, uvm_processor_id_t id) { return __test_and_set_bit(id.val, m->bitmap); }
#define UVM_ID_INVALID ((uvm_processor_id_t){ .val = 257 })

typedef struct { uvm_processor_id_t preferred_location; } uvm_va_policy_t;
struct uvm_va_block_struct {
    NvU64 start, end;
    int lock;
    uvm_va_space_t *va_space;
    struct { uvm_va_policy_t policy; } *managed_range, range_storage;
    uvm_processor_mask_t resident, mapped;           /* used by the access-counter leg (ref_access_counters_glue.c) */
    uint32_t slot;                                   /* harness: row of the flat state table */
};
typedef struct { struct { void *vma; } hmm; void *mm; } uvm_va_block_context_t;
typedef struct { uvm_va_block_t *va_block; uvm_va_block_region_t region; } uvm_reverse_map_t;
typedef struct { struct { NvU64 start, end; } node; } uvm_range_group_range_t;
static inline NvU64 uvm_va_block_size(uvm_va_block_t *b) { return b->end - b->start + 1; }
static inline size_t uvm_va_block_num_cpu_pages(uvm_va_block_t *b) { return uvm_va_block_size(b) / PAGE_SIZE; }
static uvm_page_index_t uvm_va_block_cpu_page_index(uvm_va_block_t *b, NvU64 addr) { return (addr - b->start) / PAGE_SIZE; }
static inline bool uvm_va_block_is_hmm(uvm_va_block_t *b) { (void)b; return false; }
static inline uvm_va_space_t *uvm_va_block_get_va_space(uvm_va_block_t *b) { return b->va_space; }
static inline NvU64 uvm_va_block_cpu_page_address(uvm_va_block_t *b, uvm_page_index_t i) { return b->start + PAGE_SIZE * (NvU64)i; }

#include "x_va_block_helpers.h" /* uvm_va_block.h: region + page-mask helpers, verbatim */
#include "x_prot.h"             /* uvm_hal_types.h:181-188 uvm_prot_t, :522-537 uvm_fault_access_type_to_prot */
#include "x_tools_reason.h"     /* uvm_tools.h:46-62 uvm_tools_status_to_fatal_fault_reason */

typedef enum { UVM_PERF_THRASHING_HINT_TYPE_NONE = 0, UVM_PERF_THRASHING_HINT_TYPE_PIN = 1,
               UVM_PERF_THRASHING_HINT_TYPE_THROTTLE = 2 } uvm_perf_thrashing_hint_type_t;     /* uvm_perf_thrashing.h:33-50 */
typedef struct { uvm_perf_thrashing_hint_type_t type; struct { uvm_processor_id_t residency; } pin; } uvm_perf_thrashing_hint_t;
typedef enum { UVM_SERVICE_OPERATION_REPLAYABLE_FAULTS, UVM_SERVICE_OPERATION_NON_REPLAYABLE_FAULTS,
               UVM_SERVICE_OPERATION_ACCESS_COUNTERS } uvm_service_operation_t;                /* uvm_va_block_types.h:343-348 */
typedef struct { NvU64 start, end; bool migratable; } uvm_range_group_range_iter_t;
typedef struct uvm_va_block_retry_struct uvm_va_block_retry_t;

/* uvm_service_block_context_t: the fields the loop touches (uvm_va_block_types.h:262-340) */
typedef struct uvm_service_block_context_struct {
    uvm_processor_mask_t resident_processors, update_processors;
    NvU32 thrashing_pin_count, read_duplicate_count;
    uvm_page_mask_t thrashing_pin_mask, read_duplicate_mask;
    struct { uvm_page_mask_t new_residency; } per_processor_masks[UVM_ID_MAX_PROCESSORS_S];
    uvm_va_block_context_t *block_context;
    uvm_va_block_region_t region;
    NvU32 num_retries;
    NvU8 access_type[PAGES_PER_UVM_VA_BLOCK];
} uvm_service_block_context_t;

/* the loop takes &gpu->parent->fault_buffer.replayable.block_service_context; the stand-in parent struct holds a
 * pointer to the context (its type is not known where that struct is declared), so the member name is mapped to
 * the pointee */
static uvm_service_block_context_t g_block_service_context;
static uvm_va_block_context_t g_va_block_context;
#define block_service_context block_service_context_ptr[0]

/* ---- harness state the stand-ins read ---- */
static const ref_block_state_t *g_state;
static const ref_va_layout_t *g_lay;

#define uvm_assert_mutex_locked(l) ((void)0)
#undef BUILD_BUG_ON
#define BUILD_BUG_ON(e) ((void)0)
static void uvm_range_group_range_migratability_iter_first(uvm_va_space_t *s, NvU64 start, NvU64 end, uvm_range_group_range_iter_t *it)
{ (void)s; it->start = start; it->end = end; it->migratable = true; }
static void uvm_range_group_range_migratability_iter_next(uvm_va_space_t *s, uvm_range_group_range_iter_t *it, NvU64 end)
{ (void)s; it->end = end; }
static const uvm_va_policy_t *uvm_hmm_find_policy_end(uvm_va_block_t *b, void *vma, NvU64 addr, NvU64 *end)
{ (void)vma; (void)addr; *end = b->end; return &b->managed_range->policy; }

static NV_STATUS uvm_va_block_check_logical_permissions(uvm_va_block_t *b, uvm_va_block_context_t *c, uvm_processor_id_t id,
                                                        uvm_page_index_t page, uvm_fault_access_type_t access, bool allow_migration)
{
    (void)c; (void)id; (void)page; (void)allow_migration;
    return (unsigned)uvm_fault_access_type_to_prot(access) <= g_state->logical_prot[b->slot] ? NV_OK : NV_ERR_INVALID_ACCESS_TYPE;
}
static bool uvm_va_block_page_is_gpu_authorized(uvm_va_block_t *b, uvm_page_index_t page, uvm_gpu_id_t id, uvm_prot_t prot)
{
    (void)id;
    return (g_state->gpu_pte[prot - UVM_PROT_READ_ONLY][8 * (size_t)b->slot + page / 64] >> (page % 64)) & 1;
}
/* thrashing hints: off unless ref_service_set_thrashing() attached the restatement's tables (oracle_thrashing.c, itself held
 * against the reference's uvm_perf_thrashing_get_hint by libref_thrashing_hint.so) */
#include "oracle_thrashing.h"
extern void oracle_service_set_thrashing(const thrash_cfg_t *cfg, const thrash_tables_t *th, const uint64_t *batch_time_ns);
static const thrash_cfg_t *g_thr_cfg;
static const thrash_tables_t *g_thr;
static const uint64_t *g_thr_time;
static uint32_t g_cur_batch;
void ref_service_set_thrashing(const thrash_cfg_t *cfg, const thrash_tables_t *th, const uint64_t *batch_time_ns)
{
    g_thr_cfg = cfg; g_thr = cfg ? th : NULL; g_thr_time = batch_time_ns;
    oracle_service_set_thrashing(cfg, th, batch_time_ns);     /* the back half's prefetch call reads the thrashing pages */
}
static uvm_perf_thrashing_hint_t uvm_perf_thrashing_get_hint(uvm_va_block_t *b, uvm_va_block_context_t *c, NvU64 addr, uvm_processor_id_t id)
{
    uvm_perf_thrashing_hint_t h = { UVM_PERF_THRASHING_HINT_TYPE_NONE, { { 257 } } };
    (void)c;
    if (g_thr) {
        const unsigned page = (unsigned)((addr - b->start) / PAGE_SIZE);
        const size_t w = 8 * (size_t)b->slot + page / 64;
        thrash_query_t q;
        thrash_hint_t out;
        memset(&q, 0, sizeof(q));
        q.time_stamp = g_thr_time[g_cur_batch]; q.slot = b->slot; q.page = (uint16_t)page; q.requester = (uint8_t)id.val;
        q.resident_processors = (((g_state->gpu_resident[w] >> (page % 64)) & 1) ? 1u << id.val : 0u) |
                                ((((g_state->cpu_resident[0][w] | g_state->cpu_resident[1][w]) >> (page % 64)) & 1) ? 1u : 0u);
        q.preferred_location = g_state->preferred_location[b->slot] == 257 ? 0xFFFF : g_state->preferred_location[b->slot];
        oracle_thrashing_get_hints(1, &q, g_thr_cfg, g_thr, &out);
        h.type = (uvm_perf_thrashing_hint_type_t)out.type;
        h.pin.residency.val = out.pin_residency;
    }
    return h;
}
static uvm_processor_id_t uvm_va_block_select_residency(uvm_va_block_t *b, uvm_va_block_context_t *c, uvm_page_index_t page,
                                                        uvm_processor_id_t id, NvU32 access_type_mask, const uvm_va_policy_t *policy,
                                                        const uvm_perf_thrashing_hint_t *hint, uvm_service_operation_t op,
                                                        const bool hmm_migratable, bool *read_duplicate)
{
    (void)b; (void)c; (void)page; (void)access_type_mask; (void)policy; (void)op; (void)hmm_migratable;
    *read_duplicate = false;
    return hint->type == UVM_PERF_THRASHING_HINT_TYPE_PIN ? hint->pin.residency : id;
}
static NV_STATUS uvm_va_block_set_cancel(uvm_va_block_t *b, uvm_va_block_context_t *c, uvm_gpu_t *g) { (void)b; (void)c; (void)g; return NV_OK; }

/* back half: CPU restatement (oracle_service.c), called with what the verbatim loop built */
extern void oracle_service_block_apply(const ref_va_layout_t *lay, const ref_block_state_t *st, uint32_t slot, uint64_t block_start,
                                       const uint64_t *new_residency, unsigned region_first, unsigned region_outer,
                                       ref_service_group_t *rec, ref_copy_run_t *runs, uint32_t *nrun, uint32_t run_cap);
static ref_service_group_t *g_rec;
static uint16_t *g_pin_flags;
static ref_copy_run_t *g_runs;
static uint32_t *g_nrun, g_run_cap;
static NV_STATUS uvm_va_block_service_locked(uvm_processor_id_t id, uvm_va_block_t *b, uvm_va_block_retry_t *r,
                                             uvm_service_block_context_t *ctx)
{
    (void)r;
    oracle_service_block_apply(g_lay, g_state, b->slot, b->start, (const uint64_t *)ctx->per_processor_masks[id.val].new_residency.bitmap,
                               ctx->region.first, ctx->region.outer, g_rec, g_runs, g_nrun, g_run_cap);
    return NV_OK;
}

#include "x_service.h"          /* the four extracts listed at the top, verbatim */

/*
 * Same contract as oracle_service_fault_batches (oracle_service.c): ordered / groups / batch are the
 * product-layout outputs of the preprocess stage.  entry_flags carries only what the reference keeps per entry:
 * REF_SF_FATAL | reason | cancel mode, REF_SF_INVALID_PREFETCH.
 */
uint32_t ref_service_fault_batches(uint32_t n_batches, uint32_t max_batch_size, const ref_ordered_fault_t *ordered,
                                   const ref_fault_group_t *groups, ref_fault_batch_t *batch,
                                   const ref_instance_ptr_entry_t *tbl, const ref_va_layout_t *lay,
                                   const ref_block_state_t *st, unsigned pf_enable, unsigned pf_threshold,
                                   unsigned pf_min_faults, uint16_t *entry_flags, ref_service_group_t *svc,
                                   ref_copy_run_t *runs, uint32_t run_cap, uint32_t *n_runs)
{
    extern int oracle_prefetch_init(unsigned, unsigned, unsigned);
    uvm_parent_gpu_t parent;
    uvm_gpu_t gpu;
    uvm_fault_service_batch_context_t bc;
    uvm_fault_utlb_info_t utlb;
    uint32_t b, nsvc = 0, nrun = 0;
    size_t i;
    (void)groups;

    oracle_prefetch_init(pf_enable, pf_threshold, pf_min_faults);
    g_pin_flags = calloc((size_t)n_batches * max_batch_size + 1, sizeof(uint16_t));
    memset(&parent, 0, sizeof(parent));
    gpu.id.val = lay->gpu_id;
    gpu.parent = &parent;
    for (i = 0; i < 65536; i++) g_spaces[i].index = (int)i;
    g_state = st; g_lay = lay; g_runs = runs; g_nrun = &nrun; g_run_cap = run_cap;
    g_block_service_context.block_context = &g_va_block_context;
    parent.fault_buffer.replayable.block_service_context_ptr = &g_block_service_context;

    memset(&bc, 0, sizeof(bc));
    bc.fault_cache = calloc(max_batch_size, sizeof(*bc.fault_cache));
    bc.ordered_fault_cache = calloc(max_batch_size, sizeof(*bc.ordered_fault_cache));
    bc.utlbs = &utlb;

    for (b = 0; b < n_batches; b++) {
        const size_t base = (size_t)b * max_batch_size;
        const uint32_t n = batch[b].num_coalesced_faults;
        uint32_t k, g = 0;
        if (batch[b].status != 0)
            continue;                                   /* batch is refetched, nothing serviced (:2946) */
        memset(&utlb, 0, sizeof(utlb));
        bc.num_coalesced_faults = n;
        bc.num_duplicate_faults = 0; bc.num_invalid_prefetch_faults = 0; bc.fatal_va_space = NULL; bc.fatal_gpu = NULL;
        for (k = 0; k < n; k++) {                       /* the sorted view as fault entries */
            const ref_ordered_fault_t *o = &ordered[base + k];
            uvm_fault_buffer_entry_t *e = &bc.fault_cache[k];
            memset(e, 0, sizeof(*e));
            e->fault_address = o->fault_address;
            e->fault_access_type = (uvm_fault_access_type_t)(o->fault_access_type & 7);
            e->access_type_mask = o->access_type_mask;
            e->num_instances = o->num_instances;
            e->is_fatal = (o->fault_access_type & 0x80) != 0;
            e->va_space = &g_spaces[tbl[o->ctx].va_space];
            e->gpu = &gpu;
            bc.ordered_fault_cache[k] = e;
            entry_flags[base + k] = 0;
        }
        for (k = 0; k < n; g++) {                       /* service_fault_batch :2258-2350 */
            uvm_fault_buffer_entry_t *e = bc.ordered_fault_cache[k];
            const uint32_t va_space = (uint32_t)e->va_space->index;
            const uint64_t block_start = e->fault_address & ~((1ull << 21) - 1), off = block_start - lay->va_base;
            NvU32 block_faults = 0;
            ref_service_group_t *rec;
            if (e->is_fatal) {                          /* :2294-2303 */
                k++;
                continue;
            }
            rec = &svc[nsvc++];
            memset(rec, 0, sizeof(*rec));
            rec->batch = (uint16_t)b; rec->group = g; rec->hint_residency = 257; rec->run_offset = nrun;
            rec->slot = 0xFFFFFFFFu;
            if (block_start >= lay->va_base && (off >> 21) < lay->blocks_per_space && va_space < lay->n_spaces) {
                uvm_va_block_t blk;
                memset(&blk, 0, sizeof(blk));
                blk.slot = va_space * lay->blocks_per_space + (uint32_t)(off >> 21);
                blk.start = block_start; blk.end = block_start + (1ull << 21) - 1;
                blk.va_space = e->va_space;
                blk.managed_range = (void *)&blk.range_storage;
                blk.range_storage.policy.preferred_location.val = st->preferred_location[blk.slot];
                rec->slot = blk.slot;
                g_rec = rec;
                g_block_service_context.num_retries = 0;
                g_block_service_context.thrashing_pin_count = 0;
                g_cur_batch = b;
                service_fault_batch_block_locked(&gpu, &blk, NULL, &bc, k, true, &block_faults);
                if (g_thr && g_block_service_context.thrashing_pin_count) {
                    /* thrashing_pin_mask (:1541-1545), per entry for the harness: the first entry of each pinned page */
                    uvm_page_mask_t seen;
                    NvU32 j;
                    uvm_page_mask_zero(&seen);
                    for (j = 0; j < block_faults; j++) {
                        const uvm_fault_buffer_entry_t *pe = bc.ordered_fault_cache[k + j];
                        const uvm_page_index_t page = (uvm_page_index_t)((pe->fault_address - block_start) / PAGE_SIZE);
                        if (pe->is_throttled || !uvm_page_mask_test(&g_block_service_context.thrashing_pin_mask, page) ||
                            uvm_page_mask_test(&seen, page))
                            continue;
                        uvm_page_mask_set(&seen, page);
                        g_pin_flags[base + (size_t)(pe - bc.fault_cache)] =
                            uvm_page_mask_test(&g_block_service_context.per_processor_masks[gpu.id.val].new_residency, page) &&
                            test_bit(gpu.id.val, g_block_service_context.resident_processors.bitmap)
                                ? REF_SF_PINNED : REF_SF_PIN_REMOTE;
                    }
                }
            }
            else {
                /* no va_range covers the address: one entry at a time (:2022-2031), the run = the group K1 formed */
                do {
                    NvU32 one = 0;
                    service_fault_batch_fatal_notify(&gpu, &bc, k + block_faults, NV_ERR_INVALID_ADDRESS, UVM_FAULT_CANCEL_VA_MODE_ALL, &one);
                    block_faults += one;
                } while (k + block_faults < n && !bc.ordered_fault_cache[k + block_faults]->is_fatal &&
                         bc.ordered_fault_cache[k + block_faults]->va_space == e->va_space &&
                         (bc.ordered_fault_cache[k + block_faults]->fault_address & ~((1ull << 21) - 1)) == block_start);
            }
            k += block_faults;
        }
        for (k = 0; k < n; k++) {
            const uvm_fault_buffer_entry_t *e = &bc.fault_cache[k];
            uint16_t fl = 0;
            if (e->is_fatal && !(ordered[base + k].fault_access_type & 0x80))
                fl |= REF_SF_FATAL | (uint16_t)(e->fatal_reason << 8) | (uint16_t)(e->replayable.cancel_va_mode << 12);
            if (e->is_invalid_prefetch)
                fl |= REF_SF_INVALID_PREFETCH;
            if (e->is_throttled)
                fl |= REF_SF_THROTTLED;
            entry_flags[base + k] = fl | g_pin_flags[base + k];
        }
        batch[b].num_duplicate_faults = bc.num_duplicate_faults;
        batch[b].num_invalid_prefetch_faults = bc.num_invalid_prefetch_faults;
    }
    free(bc.fault_cache);
    free(bc.ordered_fault_cache);
    free(g_pin_flags);
    g_pin_flags = NULL;
    if (n_runs)
        *n_runs = nrun;
    return nsvc;
}
