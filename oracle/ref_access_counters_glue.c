/*
 * ref_access_counters_glue.c -- TEST INFRASTRUCTURE ONLY (oracle/_ref build -> libref_access_counters.so).
 *
 * Compiles the reference's access-counter servicing VERBATIM -- line-range extracts of
 * kernel-open/nvidia-uvm/uvm_gpu_access_counters.c:
 *   :198-218    config_granularity_to_bytes
 *   :1079-1251  service_va_block_locked
 *   :1253-1260  for_each_sub_granularity_region
 *   :1287-1354  expand_notification_block
 * on top of the service leg's shims (ref_service_glue.c: service context, va_block, policy and thrashing stand-ins;
 * uvm_va_block_service_locked is routed to the CPU restatement's back half, itself pinned by libref_prefetch /
 * libref_resident).  Stand-ins added here: closest resident processor and resident masks from the flat table, the
 * CPU chunk size of the harness, processor-mask helpers.  The loop over the notifications of a va_block
 * (service_notifications_in_block :1381-1400) is restated in the harness function below.
 *
 * What this pins: N3's expansion (which pages a notification names) and its filters (GPU not mapped, pages already
 * on the GPU, region passed to the back half).
 */
#include "ref_service_glue.c"

static bool uvm_processor_mask_test(const uvm_processor_mask_t *m, uvm_processor_id_t id) { return test_bit(id.val, m->bitmap); }
static void uvm_processor_mask_set(uvm_processor_mask_t *m, uvm_processor_id_t id) { __set_bit(id.val, m->bitmap); }
static void uvm_processor_mask_clear(uvm_processor_mask_t *m, uvm_processor_id_t id) { __clear_bit(id.val, m->bitmap); }
static bool uvm_processor_mask_and(uvm_processor_mask_t *d, const uvm_processor_mask_t *a, const uvm_processor_mask_t *b)
{ return bitmap_and(d->bitmap, a->bitmap, b->bitmap, UVM_ID_MAX_PROCESSORS_S); }
static bool uvm_processor_mask_empty(const uvm_processor_mask_t *m) { return bitmap_empty(m->bitmap, UVM_ID_MAX_PROCESSORS_S); }
static uvm_processor_id_t ac_pm_next(const uvm_processor_mask_t *m, unsigned from)
{ uvm_processor_id_t r; r.val = (NvU32)find_next_bit(m->bitmap, 257, from); return r; }
#define for_each_id_in_mask(id, mask) for ((id) = ac_pm_next(mask, 0); (id).val < 257; (id) = ac_pm_next(mask, (id).val + 1))
#ifndef UVM_ID_IS_VALID
#define UVM_ID_IS_VALID(id) ((id).val < 257)
#endif
#ifndef UVM_ID_IS_GPU
#define UVM_ID_IS_GPU(id) ((id).val >= 1 && (id).val < 257)
#endif
#ifndef UVM_PAGE_SIZE_2M
#define UVM_PAGE_SIZE_2M (2ull << 20)
#endif
#ifndef NUMA_NO_NODE
#define NUMA_NO_NODE (-1)
#endif

struct mm_struct;
static bool uvm_id_equal(uvm_processor_id_t a, uvm_processor_id_t b) { return a.val == b.val; }
static const uvm_va_policy_t *uvm_va_policy_get(uvm_va_block_t *b, NvU64 addr) { (void)addr; return &b->managed_range->policy; }
static NV_STATUS uvm_hmm_find_vma(void *mm, void **vma, NvU64 addr) { (void)mm; (void)vma; (void)addr; return NV_OK; }
static NV_STATUS uvm_hmm_find_policy_vma_and_outer(uvm_va_block_t *b, void **vma, uvm_page_index_t first, const uvm_va_policy_t **policy,
                                                   uvm_page_index_t *outer)
{ (void)b; (void)vma; (void)first; (void)policy; (void)outer; return NV_OK; }

/* resident masks and closest resident processor from the flat table (one GPU, the CPU with two nodes) */
static uvm_page_mask_t g_ac_gpu_mask, g_ac_cpu_mask;
static NvU32 g_ac_gpu_id;
static NvU64 g_ac_cpu_chunk;
static uvm_gpu_t g_ac_gpu;
static void ac_load_masks(uvm_va_block_t *b)
{
    int w;
    for (w = 0; w < 8; w++) {
        ((uint64_t *)g_ac_gpu_mask.bitmap)[w] = g_state->gpu_resident[8 * (size_t)b->slot + w];
        ((uint64_t *)g_ac_cpu_mask.bitmap)[w] = g_state->cpu_resident[0][8 * (size_t)b->slot + w] | g_state->cpu_resident[1][8 * (size_t)b->slot + w];
    }
}
static uvm_page_mask_t *uvm_va_block_resident_mask_get(uvm_va_block_t *b, uvm_processor_id_t id, int nid)
{ (void)nid; ac_load_masks(b); return id.val == g_ac_gpu_id ? &g_ac_gpu_mask : &g_ac_cpu_mask; }
static uvm_processor_id_t uvm_va_block_page_get_closest_resident(uvm_va_block_t *b, uvm_va_block_context_t *c, uvm_page_index_t page,
                                                                 uvm_processor_id_t id)
{
    uvm_processor_id_t r;
    (void)c; (void)id;
    ac_load_masks(b);
    r.val = uvm_page_mask_test(&g_ac_gpu_mask, page) ? g_ac_gpu_id : (uvm_page_mask_test(&g_ac_cpu_mask, page) ? 0u : 257u);
    return r;
}
static uvm_gpu_t *uvm_gpu_get(uvm_processor_id_t id) { (void)id; return &g_ac_gpu; }
static NvU64 uvm_va_block_get_physical_size(uvm_va_block_t *b, uvm_processor_id_t id, uvm_page_index_t page)
{ (void)b; (void)id; (void)page; return g_ac_cpu_chunk; }

typedef enum {                                          /* nv_uvm_types.h:1103-1109 */
    UVM_ACCESS_COUNTER_GRANULARITY_64K = 1, UVM_ACCESS_COUNTER_GRANULARITY_2M = 2,
    UVM_ACCESS_COUNTER_GRANULARITY_16M = 3, UVM_ACCESS_COUNTER_GRANULARITY_16G = 4,
} UVM_ACCESS_COUNTER_GRANULARITY;
#define UVM_SIZE_1MB (1024ull * 1024)
#define UVM_SIZE_1GB (1024ull * 1024 * 1024)
typedef struct { uvm_gpu_t *gpu; uvm_va_space_t *va_space; } uvm_gpu_va_space_t;
typedef struct {
    struct { struct { UVM_ACCESS_COUNTER_GRANULARITY granularity; } rm; NvU64 sub_granularity_region_size;
             NvU32 sub_granularity_regions_per_translation; } current_config;
} uvm_access_counter_buffer_t;
typedef struct { NvU64 address; NvU32 sub_granularity; } uvm_access_counter_buffer_entry_t;

#include "x_access_counters.h"  /* the uvm_gpu_access_counters.c extracts listed at the top, verbatim */

/* ======================= flat harness API ======================= */
typedef struct { uint64_t address; uint16_t index, ctx; uint32_t sub_granularity; } ref_ordered_notification_t;

/* Same contract as oracle_access_counters_service (oracle_access_counters.c). */
uint32_t ref_access_counters_service(uint32_t n_batches, uint32_t B, const ref_ordered_notification_t *ordered,
                                     const ref_fault_group_t *groups, const ref_fault_batch_t *batch,
                                     const ref_instance_ptr_entry_t *tbl, const ref_va_layout_t *lay, const ref_block_state_t *st,
                                     unsigned pf_enable, unsigned pf_threshold, unsigned pf_min_faults, int granularity_2m,
                                     int cpu_chunk_2m, ref_service_group_t *svc, ref_copy_run_t *runs, uint32_t run_cap,
                                     uint32_t *n_runs)
{
    extern int oracle_prefetch_init(unsigned, unsigned, unsigned);
    static uvm_va_space_t spaces[64];
    uvm_access_counter_buffer_t ac;
    uvm_gpu_va_space_t gvs;
    uvm_processor_id_t gpu_id;
    uint32_t b, g, i, nsvc = 0, nrun = 0;
    oracle_prefetch_init(pf_enable, pf_threshold, pf_min_faults);
    g_state = st; g_lay = lay; g_runs = runs; g_nrun = &nrun; g_run_cap = run_cap;
    g_ac_gpu_id = lay->gpu_id; g_ac_gpu.id.val = lay->gpu_id; gpu_id.val = lay->gpu_id;
    g_ac_cpu_chunk = cpu_chunk_2m ? UVM_PAGE_SIZE_2M : 4096;
    /* init_access_counter_config :328-353 for the configured granularity */
    ac.current_config.rm.granularity = granularity_2m ? UVM_ACCESS_COUNTER_GRANULARITY_2M : UVM_ACCESS_COUNTER_GRANULARITY_64K;
    ac.current_config.sub_granularity_region_size = granularity_2m ? (2ull << 20) / 32 : 64 << 10;
    ac.current_config.sub_granularity_regions_per_translation = granularity_2m ? 32 : 1;
    g_block_service_context.block_context = &g_va_block_context;
    for (b = 0; b < n_batches; b++) {
        const uint32_t base = b * B;
        for (g = 0; g < batch[b].num_block_groups; g++) {
            const ref_fault_group_t *grp = &groups[base + g];
            const uint32_t va_space = tbl[grp->ctx].va_space;
            const uint64_t off = grp->block_start - lay->va_base;
            uvm_va_block_t blk;
            uvm_page_mask_t accessed;
            ref_service_group_t *rec;
            uvm_processor_id_t cpu_id = { 0 };
            int w;
            uint64_t any_pte = 0, any_gpu = 0, any_cpu = 0;
            if (!(grp->block_start >= lay->va_base && (off >> 21) < lay->blocks_per_space && va_space < lay->n_spaces))
                continue;
            memset(&blk, 0, sizeof(blk));
            blk.slot = va_space * lay->blocks_per_space + (uint32_t)(off >> 21);
            blk.start = grp->block_start; blk.end = grp->block_start + (2ull << 20) - 1;
            blk.va_space = &spaces[va_space & 63];
            blk.managed_range = (void *)&blk.range_storage;
            blk.range_storage.policy.preferred_location = UVM_ID_INVALID;
            for (w = 0; w < 8; w++) {
                any_pte |= st->gpu_pte[0][8 * (size_t)blk.slot + w];
                any_gpu |= st->gpu_resident[8 * (size_t)blk.slot + w];
                any_cpu |= st->cpu_resident[0][8 * (size_t)blk.slot + w] | st->cpu_resident[1][8 * (size_t)blk.slot + w];
            }
            if (any_pte) uvm_processor_mask_set(&blk.mapped, gpu_id);
            if (any_gpu) uvm_processor_mask_set(&blk.resident, gpu_id);
            if (any_cpu) uvm_processor_mask_set(&blk.resident, cpu_id);
            rec = &svc[nsvc++];
            memset(rec, 0, sizeof(*rec));
            rec->slot = blk.slot; rec->batch = (uint16_t)b; rec->group = g; rec->hint_residency = 257; rec->run_offset = nrun;
            g_rec = rec;
            gvs.gpu = &g_ac_gpu; gvs.va_space = blk.va_space;
            uvm_page_mask_zero(&accessed);                               /* service_notifications_in_block :1378-1400 */
            for (i = 0; i < grp->block_faults; i++) {
                uvm_access_counter_buffer_entry_t e;
                e.address = ordered[base + grp->first_fault_index + i].address;
                e.sub_granularity = ordered[base + grp->first_fault_index + i].sub_granularity;
                expand_notification_block(&gvs, &blk, &ac, &g_va_block_context, &accessed, &e);
            }
            service_va_block_locked(gpu_id, &blk, NULL, &g_block_service_context, &accessed);
            {   /* whether the back half ran: it stores the final new_residency into the record */
                uint64_t any = 0;
                for (w = 0; w < 8; w++) any |= rec->new_residency[w];
                if (any)
                    rec->page_fault_count = (uint16_t)uvm_page_mask_weight(&g_block_service_context.per_processor_masks[gpu_id.val].new_residency);
            }
        }
    }
    if (n_runs)
        *n_runs = nrun;
    return nsvc;
}
