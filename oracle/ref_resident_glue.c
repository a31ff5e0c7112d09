/*
 * ref_resident_glue.c -- TEST INFRASTRUCTURE ONLY (oracle/_ref build -> libref_resident.so).
 *
 * Compiles the reference's uvm_va_block_make_resident VERBATIM -- line-range extracts of
 * kernel-open/nvidia-uvm/uvm_va_block.c:
 *   :183-195    block_node_state_get, block_tracking_node_mask_get
 *   :584-593    block_update_cpu_resident_mask
 *   :609-701    uvm_va_block_cpu_set_resident_mask / _all_chunks, _clear_resident_mask / _region / _all_chunks
 *   :721-734    uvm_va_block_context_get_node
 *   :2406-2430  uvm_va_block_resident_mask_get
 *   :3569-3577  block_set_resident_processor
 *   :4379-4452  block_copy_resident_pages_from          :4460-4536  block_copy_resident_pages_mask
 *   :4538-4579  break_read_duplication_in_region        :4581-4619  block_copy_set_first_touch_residency
 *   :4625-4653  block_select_cpu_node_pages             :4660-4847  block_copy_resident_pages
 *   :4877-4976  uvm_va_block_make_resident_copy
 *   :4978-5114  block_make_resident_clear_evicted, block_make_resident_update_state,
 *               uvm_va_block_make_resident_finish, uvm_va_block_make_resident
 *   :5120-5162  block_prep_read_duplicate_mapping       :5164-5332  uvm_va_block_make_resident_read_duplicate
 * and of uvm_va_block_types.h (:116-129 cause enum, :149-159 page tracking, :197-242 the make_resident context,
 * :324-328 transfer mode), against stand-ins for what those functions call elsewhere:
 *   block_populate_pages                 -> CPU destination: the pages of page_mask & ~resident(dest node) are
 *                                           allocated on the destination node (the first choice of
 *                                           block_populate_pages_cpu, :1887-2018); GPU destination: nothing to record
 *   block_copy_resident_pages_between    -> its mask protocol verbatim in spirit (:4112-4117 early outs, :4368-4374
 *                                           copied_pages / migrated_pages), the CE pushes replaced by recording the
 *                                           copy runs (oracle_emit_runs)
 *   uvm_va_block_unmap_mask, _unmap      -> clears the GPU's pte bits for the mask (the only mapper in the table)
 *   uvm_va_block_revoke_prot             -> clears the GPU's WRITE and ATOMIC pte bits for the mask (CPU mappings are
 *                                           not part of the table); block_unmap_cpu -> nothing
 *   trackers, tools events, PMM marks, chunk checks -> no-ops
 *
 * What this pins: M1 / M1r / M2 / M4 -- which pages are copied from which processor / NUMA node in which order, what
 * counts as first touch, the residency, eviction and mapping state after the call -- for the three destinations of
 * the flat table (this GPU, CPU node 0 = DRAM, CPU node 1 = CXL).
 */
#include "shim/kshim.h"
#include "nvmisc.h"
#include "uvm_types.h"
#include "ref_api.h"

#include "x_va_block_types.h"   /* uvm_va_block_types.h:36-80 */

/* ---- processors (restated from uvm_processors.h:280-349,450-473): CPU = 0, the GPU of the table = its gpu_id ---- */
typedef struct { NvU32 val; } uvm_processor_id_t;
typedef uvm_processor_id_t uvm_gpu_id_t;
#define UVM_ID_MAX_PROCESSORS 257
#define UVM_ID_CPU ((uvm_processor_id_t){ .val = 0 })
#define UVM_ID_INVALID ((uvm_processor_id_t){ .val = UVM_ID_MAX_PROCESSORS })
static bool uvm_id_equal(uvm_processor_id_t a, uvm_processor_id_t b) { return a.val == b.val; }
#define UVM_ID_IS_CPU(id) uvm_id_equal(id, UVM_ID_CPU)
#define UVM_ID_IS_INVALID(id) uvm_id_equal(id, UVM_ID_INVALID)
#define UVM_ID_IS_GPU(id) (!UVM_ID_IS_CPU(id) && !UVM_ID_IS_INVALID(id))
typedef struct { DECLARE_BITMAP(bitmap, UVM_ID_MAX_PROCESSORS + 1); } uvm_processor_mask_t;
#define PMN (UVM_ID_MAX_PROCESSORS + 1)
static void uvm_processor_mask_zero(uvm_processor_mask_t *m) { bitmap_zero(m->bitmap, PMN); }
static void uvm_processor_mask_copy(uvm_processor_mask_t *d, const uvm_processor_mask_t *s) { bitmap_copy(d->bitmap, s->bitmap, PMN); }
static bool uvm_processor_mask_and(uvm_processor_mask_t *d, const uvm_processor_mask_t *a, const uvm_processor_mask_t *b) { return bitmap_and(d->bitmap, a->bitmap, b->bitmap, PMN); }
static bool uvm_processor_mask_andnot(uvm_processor_mask_t *d, const uvm_processor_mask_t *a, const uvm_processor_mask_t *b) { return bitmap_andnot(d->bitmap, a->bitmap, b->bitmap, PMN); }
static void uvm_processor_mask_set(uvm_processor_mask_t *m, uvm_processor_id_t id) { __set_bit(id.val, m->bitmap); }
static void uvm_processor_mask_clear(uvm_processor_mask_t *m, uvm_processor_id_t id) { __clear_bit(id.val, m->bitmap); }
static bool uvm_processor_mask_test(const uvm_processor_mask_t *m, uvm_processor_id_t id) { return test_bit(id.val, m->bitmap); }
static bool uvm_processor_mask_test_and_set(uvm_processor_mask_t *m, uvm_processor_id_t id) { return __test_and_set_bit(id.val, m->bitmap); }
static uvm_processor_id_t pm_next(const uvm_processor_mask_t *m, unsigned from)
{ uvm_processor_id_t r; r.val = (NvU32)find_next_bit(m->bitmap, UVM_ID_MAX_PROCESSORS, from); return r; }
#define for_each_id_in_mask(id, mask) for ((id) = pm_next(mask, 0); (id).val < UVM_ID_MAX_PROCESSORS; (id) = pm_next(mask, (id).val + 1))
#define for_each_gpu_id_in_mask(id, mask) for ((id) = pm_next(mask, 1); (id).val < UVM_ID_MAX_PROCESSORS; (id) = pm_next(mask, (id).val + 1))

/* ---- NUMA nodes: 0 = DRAM, 1 = CXL ---- */
#define NUMA_NO_NODE (-1)
#define MAX_NUMNODES 2
typedef struct { unsigned long bits; } nodemask_t;
static nodemask_t node_possible_map = { 3 };
#define nodes_clear(m) ((m).bits = 0)
#define nodes_empty(m) ((m).bits == 0)
#define node_set(n, m) ((m).bits |= 1UL << (n))
#define node_isset(n, m) ((((m).bits) >> (n)) & 1UL)
#define for_each_node_mask(n, m) for ((n) = 0; (n) < MAX_NUMNODES; (n)++) if (node_isset(n, m))
#define for_each_possible_uvm_node(n) for ((n) = 0; (n) < MAX_NUMNODES; (n)++)
/* closest node first: the start node, then the other one (uvm_va_block.c for_each_closest_uvm_node) */
#define for_each_closest_uvm_node(n, start, mask) for (int _k = 0; _k < MAX_NUMNODES && ((n) = _k == 0 ? (start) : 1 - (start), 1); _k++) if (node_isset(n, mask))
static size_t node_to_index(int nid) { UVM_ASSERT(nid >= 0 && nid < MAX_NUMNODES); return (size_t)nid; }
static int numa_mem_id(void) { return 0; }

/* ---- va_block: the fields the extracts touch (uvm_va_block.h:101-266, 274-520) ---- */
typedef struct uvm_va_space_struct { int dummy; } uvm_va_space_t;
typedef struct { uvm_processor_id_t preferred_location; int preferred_nid; } uvm_va_policy_t;
typedef struct { uvm_page_mask_t resident, evicted; uvm_page_mask_t pte_bits[3]; } uvm_va_block_gpu_state_t;
typedef struct { uvm_page_mask_t resident, allocated; } uvm_va_block_cpu_node_state_t;
typedef struct uvm_va_block_struct {
    NvU64 start, end;
    int lock;
    uvm_va_space_t *va_space;
    struct { uvm_va_policy_t policy; } *managed_range, range_storage;
    uvm_processor_mask_t resident, mapped, evicted_gpus;
    struct { uvm_page_mask_t resident; uvm_va_block_cpu_node_state_t *node_state[MAX_NUMNODES]; } cpu;
    uvm_va_block_gpu_state_t *gpus[UVM_ID_MAX_PROCESSORS];
    uvm_page_mask_t read_duplicated_pages, maybe_mapped_pages;
    int tracker;
    uint32_t slot;                                   /* harness: row of the flat state table */
} uvm_va_block_t;
typedef struct { uvm_va_block_t *va_block; uvm_va_block_region_t region; } uvm_reverse_map_t;
typedef struct { struct { NvU64 start, end; } node; } uvm_range_group_range_t;
static inline NvU64 uvm_va_block_size(uvm_va_block_t *b) { return b->end - b->start + 1; }
static inline size_t uvm_va_block_num_cpu_pages(uvm_va_block_t *b) { return uvm_va_block_size(b) / PAGE_SIZE; }
static uvm_page_index_t uvm_va_block_cpu_page_index(uvm_va_block_t *b, NvU64 addr) { return (addr - b->start) / PAGE_SIZE; }
static inline bool uvm_va_block_is_hmm(uvm_va_block_t *b) { (void)b; return false; }
static inline uvm_va_space_t *uvm_va_block_get_va_space(uvm_va_block_t *b) { return b->va_space; }
static inline NvU64 uvm_va_block_cpu_page_address(uvm_va_block_t *b, uvm_page_index_t i) { return b->start + PAGE_SIZE * (NvU64)i; }
static uvm_va_block_gpu_state_t *uvm_va_block_gpu_state_get(uvm_va_block_t *b, uvm_gpu_id_t id) { return b->gpus[id.val]; }
static const uvm_va_policy_t *uvm_va_policy_get_region(uvm_va_block_t *b, uvm_va_block_region_t r) { (void)r; return &b->managed_range->policy; }

#include "x_va_block_helpers.h" /* uvm_va_block.h: region + page-mask helpers, verbatim */

/* ---- contexts and enums, verbatim from uvm_va_block_types.h ---- */
#include "x_resident_types.h"   /* :116-129 uvm_make_resident_cause_t, :149-159 page tracking, :324-328 transfer mode */
typedef struct uvm_va_block_context_struct {
    uvm_page_mask_t scratch_page_mask;
    nodemask_t scratch_node_mask;
    void *mm;
#include "x_make_resident_ctx.h" /* :197-242: struct uvm_make_resident_context_struct { ... } make_resident; */
} uvm_va_block_context_t;
typedef struct uvm_va_block_retry_struct uvm_va_block_retry_t;
typedef struct { int dummy; } uvm_tracker_t;
#define UVM_TRACKER_INIT() { 0 }
static NV_STATUS uvm_tracker_add_tracker_safe(void *dst, uvm_tracker_t *src) { (void)dst; (void)src; return NV_OK; }
static void uvm_tracker_clear(uvm_tracker_t *t) { (void)t; }
static void uvm_tracker_deinit(uvm_tracker_t *t) { (void)t; }
#define uvm_assert_mutex_locked(l) ((void)0)

/* ---- harness state ---- */
static uvm_processor_mask_t g_can_copy_from, g_no_gpus, g_cache_mask[4];
static int g_cache_used;
static uvm_processor_mask_t *uvm_processor_mask_cache_alloc(void) { return &g_cache_mask[g_cache_used++ & 3]; }
static void uvm_processor_mask_cache_free(uvm_processor_mask_t *m) { (void)m; }
static const uvm_processor_mask_t *block_get_can_copy_from_mask(uvm_va_block_t *b, uvm_processor_id_t id) { (void)b; (void)id; return &g_can_copy_from; }
static const uvm_processor_mask_t *block_get_uvm_lite_gpus(uvm_va_block_t *b) { (void)b; return &g_no_gpus; }
static uvm_processor_id_t g_gpu_id;
/* closest first as seen from dst (uvm_processor_mask_find_closest_id, uvm_va_space.c:1879-1924, on a box where every
 * GPU has a fast link to every other and to the CPU): the processor itself, then the GPUs in ascending id ("direct
 * peers, prioritizing GPU peers over CPU"), then the CPU */
static uvm_processor_id_t closest_next(const uvm_processor_mask_t *m, uvm_processor_id_t dst, int step)
{
    uvm_processor_id_t id;
    int seen = 0;
    if (uvm_processor_mask_test(m, dst)) { if (step == 0) return dst; seen = 1; }
    for_each_gpu_id_in_mask(id, m) {
        if (uvm_id_equal(id, dst)) continue;
        if (seen == step) return id;
        seen++;
    }
    if (!UVM_ID_IS_CPU(dst) && uvm_processor_mask_test(m, UVM_ID_CPU) && seen == step)
        return UVM_ID_CPU;
    return UVM_ID_INVALID;
}
#define for_each_closest_id(id, mask, dst, va_space) \
    for (int _s = 0; ((id) = closest_next(mask, dst, _s)).val < UVM_ID_MAX_PROCESSORS; _s++)

static void block_mark_memory_used(uvm_va_block_t *b, uvm_processor_id_t id) { (void)b; (void)id; }
static bool block_check_chunks(uvm_va_block_t *b) { (void)b; return true; }
static bool block_check_cpu_chunks(uvm_va_block_t *b) { (void)b; return true; }
static void block_cleanup_temp_pinned_gpu_chunks(uvm_va_block_t *b, uvm_gpu_id_t id) { (void)b; (void)id; }
static bool block_processor_page_is_populated(uvm_va_block_t *b, uvm_processor_id_t id, uvm_page_index_t p) { (void)b; (void)id; (void)p; return true; }
static bool block_check_resident_proximity(uvm_va_block_t *b, uvm_va_block_context_t *c, uvm_page_index_t p, uvm_processor_id_t id) { (void)b; (void)c; (void)p; (void)id; return true; }
static void uvm_tools_record_read_duplicate_invalidate(uvm_va_block_t *b, uvm_processor_id_t id, uvm_va_block_region_t r, const uvm_page_mask_t *m) { (void)b; (void)id; (void)r; (void)m; }
static bool uvm_processor_mask_test_and_clear(uvm_processor_mask_t *m, uvm_processor_id_t id) { return __test_and_clear_bit(id.val, m->bitmap); }
/* block_clear_resident_processor (:3579-3603) without the PMM "root chunk unused" mark */
static void block_clear_resident_processor(uvm_va_block_t *b, uvm_processor_id_t id) { uvm_processor_mask_test_and_clear(&b->resident, id); }

/* ---- stand-ins defined after the extracts (they use functions the extracts define) ---- */
uvm_page_mask_t *uvm_va_block_resident_mask_get(uvm_va_block_t *block, uvm_processor_id_t processor, int nid);
static uvm_page_mask_t *block_resident_mask_get_alloc(uvm_va_block_t *block, uvm_processor_id_t processor, int nid);
static bool block_is_page_resident_anywhere(uvm_va_block_t *block, uvm_page_index_t page_index);
static NV_STATUS uvm_va_block_unmap_mask(uvm_va_block_t *va_block, uvm_va_block_context_t *ctx, const uvm_processor_mask_t *procs,
                                         uvm_va_block_region_t region, const uvm_page_mask_t *mask);
static NV_STATUS block_populate_pages(uvm_va_block_t *block, uvm_va_block_retry_t *retry, uvm_va_block_context_t *ctx,
                                      uvm_processor_id_t dest_id, uvm_va_block_region_t region, const uvm_page_mask_t *page_mask);
/* what the verbatim block_populate_pages (:3099-3182) calls */
typedef struct { int dummy; } uvm_memcg_context_t;
typedef struct uvm_gpu_struct uvm_gpu_t;
static uvm_gpu_t *uvm_gpu_get(uvm_gpu_id_t id) { (void)id; return NULL; }
static NV_STATUS block_populate_pages_gpu(uvm_va_block_t *b, uvm_va_block_retry_t *r, uvm_gpu_t *g, uvm_va_block_region_t region,
                                          const uvm_page_mask_t *m) { (void)b; (void)r; (void)g; (void)region; (void)m; return NV_OK; }
static void uvm_memcg_context_start(uvm_memcg_context_t *c, void *mm) { (void)c; (void)mm; }
static void uvm_memcg_context_end(uvm_memcg_context_t *c) { (void)c; }
static NV_STATUS block_populate_pages_cpu(uvm_va_block_t *block, const uvm_page_mask_t *populate_mask, uvm_va_block_region_t region,
                                          uvm_va_block_context_t *ctx, bool staged);
static NV_STATUS block_copy_resident_pages_between(uvm_va_block_t *block, uvm_va_block_context_t *block_context,
                                                   uvm_processor_id_t dst_id, int dst_nid, uvm_processor_id_t src_id, int src_nid,
                                                   uvm_va_block_region_t region, uvm_page_mask_t *copy_mask,
                                                   const uvm_page_mask_t *prefetch_page_mask,
                                                   uvm_va_block_transfer_mode_t transfer_mode, uvm_page_mask_t *migrated_pages,
                                                   NvU32 *copied_pages, uvm_tracker_t *copy_tracker);

typedef enum { UVM_PROT_NONE, UVM_PROT_READ_ONLY, UVM_PROT_READ_WRITE, UVM_PROT_READ_WRITE_ATOMIC, UVM_PROT_MAX } uvm_prot_t; /* uvm_hal_types.h:181-188 */
static NV_STATUS uvm_va_block_unmap(uvm_va_block_t *va_block, uvm_va_block_context_t *ctx, uvm_processor_id_t id,
                                    uvm_va_block_region_t region, const uvm_page_mask_t *mask, uvm_tracker_t *tracker);
static NV_STATUS uvm_va_block_revoke_prot(uvm_va_block_t *va_block, uvm_va_block_context_t *ctx, uvm_processor_id_t id,
                                          uvm_va_block_region_t region, const uvm_page_mask_t *mask, uvm_prot_t prot,
                                          uvm_tracker_t *tracker);
static void block_unmap_cpu(uvm_va_block_t *b, uvm_va_block_context_t *c, uvm_va_block_region_t r, const uvm_page_mask_t *m)
{ (void)b; (void)c; (void)r; (void)m; }
static void uvm_tools_record_read_duplicate(uvm_va_block_t *b, uvm_processor_id_t id, uvm_va_block_region_t r, const uvm_page_mask_t *m)
{ (void)b; (void)id; (void)r; (void)m; }
static bool uvm_va_block_is_dead(uvm_va_block_t *b) { (void)b; return false; }
static int g_uvm_page_mask_cache;
#define NV_UVM_GFP_FLAGS 0
static uvm_page_mask_t g_scratch_mask;
#define kmem_cache_alloc(cache, flags) ((void)(cache), &g_scratch_mask)
#define kmem_cache_free(cache, p) ((void)(cache), (void)(p))

#include "x_resident.h"         /* the uvm_va_block.c extracts listed at the top, verbatim */
#include "x_populate.h"         /* uvm_va_block.c:3099-3182 block_populate_pages, verbatim */

static uvm_page_mask_t *block_resident_mask_get_alloc(uvm_va_block_t *block, uvm_processor_id_t processor, int nid)
{
    return uvm_va_block_resident_mask_get(block, processor, nid);     /* GPU state always exists in the table */
}
static bool block_is_page_resident_anywhere(uvm_va_block_t *block, uvm_page_index_t page_index)
{
    uvm_processor_id_t id;
    if (uvm_page_mask_test(&block->cpu.resident, page_index))
        return true;
    for_each_gpu_id_in_mask(id, &block->resident)
        if (uvm_page_mask_test(&block->gpus[id.val]->resident, page_index))
            return true;
    return false;
}
/* uvm_va_block_unmap_mask (:10580-10660): the GPU of the table is the only mapper */
static uint32_t g_unmapped;
static NV_STATUS uvm_va_block_unmap_mask(uvm_va_block_t *va_block, uvm_va_block_context_t *ctx, const uvm_processor_mask_t *procs,
                                         uvm_va_block_region_t region, const uvm_page_mask_t *mask)
{
    uvm_va_block_gpu_state_t *gs = va_block->gpus[g_gpu_id.val];
    uvm_page_mask_t hit;
    int k;
    (void)ctx; (void)region;
    if (!uvm_processor_mask_test(procs, g_gpu_id))
        return NV_OK;
    if (uvm_page_mask_and(&hit, mask, &gs->pte_bits[0]))
        g_unmapped += uvm_page_mask_weight(&hit);
    for (k = 0; k < 3; k++)
        uvm_page_mask_andnot(&gs->pte_bits[k], &gs->pte_bits[k], mask);
    return NV_OK;
}
/* uvm_va_block_unmap (:10478-10578) / uvm_va_block_revoke_prot (:10662-10760) for one processor */
static NV_STATUS uvm_va_block_unmap(uvm_va_block_t *va_block, uvm_va_block_context_t *ctx, uvm_processor_id_t id,
                                    uvm_va_block_region_t region, const uvm_page_mask_t *mask, uvm_tracker_t *tracker)
{
    uvm_processor_mask_t one;
    uvm_page_mask_t clipped;
    (void)tracker;
    uvm_processor_mask_zero(&one);
    uvm_processor_mask_set(&one, id);
    uvm_page_mask_init_from_region(&clipped, region, mask);
    return uvm_va_block_unmap_mask(va_block, ctx, &one, region, &clipped);
}
static NV_STATUS uvm_va_block_revoke_prot(uvm_va_block_t *va_block, uvm_va_block_context_t *ctx, uvm_processor_id_t id,
                                          uvm_va_block_region_t region, const uvm_page_mask_t *mask, uvm_prot_t prot,
                                          uvm_tracker_t *tracker)
{
    uvm_va_block_gpu_state_t *gs = va_block->gpus[g_gpu_id.val];
    uvm_page_mask_t clipped;
    int k;
    (void)ctx; (void)tracker;
    if (!uvm_id_equal(id, g_gpu_id))
        return NV_OK;
    uvm_page_mask_init_from_region(&clipped, region, mask);
    for (k = (int)prot - 1; k < 3; k++)                               /* READ_WRITE revokes WRITE and ATOMIC */
        uvm_page_mask_andnot(&gs->pte_bits[k], &gs->pte_bits[k], &clipped);
    return NV_OK;
}
/* block_populate_pages_cpu (:1887-2018), first choice: the pages are allocated on the node the request names -- for a
 * GPU destination (staging pages) that is uvm_va_block_context_get_node's default, node 0 -- which is what
 * block_select_cpu_node_pages later finds in node_state->allocated */
static NV_STATUS block_populate_pages_cpu(uvm_va_block_t *block, const uvm_page_mask_t *populate_mask, uvm_va_block_region_t region,
                                          uvm_va_block_context_t *ctx, bool staged)
{
    const int nid = (!staged && ctx->make_resident.dest_nid != NUMA_NO_NODE) ? ctx->make_resident.dest_nid : numa_mem_id();
    uvm_va_block_cpu_node_state_t *ns = block_node_state_get(block, nid);
    uvm_page_mask_t populate;
    uvm_page_mask_init_from_region(&populate, region, populate_mask);
    uvm_page_mask_andnot(&populate, &populate, &ns->resident);
    uvm_page_mask_or(&ns->allocated, &ns->allocated, &populate);
    return NV_OK;
}

/* copy runs are recorded instead of pushed to a copy engine */
uint32_t oracle_emit_runs(const void *pages, const void *prefetch, uint32_t slot, unsigned src, unsigned dst,
                          ref_copy_run_t *runs, uint32_t at, uint32_t cap);
static ref_copy_run_t *g_runs;
static uint32_t g_nrun, g_run_cap, g_from[7];
static uvm_processor_id_t g_peer_id[4];
static uint32_t g_n_peers;
/* run source code of a processor: CPU node 0 / 1, this GPU 2, peer k 3 + k */
static unsigned src_code_of(uvm_processor_id_t id, int nid)
{
    uint32_t k;
    if (UVM_ID_IS_CPU(id)) return (unsigned)nid;
    if (uvm_id_equal(id, g_gpu_id)) return 2u;
    for (k = 0; k < g_n_peers; k++) if (uvm_id_equal(id, g_peer_id[k])) return 3u + k;
    return 2u;
}
static NV_STATUS block_copy_resident_pages_between(uvm_va_block_t *block, uvm_va_block_context_t *block_context,
                                                   uvm_processor_id_t dst_id, int dst_nid, uvm_processor_id_t src_id, int src_nid,
                                                   uvm_va_block_region_t region, uvm_page_mask_t *copy_mask,
                                                   const uvm_page_mask_t *prefetch_page_mask,
                                                   uvm_va_block_transfer_mode_t transfer_mode, uvm_page_mask_t *migrated_pages,
                                                   NvU32 *copied_pages, uvm_tracker_t *copy_tracker)
{
    uvm_page_mask_t *dst_resident_mask = uvm_va_block_resident_mask_get(block, dst_id, dst_nid);
    const uvm_make_resident_cause_t cause = block_context->make_resident.cause;
    const bool may_prefetch = (cause == UVM_MAKE_RESIDENT_CAUSE_REPLAYABLE_FAULT ||
                               cause == UVM_MAKE_RESIDENT_CAUSE_NON_REPLAYABLE_FAULT ||
                               cause == UVM_MAKE_RESIDENT_CAUSE_ACCESS_COUNTER) && !!prefetch_page_mask;   /* :4094-4096 */
    uvm_page_mask_t none;
    const unsigned src = src_code_of(src_id, src_nid), dst = UVM_ID_IS_CPU(dst_id) ? 1u + (unsigned)dst_nid : 0u;
    (void)region; (void)transfer_mode; (void)copy_tracker;
    *copied_pages = 0;
    if (!uvm_page_mask_andnot(copy_mask, copy_mask, dst_resident_mask))                  /* :4112-4113 */
        return NV_OK;
    if (migrated_pages && !uvm_page_mask_andnot(copy_mask, copy_mask, migrated_pages))   /* :4115-4116 */
        return NV_OK;
    uvm_page_mask_zero(&none);
    g_nrun += oracle_emit_runs(copy_mask->bitmap, may_prefetch ? prefetch_page_mask->bitmap : none.bitmap, block->slot, src, dst,
                               g_runs, g_nrun, g_run_cap);
    uvm_processor_mask_set(&block_context->make_resident.all_involved_processors, src_id);   /* block_copy_begin_push :3470-3476 */
    uvm_processor_mask_set(&block_context->make_resident.all_involved_processors, dst_id);
    *copied_pages = uvm_page_mask_weight(copy_mask);                                     /* :4368-4374 */
    g_from[src] += *copied_pages;
    if (*copied_pages && migrated_pages)
        uvm_page_mask_or(migrated_pages, migrated_pages, copy_mask);
    return NV_OK;
}

/* ======================= flat harness API ======================= */
static void load_mask(uvm_page_mask_t *m, const uint64_t *base, uint32_t slot) { memcpy(m->bitmap, base + 8 * (size_t)slot, 64); }
static void store_mask(uint64_t *base, uint32_t slot, const uvm_page_mask_t *m) { memcpy(base + 8 * (size_t)slot, m->bitmap, 64); }

/* Same contract as oracle_make_resident (oracle_resident.c). */
uint32_t ref_make_resident(uint32_t n_requests, const ref_make_resident_req_t *reqs, const uint64_t *page_masks,
                           const uint64_t *prefetch_masks, uint32_t n_slots, uint32_t gpu_id, const ref_block_state_t *st,
                           ref_make_resident_result_t *results, ref_copy_run_t *runs, uint32_t run_cap)
{
    static uvm_va_space_t space;
    uint32_t q;
    g_gpu_id.val = gpu_id;
    g_runs = runs; g_nrun = 0; g_run_cap = run_cap;
    uvm_processor_mask_zero(&g_can_copy_from);
    uvm_processor_mask_zero(&g_no_gpus);
    uvm_processor_mask_set(&g_can_copy_from, UVM_ID_CPU);
    uvm_processor_mask_set(&g_can_copy_from, g_gpu_id);
    g_n_peers = st->n_peers;
    for (q = 0; q < st->n_peers; q++) {
        g_peer_id[q].val = st->peer_gpu_id[q];
        if ((st->peer_direct >> q) & 1u)
            uvm_processor_mask_set(&g_can_copy_from, g_peer_id[q]);
    }

    for (q = 0; q < n_requests; q++) {
        const ref_make_resident_req_t *rq = &reqs[q];
        ref_make_resident_result_t *res = &results[q];
        const uint32_t slot = rq->slot;
        uvm_va_block_t blk;
        uvm_va_block_gpu_state_t gs, peer_gs[4];
        uvm_va_block_cpu_node_state_t ns[2];
        uvm_va_block_context_t ctx;
        uvm_page_mask_t tracking[2], *tracking_ptr[2] = { &tracking[0], &tracking[1] }, pm, pf, touch;
        uvm_va_block_region_t region;
        uvm_processor_id_t dest;
        int k;

        memset(res, 0, sizeof(*res));
        if (slot >= n_slots || rq->region_first > rq->region_outer || rq->region_outer > PAGES_PER_UVM_VA_BLOCK || rq->dest_tier > 2 ||
            ((rq->flags & 8) && (!st->read_duplicated || rq->cause == UVM_MAKE_RESIDENT_CAUSE_EVICTION || (rq->flags & 4) || st->n_peers))) {
            res->slot = 0xFFFFFFFFu;
            continue;
        }
        res->slot = slot;
        res->run_offset = g_nrun;
        memset(&blk, 0, sizeof(blk));
        memset(&ctx, 0, sizeof(ctx));
        memset(tracking, 0, sizeof(tracking));
        blk.slot = slot; blk.start = (NvU64)slot << 21; blk.end = blk.start + (1ull << 21) - 1;
        blk.va_space = &space;
        blk.managed_range = (void *)&blk.range_storage;
        blk.range_storage.policy.preferred_location = UVM_ID_INVALID;
        blk.range_storage.policy.preferred_nid = NUMA_NO_NODE;
        blk.gpus[gpu_id] = &gs;
        blk.cpu.node_state[0] = &ns[0]; blk.cpu.node_state[1] = &ns[1];
        load_mask(&gs.resident, st->gpu_resident, slot);
        for (k = 0; k < (int)st->n_peers; k++) {                      /* the other GPUs: residency only */
            memset(&peer_gs[k], 0, sizeof(peer_gs[k]));
            load_mask(&peer_gs[k].resident, st->peer_resident[k], slot);
            blk.gpus[g_peer_id[k].val] = &peer_gs[k];
            if (!uvm_page_mask_empty(&peer_gs[k].resident)) uvm_processor_mask_set(&blk.resident, g_peer_id[k]);
        }
        for (k = 0; k < 3; k++) load_mask(&gs.pte_bits[k], st->gpu_pte[k], slot);
        uvm_page_mask_zero(&gs.evicted);
        if (st->gpu_evicted) load_mask(&gs.evicted, st->gpu_evicted, slot);
        if (st->read_duplicated) load_mask(&blk.read_duplicated_pages, st->read_duplicated, slot);
        for (k = 0; k < 2; k++) {
            load_mask(&ns[k].resident, st->cpu_resident[k], slot);
            uvm_page_mask_copy(&ns[k].allocated, &ns[k].resident);    /* resident pages have chunks */
        }
        block_update_cpu_resident_mask(&blk);
        if (!uvm_page_mask_empty(&blk.cpu.resident)) uvm_processor_mask_set(&blk.resident, UVM_ID_CPU);
        if (!uvm_page_mask_empty(&gs.resident)) uvm_processor_mask_set(&blk.resident, g_gpu_id);
        uvm_processor_mask_set(&blk.mapped, g_gpu_id);                /* the unmap stand-in looks at the pte bits */
        if (!uvm_page_mask_empty(&gs.evicted)) uvm_processor_mask_set(&blk.evicted_gpus, g_gpu_id);

        region = uvm_va_block_region(rq->region_first, rq->region_outer);
        if (rq->flags & 1) load_mask(&pm, page_masks, rq->mask_row);
        if (rq->flags & 4) {                                          /* uvm_va_block_evict_chunks :13055-13080 */
            if (!(rq->flags & 1)) uvm_page_mask_init_from_region(&pm, region, NULL);
            uvm_page_mask_and(&pm, &pm, &gs.resident);
        }
        if (rq->flags & 2) load_mask(&pf, prefetch_masks, rq->mask_row);
        dest = rq->dest_tier == 0 ? g_gpu_id : UVM_ID_CPU;
        ctx.make_resident.dest_nid = rq->dest_tier == 0 ? NUMA_NO_NODE : (int)rq->dest_tier - 1;
        ctx.make_resident.cpu_pages_used.node_masks = tracking_ptr;
        g_unmapped = 0; memset(g_from, 0, sizeof(g_from));

        if (rq->flags & 8)
            uvm_va_block_make_resident_read_duplicate(&blk, NULL, &ctx, dest, region, (rq->flags & 1) ? &pm : NULL,
                                                      (rq->flags & 2) ? &pf : NULL, (uvm_make_resident_cause_t)rq->cause);
        else
            uvm_va_block_make_resident(&blk, NULL, &ctx, dest, region, (rq->flags & 5) ? &pm : NULL, (rq->flags & 2) ? &pf : NULL,
                                       (uvm_make_resident_cause_t)rq->cause);

        /* pages resident nowhere were populated (zeroed) at the destination: the first-touch runs come last */
        uvm_page_mask_andnot(&touch, &ctx.make_resident.pages_changed_residency, &ctx.make_resident.pages_migrated);
        {
            uvm_page_mask_t none;
            uvm_page_mask_zero(&none);
            g_nrun += oracle_emit_runs(touch.bitmap, (rq->flags & 2) ? pf.bitmap : none.bitmap, slot, 7, rq->dest_tier, runs, g_nrun, run_cap);
        }
        for (k = 0; k < 3; k++) res->migrated[k] = (uint16_t)g_from[k];
        for (k = 0; k < 4; k++) res->migrated_peer[k] = (uint16_t)g_from[3 + k];
        res->first_touch = (uint16_t)uvm_page_mask_weight(&touch);
        res->run_count = (uint16_t)(g_nrun - res->run_offset);
        res->unmapped = (uint16_t)g_unmapped;
        memcpy(res->pages_migrated, ctx.make_resident.pages_migrated.bitmap, 64);
        memcpy(res->pages_changed_residency, ctx.make_resident.pages_changed_residency.bitmap, 64);

        store_mask(st->gpu_resident, slot, &gs.resident);
        for (k = 0; k < (int)st->n_peers; k++) store_mask(st->peer_resident[k], slot, &peer_gs[k].resident);
        for (k = 0; k < 3; k++) store_mask(st->gpu_pte[k], slot, &gs.pte_bits[k]);
        if (st->gpu_evicted) store_mask(st->gpu_evicted, slot, &gs.evicted);
        if (st->read_duplicated) store_mask(st->read_duplicated, slot, &blk.read_duplicated_pages);
        for (k = 0; k < 2; k++) store_mask(st->cpu_resident[k], slot, &ns[k].resident);
    }
    return g_nrun;
}
