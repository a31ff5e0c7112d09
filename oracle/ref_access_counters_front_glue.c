/*
 * ref_access_counters_front_glue.c -- TEST INFRASTRUCTURE ONLY (oracle/_ref build -> libref_access_counters_front.so).
 *
 * Compiles the reference's access-counter FRONT HALF VERBATIM:
 *   kernel-open/nvidia-uvm/uvm_turing_access_counter_buffer.c
 *     :88-198    get_access_counter_inst_aperture, get_address, get_access_counter_buffer_entry, entry_is_valid,
 *                entry_clear_valid, uvm_hal_turing_access_counter_buffer_parse_entry
 *   kernel-open/nvidia-uvm/uvm_gpu_access_counters.c
 *     :798-821   cmp_access_counter_instance_ptr, cmp_sort_notifications_by_instance_ptr
 *     :834-860   cmp_sort_notifications_by_va_space_gpu_address, notification_fetch_mode_t
 *                (cmp_gpu :824-830 is byte-identical to the copy of uvm_gpu_replayable_faults.c:713-719 that the fault
 *                leg compiles; it is taken from there)
 *     :862-940   fetch_access_counter_buffer_entries
 *     :942-969   translate_notifications_instance_ptrs       :974-992  preprocess_notifications
 *   kernel-open/nvidia-uvm/uvm_hal_types.h :481-520  uvm_access_counter_buffer_entry_t
 * on top of the fault leg's shims (ref_faults_glue.c: parent GPU, canonical address, the Hopper get_ve_id, a stable
 * sort()).  Stand-ins: the instance-pointer translation is the harness's table lookup; write_get records GET.
 *
 * What this pins: N3's packet parse, fetch (stop at the first entry that is not ready, VALID cleared), translation of
 * untranslatable entries to a NULL va_space, and the order of the sorted notifications.
 */
#include "ref_faults_glue.c"
#include "clc365.h"

typedef struct { NvU8 bufferEntry[NVC365_NOTIFY_BUF_SIZE]; } access_counter_buffer_entry_c365_t;   /* turing file :30-32 */
typedef struct uvm_access_counter_buffer_entry_struct uvm_access_counter_buffer_entry_t;
#include "x_ac_entry.h"         /* uvm_hal_types.h:481-520 */

typedef struct {                /* uvm_access_counter_service_batch_context_t: the fields the extracts touch (uvm_gpu.h) */
    uvm_access_counter_buffer_entry_t *notification_cache;
    uvm_access_counter_buffer_entry_t **notifications;
    NvU32 num_notifications;
    bool is_single_instance_ptr;
} uvm_access_counter_service_batch_context_t;
typedef struct uvm_access_counter_buffer_struct {
    uvm_parent_gpu_t *parent_gpu;
    NvU32 index, max_notifications, max_batch_size, cached_get, cached_put;
    struct { void *bufferAddress; NvU32 *pAccessCntrBufferPut; } rm_info;
    uvm_access_counter_service_batch_context_t batch_service_context;
    NvU32 written_get;          /* harness */
} uvm_access_counter_buffer_t;
struct uvm_access_counter_buffer_hal_struct {
    bool (*entry_is_valid)(uvm_access_counter_buffer_t *, NvU32);
    void (*parse_entry)(uvm_access_counter_buffer_t *, NvU32, uvm_access_counter_buffer_entry_t *);
};
#ifndef UVM_GPU_READ_ONCE
#define UVM_GPU_READ_ONCE(x) (*(volatile __typeof__(x) *)&(x))
#endif
/* write_get is a static function of each of the two reference files; here both live in one translation unit */
static void ac_write_get(uvm_access_counter_buffer_t *ac, NvU32 get) { ac->cached_get = get; ac->written_get = get; }

/* uvm_parent_gpu_access_counter_entry_to_va_space (uvm_gpu.c): the harness's table, first matching row wins */
static const ref_instance_ptr_entry_t *g_ac_tbl;
static uint32_t g_ac_rows;
static uvm_va_space_t g_ac_spaces[256];
static uvm_gpu_t g_ac_gpus[256];
static NV_STATUS uvm_parent_gpu_access_counter_entry_to_va_space(uvm_parent_gpu_t *p, const uvm_access_counter_buffer_entry_t *e,
                                                                 uvm_va_space_t **va_space, uvm_gpu_t **gpu);

#include "x_ac_turing.h"        /* uvm_turing_access_counter_buffer.c:88-198 */
#define write_get ac_write_get
#include "x_ac_front.h"         /* uvm_gpu_access_counters.c:798-821, 834-860, 862-940, 942-969, 974-992 */
#undef write_get

static uvm_access_counter_buffer_entry_t *g_ac_cache;
static uint16_t g_ac_row_of[1024];
static NV_STATUS uvm_parent_gpu_access_counter_entry_to_va_space(uvm_parent_gpu_t *p, const uvm_access_counter_buffer_entry_t *e,
                                                                 uvm_va_space_t **va_space, uvm_gpu_t **gpu)
{
    uint32_t k;
    (void)p;
    *va_space = NULL;
    *gpu = NULL;
    for (k = 0; k < g_ac_rows; k++)
        if (g_ac_tbl[k].instance_ptr == e->instance_ptr.address && g_ac_tbl[k].aperture == (uint8_t)e->instance_ptr.aperture &&
            (g_ac_tbl[k].ve_id == REF_VE_ID_ANY || g_ac_tbl[k].ve_id == e->ve_id)) {
            *va_space = &g_ac_spaces[g_ac_tbl[k].va_space & 255];
            *gpu = &g_ac_gpus[g_ac_tbl[k].gpu_id & 255];
            (*gpu)->id.val = g_ac_tbl[k].gpu_id;
            g_ac_row_of[e - g_ac_cache] = (uint16_t)k;
            return NV_OK;
        }
    return NV_ERR_INVALID_CHANNEL;
}

/* ======================= flat harness API ======================= */
typedef struct { uint64_t address; uint16_t index, ctx; uint32_t sub_granularity; } ref_ordered_notification_t;
typedef struct {
    uint64_t address, instance_ptr;
    uint32_t instance_aperture, mmu_engine_id, ve_id, counter_value, sub_granularity, bank, tag, valid;
} ref_ac_entry_t;

/* Same contract as oracle_access_counters_fetch_preprocess; entries[] is filled for fetched notifications only. */
uint32_t ref_access_counters_fetch_preprocess(uint32_t *ring, uint32_t ring_entries, uint32_t get, uint32_t put, uint32_t B,
                                              unsigned va_bits, const ref_instance_ptr_entry_t *tbl, uint32_t n_rows,
                                              int clear_valid, ref_ordered_notification_t *ordered, ref_fault_group_t *groups,
                                              ref_fault_batch_t *batch, ref_ac_entry_t *entries)
{
    static struct uvm_access_counter_buffer_hal_struct hal = { uvm_hal_turing_access_counter_buffer_entry_is_valid,
                                                               uvm_hal_turing_access_counter_buffer_parse_entry };
    const uint32_t n_total = put >= get ? put - get : ring_entries - get + put, nb = (n_total + B - 1) / B;
    uvm_parent_gpu_t gpu;
    uvm_access_counter_buffer_t ac;
    uint32_t b, i, *saved = NULL;
    memset(&gpu, 0, sizeof(gpu));
    memset(&ac, 0, sizeof(ac));
    gpu.va_bits = va_bits;
    gpu.fault_buffer_hal = &g_hal;
    gpu.access_counter_buffer_hal = &hal;
    ac.parent_gpu = &gpu;
    ac.max_notifications = ring_entries;
    ac.max_batch_size = B;
    ac.rm_info.bufferAddress = ring;
    ac.rm_info.pAccessCntrBufferPut = &ac.cached_put;
    ac.batch_service_context.notification_cache = (uvm_access_counter_buffer_entry_t *)calloc(B, sizeof(uvm_access_counter_buffer_entry_t));
    ac.batch_service_context.notifications = (uvm_access_counter_buffer_entry_t **)calloc(B, sizeof(void *));
    g_ac_cache = ac.batch_service_context.notification_cache;
    g_ac_tbl = tbl; g_ac_rows = n_rows;
    if (!clear_valid) {                                   /* parse_entry always clears VALID: put the words back afterwards */
        saved = (uint32_t *)malloc((size_t)ring_entries * 32);
        memcpy(saved, ring, (size_t)ring_entries * 32);
    }
    for (b = 0; b < nb; b++) {
        const uint32_t base = b * B, in_batch = n_total - base < B ? n_total - base : B;
        uvm_access_counter_service_batch_context_t *bc = &ac.batch_service_context;
        uint32_t n, ng = 0;
        memset(&batch[b], 0, sizeof(batch[b]));
        ac.cached_get = (get + base) % ring_entries;
        ac.cached_put = (get + base + in_batch) % ring_entries;
        bc->num_notifications = 0;
        bc->is_single_instance_ptr = true;
        /* the reference spins until the first entry of a fetch is ready (:893-907); a batch that starts with an
         * entry that is not ready is empty here */
        n = (ring[8 * (size_t)ac.cached_get + 7] >> 31) ? fetch_access_counter_buffer_entries(&ac, NOTIFICATION_FETCH_MODE_BATCH_READY) : 0;
        if (n)
            preprocess_notifications(&gpu, bc);
        for (i = 0; i < n; i++) {
            const uvm_access_counter_buffer_entry_t *e = bc->notifications[i], *c = &bc->notification_cache[i];
            ref_ordered_notification_t *o = &ordered[base + i];
            o->address = e->address; o->index = (uint16_t)(e - bc->notification_cache); o->sub_granularity = e->sub_granularity;
            o->ctx = 0xFFFF;
            if (e->va_space) {                            /* the table row behind the translation (the reference reuses the
                                                             previous entry's translation when the instance pointers match) */
                uvm_va_space_t *vs; uvm_gpu_t *gp;
                uvm_parent_gpu_access_counter_entry_to_va_space(&gpu, e, &vs, &gp);
                o->ctx = g_ac_row_of[o->index];
            }
            if (entries) {
                ref_ac_entry_t *x = &entries[base + i];
                x->address = c->address; x->instance_ptr = c->instance_ptr.address; x->instance_aperture = c->instance_ptr.aperture;
                x->mmu_engine_id = c->mmu_engine_id; x->ve_id = c->ve_id; x->counter_value = c->counter_value;
                x->sub_granularity = c->sub_granularity; x->bank = c->bank; x->tag = c->tag; x->valid = 1;
            }
            if (!e->va_space)
                continue;
            /* service_notifications_in_block :1381-1384: a run ends where va_space, gpu or the va_block changes */
            if (i == 0 || bc->notifications[i - 1]->va_space != e->va_space || bc->notifications[i - 1]->gpu != e->gpu ||
                (bc->notifications[i - 1]->address >> 21) != (e->address >> 21)) {
                ref_fault_group_t *g = &groups[base + ng++];
                g->block_start = e->address & ~((2ull << 20) - 1);
                g->ctx = o->ctx; g->first_fault_index = (uint16_t)i; g->block_faults = 0; g->flags = 0;
            }
            groups[base + ng - 1].block_faults++;
        }
        batch[b].num_cached_faults = n;
        batch[b].num_coalesced_faults = n;
        batch[b].is_single_instance_ptr = n ? bc->is_single_instance_ptr : 1;
        batch[b].num_block_groups = ng;
    }
    if (saved) { memcpy(ring, saved, (size_t)ring_entries * 32); free(saved); }
    free(ac.batch_service_context.notification_cache);
    free(ac.batch_service_context.notifications);
    return nb;
}
