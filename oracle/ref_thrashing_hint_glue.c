/*
 * ref_thrashing_hint_glue.c -- TEST INFRASTRUCTURE ONLY (oracle/_ref build -> libref_thrashing_hint.so).
 *
 * Compiles the WHOLE thrashing heuristic of the reference verbatim -- line-range extracts of
 * kernel-open/nvidia-uvm/uvm_perf_thrashing.c, made by build_ref.sh:
 *   :36-202     bit widths, page / block / pinned-page / va_space tracking structures
 *   :376-394    page time stamp and throttling end time stamp get / set
 *   :721-1021   state checks, throttle update / start / end / reset, find_pinned_page, pin, unpin
 *   :1023-1093  thrashing_detected, thrashing_reset_page, thrashing_reset_pages_in_region
 *   :1172-1226  migrating_wrong_pages, is_migration_pinned_pages_update
 *   :1230-1380  thrashing_event_cb
 *   :1382-1780  the hint: thrashing_processors_*, get_hint_for_migration_thrashing, uvm_perf_thrashing_get_hint
 *   :1822-1906  thrashing_unpin_pages (the delayed worker that lets pins expire)
 * and uvm_va_space.c:1879-1934 (uvm_processor_mask_find_closest_id), uvm_perf_utils.h:30-43, against stand-ins for the
 * va_block, the va_space's processor masks, the policy, the perf-event payload, the slab cache, the lists' lock and
 * the work queue.  Unlike ref_thrashing_glue.c the tracking state lives in a context that survives between calls, so
 * events, hints and expiries can be interleaved the way a driver sees them.  NV_GETTIME() is the time stamp of the
 * event / query being replayed.
 *
 * What this pins: N2 -- detector and hint (throttle, pin, un-pin, epoch reset) -- field by field.
 */
#include "shim/kshim.h"
#include "ref_api.h"
#include <stdlib.h>

/* ---- processors and masks (uvm_processors.h) ---- */
typedef struct { NvU32 val; } uvm_processor_id_t;
typedef uvm_processor_id_t uvm_gpu_id_t;
typedef struct { NvU32 val; } uvm_parent_gpu_id_t;
#define UVM_ID_MAX_PROCESSORS 32
#define UVM_PARENT_ID_MAX_SUB_PROCESSORS 8
typedef struct { DECLARE_BITMAP(bitmap, UVM_ID_MAX_PROCESSORS); } uvm_processor_mask_t;
typedef struct { DECLARE_BITMAP(bitmap, 512); } uvm_page_mask_t;
typedef NvU16 uvm_page_index_t;
typedef struct { uvm_page_index_t first, outer; } uvm_va_block_region_t;
static const uvm_processor_id_t UVM_ID_INVALID = { 0xffff };
static const uvm_processor_id_t UVM_ID_CPU = { 0 };
#define UVM_ID_IS_CPU(id) ((id).val == 0)
#define UVM_ID_IS_INVALID(id) ((id).val == 0xffff)
#define UVM_ID_IS_VALID(id) (!UVM_ID_IS_INVALID(id))
#define UVM_ID_IS_GPU(id) (!UVM_ID_IS_CPU(id) && !UVM_ID_IS_INVALID(id))
static NvU32 uvm_id_value(uvm_processor_id_t id) { return id.val; }
static bool uvm_id_equal(uvm_processor_id_t a, uvm_processor_id_t b) { return a.val == b.val; }
static uvm_parent_gpu_id_t uvm_parent_gpu_id_from_gpu_id(uvm_gpu_id_t id) { uvm_parent_gpu_id_t p = { 1 + (id.val - 1) / UVM_PARENT_ID_MAX_SUB_PROCESSORS }; return p; }
static uvm_gpu_id_t uvm_gpu_id_from_sub_processor(uvm_parent_gpu_id_t id, NvU32 sub) { uvm_gpu_id_t g = { 1 + (id.val - 1) * UVM_PARENT_ID_MAX_SUB_PROCESSORS + sub }; return g; }
static void uvm_processor_mask_set(uvm_processor_mask_t *m, uvm_processor_id_t id) { __set_bit(id.val, m->bitmap); }
static void uvm_processor_mask_clear(uvm_processor_mask_t *m, uvm_processor_id_t id) { __clear_bit(id.val, m->bitmap); }
static bool uvm_processor_mask_test(const uvm_processor_mask_t *m, uvm_processor_id_t id) { return id.val < UVM_ID_MAX_PROCESSORS && test_bit(id.val, m->bitmap); }
static bool uvm_processor_mask_test_and_set(uvm_processor_mask_t *m, uvm_processor_id_t id) { return __test_and_set_bit(id.val, m->bitmap); }
static void uvm_processor_mask_zero(uvm_processor_mask_t *m) { bitmap_zero(m->bitmap, UVM_ID_MAX_PROCESSORS); }
static bool uvm_processor_mask_empty(const uvm_processor_mask_t *m) { return bitmap_empty(m->bitmap, UVM_ID_MAX_PROCESSORS); }
static void uvm_processor_mask_copy(uvm_processor_mask_t *d, const uvm_processor_mask_t *s) { *d = *s; }
static bool uvm_processor_mask_and(uvm_processor_mask_t *d, const uvm_processor_mask_t *a, const uvm_processor_mask_t *b)
{ return bitmap_and(d->bitmap, a->bitmap, b->bitmap, UVM_ID_MAX_PROCESSORS); }
static bool uvm_processor_mask_subset(const uvm_processor_mask_t *a, const uvm_processor_mask_t *b)
{ return bitmap_subset(a->bitmap, b->bitmap, UVM_ID_MAX_PROCESSORS); }
static void uvm_processor_mask_range_fill(uvm_processor_mask_t *m, uvm_processor_id_t start, NvU32 n)
{
    NvU32 i;
    for (i = 0; i < n && start.val + i < UVM_ID_MAX_PROCESSORS; i++) __set_bit(start.val + i, m->bitmap);
}
static uvm_processor_id_t uvm_processor_mask_find_first_id(const uvm_processor_mask_t *m)
{
    uvm_processor_id_t id = { (NvU32)find_first_bit(m->bitmap, UVM_ID_MAX_PROCESSORS) };
    return id.val >= UVM_ID_MAX_PROCESSORS ? UVM_ID_INVALID : id;
}
static uvm_processor_id_t uvm_processor_mask_find_first_gpu_id(const uvm_processor_mask_t *m)
{
    uvm_processor_id_t id = { (NvU32)find_next_bit(m->bitmap, UVM_ID_MAX_PROCESSORS, 1) };
    return id.val >= UVM_ID_MAX_PROCESSORS ? UVM_ID_INVALID : id;
}
#define for_each_id_in_mask(id, mask) \
    for ((id).val = (NvU32)find_first_bit((mask)->bitmap, UVM_ID_MAX_PROCESSORS); (id).val < UVM_ID_MAX_PROCESSORS; \
         (id).val = (NvU32)find_next_bit((mask)->bitmap, UVM_ID_MAX_PROCESSORS, (id).val + 1))
static bool uvm_page_mask_test(const uvm_page_mask_t *m, uvm_page_index_t p) { return test_bit(p, m->bitmap); }
static void uvm_page_mask_set(uvm_page_mask_t *m, uvm_page_index_t p) { __set_bit(p, m->bitmap); }
static void uvm_page_mask_clear(uvm_page_mask_t *m, uvm_page_index_t p) { __clear_bit(p, m->bitmap); }
static void uvm_page_mask_zero(uvm_page_mask_t *m) { bitmap_zero(m->bitmap, 512); }
static bool uvm_page_mask_empty(const uvm_page_mask_t *m) { return bitmap_empty(m->bitmap, 512); }
static bool uvm_page_mask_subset(const uvm_page_mask_t *a, const uvm_page_mask_t *b) { return bitmap_subset(a->bitmap, b->bitmap, 512); }
static bool uvm_page_mask_test_and_set(uvm_page_mask_t *m, uvm_page_index_t p) { return __test_and_set_bit(p, m->bitmap); }
static bool uvm_page_mask_test_and_clear(uvm_page_mask_t *m, uvm_page_index_t p) { return __test_and_clear_bit(p, m->bitmap); }
static bool uvm_page_mask_region_full(const uvm_page_mask_t *m, uvm_va_block_region_t r)
{
    uvm_page_index_t p;
    for (p = r.first; p < r.outer; p++) if (!test_bit(p, m->bitmap)) return false;
    return true;
}

/* ---- va_space, va_block, policy ---- */
typedef struct { int dummy; } uvm_spinlock_t;
typedef struct { int dummy; } uvm_mutex_t;
struct delayed_work { int dummy; };
struct work_struct { int dummy; };
typedef struct { uvm_processor_mask_t fast_access_mask, scratch_processor_mask; } uvm_va_block_context_t;
typedef struct { int read_duplication; uvm_processor_id_t preferred_location; } uvm_va_policy_t;
typedef struct { uvm_va_policy_t policy; } uvm_va_range_managed_t;
typedef struct uvm_va_space_struct uvm_va_space_t;
typedef struct uvm_va_block_struct {
    NvU64 start, end;
    uvm_va_range_managed_t *managed_range;
    uvm_va_range_managed_t range;                  /* one policy per va_block (preferred_location comes with the query) */
    uvm_va_space_t *va_space;
    void *thrashing;                               /* perf_modules_data[UVM_PERF_MODULE_TYPE_THRASHING] */
    uvm_mutex_t lock;
    uvm_processor_mask_t page_resident;            /* where the page being asked about lives (set per query) */
} uvm_va_block_t;
enum { UVM_READ_DUPLICATION_UNSET, UVM_READ_DUPLICATION_ENABLED, UVM_READ_DUPLICATION_DISABLED };
typedef int uvm_read_duplication_policy_t;
typedef enum {
    UVM_MAKE_RESIDENT_CAUSE_REPLAYABLE_FAULT, UVM_MAKE_RESIDENT_CAUSE_NON_REPLAYABLE_FAULT, UVM_MAKE_RESIDENT_CAUSE_ACCESS_COUNTER,
    UVM_MAKE_RESIDENT_CAUSE_PREFETCH, UVM_MAKE_RESIDENT_CAUSE_EVICTION, UVM_MAKE_RESIDENT_CAUSE_API_TOOLS,
    UVM_MAKE_RESIDENT_CAUSE_API_MIGRATE, UVM_MAKE_RESIDENT_CAUSE_API_SET_RANGE_GROUP, UVM_MAKE_RESIDENT_CAUSE_API_HINT
} uvm_make_resident_cause_t;
typedef enum { UVM_VA_BLOCK_TRANSFER_MODE_MOVE = 1, UVM_VA_BLOCK_TRANSFER_MODE_COPY = 2 } uvm_va_block_transfer_mode_t;
typedef struct { uvm_processor_id_t dest_id; } uvm_make_resident_context_t;
typedef enum { UVM_PERF_EVENT_MIGRATION = 3, UVM_PERF_EVENT_REVOCATION = 4 } uvm_perf_event_t;
typedef union {
    struct {
        uvm_va_block_t *block; NvU64 address, bytes; uvm_processor_id_t dst; uvm_make_resident_cause_t cause;
        uvm_make_resident_context_t *make_resident_context; uvm_va_block_transfer_mode_t transfer_mode;
    } migration;
    struct { uvm_va_block_t *block; NvU64 address, bytes; uvm_processor_id_t proc_id; } revocation;
} uvm_perf_event_data_t;
typedef enum { UVM_PERF_THRASHING_HINT_TYPE_NONE = 0, UVM_PERF_THRASHING_HINT_TYPE_PIN = 1, UVM_PERF_THRASHING_HINT_TYPE_THROTTLE = 2 } uvm_perf_thrashing_hint_type_t;
typedef struct {
    uvm_perf_thrashing_hint_type_t type;
    union {
        struct { uvm_processor_id_t residency; uvm_processor_mask_t processors; } pin;
        struct { NvU64 end_time_stamp; } throttle;
    };
} uvm_perf_thrashing_hint_t;

static NvU64 g_now;
static NvU64 g_detections, g_resets, g_throttles, g_pins_local, g_pins_remote;
static bool g_uvm_perf_thrashing_enable = true;
static uvm_processor_mask_t g_has_memory;
#define NV_GETTIME() g_now
#define NV_UVM_GFP_FLAGS 0
#define uvm_kvmalloc_zero(n) calloc(1, n)
#define uvm_kvfree(p) free(p)
static struct kmem_cache *g_pinned_page_cache;
#define nv_kmem_cache_zalloc(cache, flags) calloc(1, sizeof(pinned_page_t))
#define kmem_cache_free(cache, p) free(p)
#define uvm_spin_lock(l) ((void)(l))
#define uvm_spin_unlock(l) ((void)(l))
#define uvm_mutex_lock(l) ((void)(l))
#define uvm_mutex_unlock(l) ((void)(l))
#define uvm_va_space_down_read(vs) ((void)(vs))
#define uvm_va_space_up_read(vs) ((void)(vs))
static inline int schedule_delayed_work(struct delayed_work *w, unsigned long j) { (void)w; (void)j; return 1; }
#define cancel_delayed_work(w) ((void)(w))
#define usecs_to_jiffies(u) (u)
#define to_delayed_work(w) ((struct delayed_work *)(w))
#define uvm_hmm_migrate_begin_wait(b) ((void)(b))
#define uvm_hmm_migrate_finish(b) ((void)(b))
#define uvm_va_block_context_init(c, mm) ((void)(c))
#define uvm_perf_thrashing_unmap_remote_pinned_pages_all(b, c, r) ((void)0)      /* PTE work: out of scope */
#define uvm_va_block_region_for_page(p) (p)
#define container_of(ptr, type, member) ((type *)((char *)(ptr) - offsetof(type, member)))
#define list_entry(ptr, type, member) container_of(ptr, type, member)
#define list_first_entry_or_null(head, type, member) (list_empty(head) ? NULL : list_entry((head)->next, type, member))
#define list_for_each_entry(pos, head, member) \
    for (pos = list_entry((head)->next, typeof(*pos), member); &pos->member != (head); pos = list_entry(pos->member.next, typeof(*pos), member))
static inline void list_del_init(struct list_head *e) { e->prev->next = e->next; e->next->prev = e->prev; INIT_LIST_HEAD(e); }
static inline int list_is_singular(const struct list_head *h) { return !list_empty(h) && h->next == h->prev; }
#define PROCESSOR_THRASHING_STATS_INC(proc, field) ((void)(proc), g_stat_##field++)
static NvU64 g_stat_num_thrashing, g_stat_num_throttle, g_stat_num_pin_local, g_stat_num_pin_remote;
#define uvm_tools_record_thrashing(vs, addr, size, procs) ((void)(vs), (void)(addr), (void)(procs))
#define uvm_tools_record_throttling_start(vs, addr, proc) ((void)(vs), (void)(addr), (void)(proc))
#define uvm_tools_record_throttling_end(vs, addr, proc) ((void)(vs), (void)(addr), (void)(proc))
static uvm_va_space_t *uvm_va_block_get_va_space(uvm_va_block_t *b) { return b->va_space; }
static NvU64 uvm_va_block_cpu_page_address(uvm_va_block_t *b, uvm_page_index_t p) { return b->start + (NvU64)p * PAGE_SIZE; }
static uvm_page_index_t uvm_va_block_cpu_page_index(uvm_va_block_t *b, NvU64 addr) { return (uvm_page_index_t)((addr - b->start) / PAGE_SIZE); }
static NvU64 uvm_va_block_size(uvm_va_block_t *b) { return b->end - b->start + 1; }
static bool uvm_va_block_is_hmm(uvm_va_block_t *b) { (void)b; return false; }
static const uvm_va_policy_t *uvm_va_policy_get(uvm_va_block_t *b, NvU64 addr) { (void)addr; return &b->managed_range->policy; }
static bool uvm_processor_has_memory(uvm_processor_id_t id) { return uvm_processor_mask_test(&g_has_memory, id); }
static uvm_va_block_region_t uvm_va_block_region_from_start_size(uvm_va_block_t *b, NvU64 address, NvU64 bytes)
{
    uvm_va_block_region_t r;
    r.first = (uvm_page_index_t)((address - b->start) / PAGE_SIZE);
    r.outer = (uvm_page_index_t)(r.first + bytes / PAGE_SIZE);
    return r;
}
#define for_each_va_block_page_in_region(p, region) for ((p) = (region).first; (p) < (region).outer; (p)++)
#define for_each_va_block_page_in_region_mask(p, mask, region) \
    for ((p) = (region).first; (p) < (region).outer; (p)++) if (uvm_page_mask_test(mask, p))
#define for_each_va_block_page_in_mask(p, mask, b) for ((p) = 0; (p) < 512; (p)++) if (uvm_page_mask_test(mask, p))

#include "x_perf_saturating.h"      /* uvm_perf_utils.h:30-43 */
#include "x_th_types.h"             /* uvm_perf_thrashing.c:36-202 */

struct uvm_va_space_struct {
    va_space_thrashing_info_t thrashing;
    uvm_processor_mask_t accessible_from[UVM_ID_MAX_PROCESSORS], can_access[UVM_ID_MAX_PROCESSORS];
    uvm_processor_mask_t has_fast_link[UVM_ID_MAX_PROCESSORS], has_native_atomics[UVM_ID_MAX_PROCESSORS];
    struct { uvm_processor_mask_t mask; uvm_mutex_t mask_mutex; } closest_processors;
};
static va_space_thrashing_info_t *va_space_thrashing_info_get(uvm_va_space_t *vs) { return &vs->thrashing; }
static block_thrashing_info_t *thrashing_info_get(uvm_va_block_t *b) { return b->thrashing; }
static block_thrashing_info_t *thrashing_info_get_create(uvm_va_block_t *b)
{
    if (!b->thrashing) {                            /* :645-665 */
        block_thrashing_info_t *bt = calloc(1, sizeof(*bt));
        bt->last_processor = UVM_ID_INVALID;
        INIT_LIST_HEAD(&bt->pinned_pages.list);
        b->thrashing = bt;
    }
    return b->thrashing;
}

#include "x_vs_closest.h"           /* uvm_va_space.c:1879-1934 uvm_processor_mask_find_closest_id */

/* the page the hint is asked about lives on va_block->page_resident (set per query by the harness) */
static void uvm_va_block_page_resident_processors(uvm_va_block_t *b, uvm_page_index_t p, uvm_processor_mask_t *out) { (void)p; *out = b->page_resident; }
static uvm_processor_id_t uvm_va_block_page_get_closest_resident(uvm_va_block_t *b, uvm_va_block_context_t *c, uvm_page_index_t p,
                                                                uvm_processor_id_t processor)
{
    (void)c; (void)p;
    if (uvm_processor_mask_empty(&b->page_resident))
        return UVM_ID_INVALID;
    return uvm_processor_mask_find_closest_id(b->va_space, &b->page_resident, processor);   /* == for_each_closest_id's first hit */
}

#include "x_th_stamps.h"            /* :376-394 */
#include "x_th_throttle_pin.h"      /* :721-1021 */
#define thrashing_reset_page thrashing_reset_page_ref
#include "x_th_detect.h"            /* :1023-1073 */
#undef thrashing_reset_page
static void thrashing_reset_page(va_space_thrashing_info_t *vs, uvm_va_block_t *b, block_thrashing_info_t *bt, uvm_page_index_t p)
{
    g_resets++;
    thrashing_reset_page_ref(vs, b, bt, p);
}
#include "x_th_region.h"            /* :1077-1093 */
#include "x_th_pinned_update.h"     /* :1172-1226 */
void thrashing_event_cb(uvm_perf_event_t event_id, uvm_perf_event_data_t *event_data);
#include "x_th_event.h"             /* :1230-1380 */
#include "x_th_hint.h"              /* :1382-1780 */
#include "x_th_unpin_worker.h"      /* :1822-1906 */

/* ---- harness ---- */
typedef struct { uint64_t time_stamp; uint32_t slot; uint16_t first_page, num_pages; uint16_t processor; uint8_t kind, flags; uint32_t reserved; } h_event_t;
typedef struct { uint64_t last_time_stamp, last_thrashing_time_stamp; uint16_t last_processor, num_thrashing_pages; uint32_t tracked;
                 uint16_t pinned_count; uint8_t reset_count, pad; uint32_t throttling_count; } h_block_t;
typedef struct { uint64_t lapse_ns, nap_ns, epoch_ns, pin_ns; uint32_t threshold, pin_threshold, max_resets, has_memory;
                 uint32_t accessible_from[32], can_access[32], has_fast_link[32], has_native_atomics[32]; } h_cfg_t;
typedef struct { uint64_t time_stamp; uint32_t slot; uint16_t page; uint8_t requester, pad; uint32_t resident_processors;
                 uint16_t preferred_location, pad2; } h_query_t;
typedef struct { uint64_t throttle_end; uint32_t pin_processors; uint8_t type, pin_residency; uint16_t pad; } h_hint_t;

typedef struct { uvm_va_space_t vs; uvm_va_block_t *vb; uint32_t n_slots; uvm_va_block_context_t ctx; } h_ctx_t;

static void set_mask(uvm_processor_mask_t *m, uint32_t bits) { uvm_processor_mask_zero(m); m->bitmap[0] = bits; }

void *ref_thrash_open(uint32_t n_slots, const h_cfg_t *cfg)
{
    h_ctx_t *h = calloc(1, sizeof(*h));
    uint32_t s, i;
    h->n_slots = n_slots;
    h->vb = calloc(n_slots, sizeof(*h->vb));
    h->vs.thrashing.params.enable = true;
    h->vs.thrashing.params.threshold = cfg->threshold;
    h->vs.thrashing.params.pin_threshold = cfg->pin_threshold;
    h->vs.thrashing.params.lapse_ns = cfg->lapse_ns;
    h->vs.thrashing.params.nap_ns = cfg->nap_ns;
    h->vs.thrashing.params.epoch_ns = cfg->epoch_ns;
    h->vs.thrashing.params.max_resets = cfg->max_resets;
    h->vs.thrashing.params.pin_ns = cfg->pin_ns;
    h->vs.thrashing.va_space = &h->vs;
    h->vs.thrashing.pinned_pages.va_block_context = &h->ctx;
    INIT_LIST_HEAD(&h->vs.thrashing.pinned_pages.list);
    for (i = 0; i < 32; i++) {
        set_mask(&h->vs.accessible_from[i], cfg->accessible_from[i]);
        set_mask(&h->vs.can_access[i], cfg->can_access[i]);
        set_mask(&h->vs.has_fast_link[i], cfg->has_fast_link[i]);
        set_mask(&h->vs.has_native_atomics[i], cfg->has_native_atomics[i]);
    }
    set_mask(&g_has_memory, cfg->has_memory);
    for (s = 0; s < n_slots; s++) {
        uvm_va_block_t *b = &h->vb[s];
        b->start = (NvU64)s << 21; b->end = b->start + (2u << 20) - 1;
        b->range.policy.read_duplication = UVM_READ_DUPLICATION_DISABLED;
        b->range.policy.preferred_location = UVM_ID_INVALID;
        b->managed_range = &b->range; b->va_space = &h->vs;
    }
    g_resets = g_stat_num_thrashing = g_stat_num_throttle = g_stat_num_pin_local = g_stat_num_pin_remote = 0;
    return h;
}

void ref_thrash_events(void *hv, uint32_t n_events, const h_event_t *ev)
{
    h_ctx_t *h = hv;
    uvm_make_resident_context_t mrc;
    uint32_t i;
    set_mask(&g_has_memory, g_has_memory.bitmap[0]);
    for (i = 0; i < n_events; i++) {
        uvm_perf_event_data_t d;
        uvm_va_block_t *b = &h->vb[ev[i].slot];
        memset(&d, 0, sizeof(d));
        g_now = ev[i].time_stamp;
        if (ev[i].kind == 0) {
            d.migration.block = b; d.migration.address = b->start + (NvU64)ev[i].first_page * PAGE_SIZE;
            d.migration.bytes = (NvU64)ev[i].num_pages * PAGE_SIZE; d.migration.dst.val = ev[i].processor;
            d.migration.cause = (ev[i].flags & 1) ? UVM_MAKE_RESIDENT_CAUSE_API_MIGRATE
                              : (ev[i].flags & 2) ? UVM_MAKE_RESIDENT_CAUSE_PREFETCH : UVM_MAKE_RESIDENT_CAUSE_REPLAYABLE_FAULT;
            mrc.dest_id = d.migration.dst; d.migration.make_resident_context = &mrc;
            d.migration.transfer_mode = UVM_VA_BLOCK_TRANSFER_MODE_MOVE;
            thrashing_event_cb(UVM_PERF_EVENT_MIGRATION, &d);
        }
        else {
            d.revocation.block = b; d.revocation.address = b->start + (NvU64)ev[i].first_page * PAGE_SIZE;
            d.revocation.bytes = (NvU64)ev[i].num_pages * PAGE_SIZE; d.revocation.proc_id.val = ev[i].processor;
            thrashing_event_cb(UVM_PERF_EVENT_REVOCATION, &d);
        }
    }
}

void ref_thrash_hints(void *hv, uint32_t n, const h_query_t *q, h_hint_t *out)
{
    h_ctx_t *h = hv;
    uint32_t i;
    for (i = 0; i < n; i++) {
        uvm_va_block_t *b = &h->vb[q[i].slot];
        uvm_processor_id_t req = { q[i].requester };
        uvm_perf_thrashing_hint_t hint;
        g_now = q[i].time_stamp;
        set_mask(&b->page_resident, q[i].resident_processors);
        b->range.policy.preferred_location.val = q[i].preferred_location;
        hint = uvm_perf_thrashing_get_hint(b, &h->ctx, b->start + (NvU64)q[i].page * PAGE_SIZE, req);
        memset(&out[i], 0, sizeof(out[i]));
        out[i].type = (uint8_t)hint.type;
        out[i].pin_residency = 0xff;
        if (hint.type == UVM_PERF_THRASHING_HINT_TYPE_PIN) {
            out[i].pin_residency = (uint8_t)hint.pin.residency.val;
            out[i].pin_processors = (uint32_t)hint.pin.processors.bitmap[0];
        }
        else if (hint.type == UVM_PERF_THRASHING_HINT_TYPE_THROTTLE)
            out[i].throttle_end = hint.throttle.end_time_stamp;
    }
}

/* The delayed worker at time `now`: pins whose deadline has passed (20 us of timer slack, :1822,1849) are dropped. */
void ref_thrash_unpin_expired(void *hv, uint64_t now)
{
    h_ctx_t *h = hv;
    g_now = now;
    thrashing_unpin_pages((struct work_struct *)&h->vs.thrashing.pinned_pages.dwork);
}

void ref_thrash_export(void *hv, h_block_t *blocks, uint64_t *page_word, uint32_t *page_procs, uint32_t *thrashing_pages,
                       uint64_t *page_throttle, uint32_t *page_hint, uint32_t *page_throttled, uint64_t *page_deadline,
                       uint32_t *pinned_pages, uint32_t *counters)
{
    h_ctx_t *h = hv;
    uint32_t s, p;
    counters[0] = (uint32_t)g_stat_num_thrashing; counters[1] = (uint32_t)g_resets; counters[2] = 0;
    counters[3] = (uint32_t)g_stat_num_throttle; counters[4] = (uint32_t)g_stat_num_pin_local; counters[5] = (uint32_t)g_stat_num_pin_remote;
    for (s = 0; s < h->n_slots; s++) {
        block_thrashing_info_t *bt = h->vb[s].thrashing;
        memset(&blocks[s], 0, sizeof(blocks[s]));
        for (p = 0; p < 16; p++) thrashing_pages[(size_t)s * 16 + p] = pinned_pages[(size_t)s * 16 + p] = 0;
        for (p = 0; p < 512; p++) {
            const size_t at = (size_t)s * 512 + p;
            page_word[at] = 0; page_procs[at] = 0; page_throttle[at] = 0; page_hint[at] = 0; page_throttled[at] = 0; page_deadline[at] = 0;
        }
        if (!bt)
            continue;
        blocks[s].last_time_stamp = bt->last_time_stamp;
        blocks[s].last_thrashing_time_stamp = bt->last_thrashing_time_stamp;
        blocks[s].last_processor = (uint16_t)bt->last_processor.val;
        blocks[s].num_thrashing_pages = bt->num_thrashing_pages;
        blocks[s].tracked = bt->pages != NULL;
        blocks[s].pinned_count = (uint16_t)bt->pinned_pages.count;
        blocks[s].reset_count = bt->thrashing_reset_count;
        blocks[s].throttling_count = bt->throttling_count;
        for (p = 0; p < 16; p++) {
            thrashing_pages[(size_t)s * 16 + p] = (uint32_t)(bt->thrashing_pages.bitmap[p / 2] >> (32 * (p & 1)));
            pinned_pages[(size_t)s * 16 + p] = (uint32_t)(bt->pinned_pages.mask.bitmap[p / 2] >> (32 * (p & 1)));
        }
        if (bt->pages)
            for (p = 0; p < 512; p++) {
                page_thrashing_info_t *pt = &bt->pages[p];
                const size_t at = (size_t)s * 512 + p;
                pinned_page_t *pp;
                page_word[at] = (uint64_t)pt->last_time_stamp | ((uint64_t)pt->has_migration_events << 58) |
                                ((uint64_t)pt->has_revocation_events << 59) | ((uint64_t)pt->num_thrashing_events << 60) |
                                ((uint64_t)pt->pinned << 63);
                page_procs[at] = (uint32_t)pt->processors.bitmap[0];
                page_throttle[at] = pt->throttling_end_time_stamp;
                page_hint[at] = (uint32_t)pt->throttling_count |
                                ((UVM_ID_IS_INVALID(pt->pinned_residency_id) ? 0u : pt->pinned_residency_id.val + 1u) << 8) |
                                ((UVM_ID_IS_INVALID(pt->do_not_throttle_processor_id) ? 0u : pt->do_not_throttle_processor_id.val + 1u) << 16);
                page_throttled[at] = (uint32_t)pt->throttled_processors.bitmap[0];
                if (pt->pinned && h->vs.thrashing.params.pin_ns > 0 && (pp = find_pinned_page(bt, (uvm_page_index_t)p)))
                    page_deadline[at] = pp->deadline;
            }
    }
}

void ref_thrash_close(void *hv)
{
    h_ctx_t *h = hv;
    uint32_t s;
    for (s = 0; s < h->n_slots; s++) {
        block_thrashing_info_t *bt = h->vb[s].thrashing;
        if (bt) { free(bt->pages); free(bt); }
    }
    free(h->vb);
    free(h);
}
