/*
 * ref_thrashing_glue.c -- TEST INFRASTRUCTURE ONLY (oracle/_ref build -> libref_thrashing.so).
 *
 * Compiles the detection half of the reference's thrashing heuristic VERBATIM -- line-range extracts of
 * kernel-open/nvidia-uvm/uvm_perf_thrashing.c:
 *   :36-119     bit widths, page_thrashing_info_t, block_thrashing_info_t
 *   :376-384    page time stamp get / set
 *   :1023-1073  thrashing_detected, thrashing_reset_page (wrapped by a counter)
 *   :1077-1093  thrashing_reset_pages_in_region
 *   :1230-1380  thrashing_event_cb
 * and uvm_perf_utils.h:30-43 (UVM_PERF_SATURATING_ADD / _INC), against stand-ins for the va_block, the va_space's
 * parameters, the perf-event payload and the page / processor masks.  The hint side (pin, throttle) is stubbed: no
 * page is pinned or throttled, so the two calls thrashing_reset_page makes into it are no-ops.  NV_GETTIME() is the
 * event's own time stamp.
 *
 * What this pins: N2's detector -- the va_block gate, the per-page window / counter / processors update, detection
 * and reset, the thrashing_pages mask.
 */
#include "shim/kshim.h"
#include "ref_api.h"
#include <stdlib.h>

typedef struct { NvU32 val; } uvm_processor_id_t;
#define UVM_ID_MAX_PROCESSORS 32
typedef struct { DECLARE_BITMAP(bitmap, UVM_ID_MAX_PROCESSORS); } uvm_processor_mask_t;
typedef struct { DECLARE_BITMAP(bitmap, 512); } uvm_page_mask_t;
typedef NvU16 uvm_page_index_t;
typedef struct { uvm_page_index_t first, outer; } uvm_va_block_region_t;
static const uvm_processor_id_t UVM_ID_INVALID = { 0xffff };
static bool uvm_id_equal(uvm_processor_id_t a, uvm_processor_id_t b) { return a.val == b.val; }
static void uvm_processor_mask_set(uvm_processor_mask_t *m, uvm_processor_id_t id) { __set_bit(id.val, m->bitmap); }
static void uvm_processor_mask_zero(uvm_processor_mask_t *m) { bitmap_zero(m->bitmap, UVM_ID_MAX_PROCESSORS); }
static bool uvm_processor_mask_empty(const uvm_processor_mask_t *m) { return bitmap_empty(m->bitmap, UVM_ID_MAX_PROCESSORS); }
static bool uvm_page_mask_test(const uvm_page_mask_t *m, uvm_page_index_t p) { return test_bit(p, m->bitmap); }
static bool uvm_page_mask_test_and_set(uvm_page_mask_t *m, uvm_page_index_t p) { return __test_and_set_bit(p, m->bitmap); }
static bool uvm_page_mask_test_and_clear(uvm_page_mask_t *m, uvm_page_index_t p) { return __test_and_clear_bit(p, m->bitmap); }

typedef struct { int read_duplication; } uvm_va_policy_t;
typedef struct { uvm_va_policy_t policy; } uvm_va_range_managed_t;
typedef struct uvm_va_space_struct uvm_va_space_t;
typedef struct uvm_va_block_struct {
    NvU64 start, end;
    uvm_va_range_managed_t *managed_range;
    uvm_va_space_t *va_space;
    void *thrashing;                               /* perf_modules_data[UVM_PERF_MODULE_TYPE_THRASHING] */
    int lock;
} uvm_va_block_t;
typedef struct { struct { bool enable; unsigned threshold; NvU64 lapse_ns; } params; } va_space_thrashing_info_t;
struct uvm_va_space_struct { va_space_thrashing_info_t thrashing; };
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

static NvU64 g_now;
static NvU64 g_detections, g_resets;
static bool g_uvm_perf_thrashing_enable = true;
#define NV_GETTIME() g_now
#define uvm_kvmalloc_zero(n) calloc(1, n)
#define PROCESSOR_THRASHING_STATS_INC(proc, field) ((void)(proc), g_detections++)
#define uvm_tools_record_thrashing(vs, addr, size, procs) ((void)(vs), (void)(addr), (void)(procs))
static uvm_va_space_t *uvm_va_block_get_va_space(uvm_va_block_t *b) { return b->va_space; }
static NvU64 uvm_va_block_cpu_page_address(uvm_va_block_t *b, uvm_page_index_t p) { return b->start + (NvU64)p * PAGE_SIZE; }
static NvU64 uvm_va_block_size(uvm_va_block_t *b) { return b->end - b->start + 1; }
static bool uvm_va_block_is_hmm(uvm_va_block_t *b) { (void)b; return false; }
static va_space_thrashing_info_t *va_space_thrashing_info_get(uvm_va_space_t *vs) { return &vs->thrashing; }
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

#include "x_perf_saturating.h"     /* uvm_perf_utils.h:30-43 */
#include "x_thrashing_types.h"     /* uvm_perf_thrashing.c:36-119 */
#include "x_thrashing_stamps.h"    /* :376-384 */

static bool thrashing_state_checks(uvm_va_block_t *b, block_thrashing_info_t *bt, page_thrashing_info_t *pt, uvm_page_index_t p)
{ (void)b; (void)bt; (void)pt; (void)p; return true; }
static void thrashing_throttling_reset_page(uvm_va_block_t *b, block_thrashing_info_t *bt, page_thrashing_info_t *pt, uvm_page_index_t p)
{ (void)b; (void)bt; (void)pt; (void)p; }
static void thrashing_unpin_page(va_space_thrashing_info_t *vs, uvm_va_block_t *b, block_thrashing_info_t *bt, page_thrashing_info_t *pt,
                                 uvm_page_index_t p)
{ (void)vs; (void)b; (void)bt; (void)pt; (void)p; }
static block_thrashing_info_t *thrashing_info_get(uvm_va_block_t *b) { return b->thrashing; }
static block_thrashing_info_t *thrashing_info_get_create(uvm_va_block_t *b)
{
    if (!b->thrashing) {                           /* :645-665 */
        block_thrashing_info_t *bt = calloc(1, sizeof(*bt));
        bt->last_processor = UVM_ID_INVALID;
        b->thrashing = bt;
    }
    return b->thrashing;
}
static bool migrating_wrong_pages(uvm_va_block_t *b, NvU64 a, NvU64 n, uvm_processor_id_t p, uvm_make_resident_cause_t c)
{ (void)b; (void)a; (void)n; (void)p; (void)c; return false; }
static bool is_migration_pinned_pages_update(uvm_va_block_t *b, uvm_perf_event_data_t *d, NvU64 a, NvU64 n)
{ (void)b; (void)d; (void)a; (void)n; return false; }
static void thrashing_reset_pages_in_region(uvm_va_block_t *va_block, NvU64 address, NvU64 bytes);

#define thrashing_reset_page thrashing_reset_page_ref
#include "x_thrashing_detect.h"    /* :1023-1073 */
#undef thrashing_reset_page
static void thrashing_reset_page(va_space_thrashing_info_t *vs, uvm_va_block_t *b, block_thrashing_info_t *bt, uvm_page_index_t p)
{
    g_resets++;                                    /* counted for the harness, then the reference's function */
    thrashing_reset_page_ref(vs, b, bt, p);
}
#include "x_thrashing_region.h"    /* :1077-1093 */
void thrashing_event_cb(uvm_perf_event_t event_id, uvm_perf_event_data_t *event_data);
#include "x_thrashing_event.h"     /* :1230-1380 */

/* ---- harness: same event / state layout as oracle_thrashing.c ---- */
typedef struct { uint64_t time_stamp; uint32_t slot; uint16_t first_page, num_pages; uint16_t processor; uint8_t kind, flags; uint32_t reserved; } h_event_t;
typedef struct { uint64_t last_time_stamp, last_thrashing_time_stamp; uint16_t last_processor, num_thrashing_pages; uint32_t tracked; uint64_t reserved; } h_block_t;

/* Runs the events through thrashing_event_cb on fresh va_blocks (all-zero tracking state) and writes the state out. */
void ref_thrashing_note_events(uint32_t n_slots, uint32_t n_events, const h_event_t *ev, uint64_t lapse_ns, uint32_t threshold,
                               h_block_t *blocks, uint64_t *page_word, uint32_t *page_procs, uint32_t *thrashing_pages,
                               uint32_t *counters)
{
    uvm_va_space_t vs;
    uvm_va_range_managed_t range;
    uvm_va_block_t *vb = calloc(n_slots, sizeof(*vb));
    uvm_make_resident_context_t mrc;
    uint32_t i, s, p;
    vs.thrashing.params.enable = true;
    vs.thrashing.params.threshold = threshold;
    vs.thrashing.params.lapse_ns = lapse_ns;
    range.policy.read_duplication = UVM_READ_DUPLICATION_DISABLED;
    for (s = 0; s < n_slots; s++) {
        vb[s].start = (NvU64)s << 21; vb[s].end = vb[s].start + (2u << 20) - 1;
        vb[s].managed_range = &range; vb[s].va_space = &vs;
    }
    g_detections = g_resets = 0;
    for (i = 0; i < n_events; i++) {
        uvm_perf_event_data_t d;
        uvm_va_block_t *b = &vb[ev[i].slot];
        memset(&d, 0, sizeof(d));
        g_now = ev[i].time_stamp;
        if (ev[i].kind == 0) {
            d.migration.block = b; d.migration.address = b->start + (NvU64)ev[i].first_page * PAGE_SIZE;
            d.migration.bytes = (NvU64)ev[i].num_pages * PAGE_SIZE; d.migration.dst.val = ev[i].processor;
            d.migration.cause = (ev[i].flags & 1) ? UVM_MAKE_RESIDENT_CAUSE_API_MIGRATE : UVM_MAKE_RESIDENT_CAUSE_REPLAYABLE_FAULT;
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
    counters[0] = (uint32_t)g_detections;
    counters[1] = (uint32_t)g_resets;
    for (s = 0; s < n_slots; s++) {
        block_thrashing_info_t *bt = vb[s].thrashing;
        memset(&blocks[s], 0, sizeof(blocks[s]));
        if (!bt)
            continue;
        blocks[s].last_time_stamp = bt->last_time_stamp;
        blocks[s].last_thrashing_time_stamp = bt->last_thrashing_time_stamp;
        blocks[s].last_processor = (uint16_t)bt->last_processor.val;
        blocks[s].num_thrashing_pages = bt->num_thrashing_pages;
        blocks[s].tracked = bt->pages != NULL;
        for (p = 0; p < 16; p++)
            thrashing_pages[(size_t)s * 16 + p] = (uint32_t)(bt->thrashing_pages.bitmap[p / 2] >> (32 * (p & 1)));
        if (bt->pages)
            for (p = 0; p < 512; p++) {
                page_thrashing_info_t *pt = &bt->pages[p];
                page_word[(size_t)s * 512 + p] = (uint64_t)pt->last_time_stamp | ((uint64_t)pt->has_migration_events << 58) |
                                                 ((uint64_t)pt->has_revocation_events << 59) |
                                                 ((uint64_t)pt->num_thrashing_events << 60) | ((uint64_t)pt->pinned << 63);
                page_procs[(size_t)s * 512 + p] = (uint32_t)pt->processors.bitmap[0];
            }
        free(bt->pages);
        free(bt);
    }
    free(vb);
}
