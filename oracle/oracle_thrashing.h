/* oracle_thrashing.h -- TEST INFRASTRUCTURE ONLY: the harness layouts of the thrashing restatement (oracle_thrashing.c),
 * shared with oracle_service.c, which asks it for hints inside the fault loop. */
#ifndef ORACLE_THRASHING_H
#define ORACLE_THRASHING_H
#include <stdint.h>

typedef struct {
    uint64_t time_stamp;
    uint32_t slot;
    uint16_t first_page, num_pages;
    uint16_t processor;
    uint8_t kind;                 /* 0 = migration, 1 = revocation */
    uint8_t flags;                /* 1 = a migration that clears the tracking state (:1272-1285) */
    uint32_t reserved;
} thrash_event_t;                 /* 24 bytes, == uvmtier_perf_event_t */

typedef struct {
    uint64_t last_time_stamp;
    uint64_t last_thrashing_time_stamp;
    uint16_t last_processor;
    uint16_t num_thrashing_pages;
    uint32_t tracked;             /* block_thrashing->pages != NULL */
    uint16_t pinned_count;        /* pinned_pages.count */
    uint8_t reset_count;          /* thrashing_reset_count */
    uint8_t pad;
    uint32_t throttling_count;    /* stats */
} thrash_block_t;                 /* 32 bytes, == uvmtier_thrashing_block_t */

/* The hint half's per-page state (all zero = the reference's freshly allocated page: ids are stored + 1):
 *   page_throttle   throttling_end_time_stamp in 64 ns units
 *   page_hint       bits 0-7 throttling_count, 8-15 pinned_residency_id + 1, 16-23 do_not_throttle_processor_id + 1
 *   page_throttled  throttled_processors
 *   page_deadline   the pinned_page_t's deadline (pin_ns > 0), 0 = not on the list
 *   pinned_pages    pinned_pages.mask, 16 words per va_block */
typedef struct {
    thrash_block_t *blocks;
    uint64_t *page_word;
    uint32_t *page_procs;
    uint32_t *thrashing_pages;
    uint64_t *page_throttle;      /* NULL: detector only */
    uint32_t *page_hint;
    uint32_t *page_throttled;
    uint64_t *page_deadline;
    uint32_t *pinned_pages;
    uint32_t *counters;           /* [0] detections [1] resets [2] dropped [3] throttles [4] pins local [5] pins remote */
} thrash_tables_t;

typedef struct {
    uint64_t lapse_ns, nap_ns, epoch_ns, pin_ns;
    uint32_t threshold, pin_threshold, max_resets, has_memory;
    uint32_t accessible_from[32], can_access[32], has_fast_link[32], has_native_atomics[32];
} thrash_cfg_t;                   /* == uvmtier_thrashing_hint_config_t */
typedef struct {
    uint64_t time_stamp; uint32_t slot; uint16_t page; uint8_t requester, pad; uint32_t resident_processors;
    uint16_t preferred_location, pad2;
} thrash_query_t;                 /* 24 bytes, == uvmtier_thrashing_query_t */
typedef struct { uint64_t throttle_end; uint32_t pin_processors; uint8_t type, pin_residency; uint16_t pad; } thrash_hint_t;   /* 16 bytes */
#define ID_NONE 0xffffu


void oracle_thrashing_get_hints(uint32_t n, const thrash_query_t *q, const thrash_cfg_t *cfg, const thrash_tables_t *th, thrash_hint_t *out);
#endif
