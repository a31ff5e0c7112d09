/* oracle_thrashing.c -- TEST INFRASTRUCTURE ONLY (CPU restatement, never on the product path).
 *
 * The detection half of the thrashing heuristic: what the reference's perf-event callback does to its per-page
 * and per-va_block tracking state for one migration / revocation event
 * (kernel-open/nvidia-uvm/uvm_perf_thrashing.c):
 *   thrashing_event_cb                :1230-1380  (block gate :1320-1337, per-page update :1342-1375)
 *   thrashing_detected                :1023-1040
 *   thrashing_reset_page              :1045-1073
 *   thrashing_reset_pages_in_region   :1077-1093
 *   page time stamps (64 ns units)    :376-384, counter width :38-39,58
 * Not covered: the hint (pin / throttle, :1455-1800) -- no page is ever pinned or throttled here, so the two
 * calls thrashing_reset_page makes into that state (:1057-1061) have nothing to undo.
 *
 * Pinning: tests/test_oracle_thrashing.py runs the same event streams through oracle/_ref/libref_thrashing.so
 * (the reference's own lines, compiled against shims by build_ref.sh) and compares every field.
 */
#include <stdint.h>
#include <string.h>

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
    uint64_t reserved;
} thrash_block_t;                 /* 32 bytes, == uvmtier_thrashing_block_t */

#define TS_MASK ((1ull << 58) - 1)
#define W_MIG (1ull << 58)
#define W_REV (1ull << 59)
#define W_NUM_SHIFT 60
#define W_PINNED (1ull << 63)

static void reset_page(thrash_block_t *blk, uint64_t *word, uint32_t *procs, uint32_t *mask, uint32_t page, uint32_t *counters)
{
    /* :1063-1070 (pinned is never set here, so it stays as it is: zero) */
    *word &= W_PINNED;
    *procs = 0;
    if (mask[page >> 5] & (1u << (page & 31))) {
        mask[page >> 5] &= ~(1u << (page & 31));
        blk->num_thrashing_pages--;
    }
    counters[1]++;
}

void oracle_thrashing_note_events(uint32_t n_events, const thrash_event_t *ev, uint64_t lapse_ns, uint32_t threshold,
                                  thrash_block_t *blocks, uint64_t *page_word, uint32_t *page_procs,
                                  uint32_t *thrashing_pages, uint32_t *counters)
{
    for (uint32_t i = 0; i < n_events; i++) {
        const thrash_event_t *e = &ev[i];
        thrash_block_t *blk = &blocks[e->slot];
        uint64_t *words = page_word + (size_t)e->slot * 512;
        uint32_t *procs = page_procs + (size_t)e->slot * 512;
        uint32_t *mask = thrashing_pages + (size_t)e->slot * 16;
        const uint32_t first = e->first_page, outer = (uint32_t)e->first_page + e->num_pages;
        const uint64_t ts = e->time_stamp;

        if (e->kind == 0 && (e->flags & 1)) {                            /* :1272-1285 -> :1077-1093 */
            if (blk->tracked)
                for (uint32_t p = first; p < outer; p++)
                    if (mask[p >> 5] & (1u << (p & 31)))
                        reset_page(blk, &words[p], &procs[p], mask, p, counters);
            continue;                                                    /* returns before `done:` */
        }
        if (!blk->tracked) {                                             /* :1316-1337 */
            if (blk->last_time_stamp == 0 || blk->last_processor == e->processor || ts - blk->last_time_stamp > lapse_ns)
                goto done;
            blk->tracked = 1;
        }
        for (uint32_t p = first; p < outer; p++) {                       /* :1342-1375 */
            uint64_t w = words[p];
            const uint64_t last = (w & TS_MASK) << 6;                    /* :376-379 */
            uint32_t num = (uint32_t)(w >> W_NUM_SHIFT) & 7u;
            procs[p] |= 1u << e->processor;
            w = (w & ~TS_MASK) | (ts >> 6);                              /* :381-384 */
            if (last != 0) {
                if (ts - last <= lapse_ns) {
                    if (num < 7) num++;                                  /* 3-bit saturating counter */
                    w = (w & ~(7ull << W_NUM_SHIFT)) | ((uint64_t)num << W_NUM_SHIFT);
                    if (num == threshold) {                              /* thrashing_detected :1023-1040 */
                        if (!(mask[p >> 5] & (1u << (p & 31)))) {
                            mask[p >> 5] |= 1u << (p & 31);
                            blk->num_thrashing_pages++;
                        }
                        counters[0]++;
                    }
                    if (num >= threshold)
                        blk->last_thrashing_time_stamp = ts;
                    w |= e->kind == 0 ? W_MIG : W_REV;
                }
                else if (num >= threshold && !(w & W_PINNED)) {
                    words[p] = w;
                    reset_page(blk, &words[p], &procs[p], mask, p, counters);
                    continue;
                }
            }
            words[p] = w;
        }
done:
        blk->last_time_stamp = ts;
        blk->last_processor = e->processor;
    }
}
