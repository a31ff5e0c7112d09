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
 * and the hint half (oracle_thrashing_get_hints, oracle_thrashing_unpin_expired below):
 *   uvm_perf_thrashing_get_hint       :1615-1780  (epoch reset :1644-1674, throttled requester :1684-1694, window :1702-1705)
 *   get_hint_for_migration_thrashing  :1455-1595, thrashing_processors_* :1382-1453
 *   thrashing_throttle_update / _processor / _end_processor / thrashing_throttling_reset_page  :772-881
 *   thrashing_pin_page / thrashing_unpin_page :902-1021, the expiry worker thrashing_unpin_pages :1822-1906
 *   is_migration_pinned_pages_update  :1198-1226 (a fault-driven migration onto pinned pages is not an event)
 *   uvm_processor_mask_find_closest_id  uvm_va_space.c:1879-1934
 * With hint tables (th->page_throttle != NULL) thrashing_reset_page also ends the throttling and the pin (:1057-1061).
 *
 * Pinning: tests/test_oracle_thrashing.py runs the same event streams through oracle/_ref/libref_thrashing.so, and
 * interleaved events / hint queries / expiries through oracle/_ref/libref_thrashing_hint.so (the reference's own
 * lines, compiled against shims by build_ref.sh), and compares every field.
 */
#include <stdint.h>
#include <string.h>

#include "oracle_thrashing.h"

#define TS_MASK ((1ull << 58) - 1)
#define W_MIG (1ull << 58)
#define W_REV (1ull << 59)
#define W_NUM_SHIFT 60
#define W_PINNED (1ull << 63)

/* thrashing_throttling_reset_page :865-881 -> thrashing_throttle_end_processor :841-861 for every throttled processor */
static void throttling_reset(const thrash_tables_t *th, size_t at)
{
    if (th->page_throttled[at]) {                                        /* the end time stamp goes with the last throttled processor */
        th->page_throttled[at] = 0;
        th->page_throttle[at] = 0;
    }
}

/* thrashing_unpin_page :974-1021 */
static void unpin_page(const thrash_tables_t *th, thrash_block_t *blk, uint32_t slot, uint32_t page)
{
    const size_t at = (size_t)slot * 512 + page;
    th->page_deadline[at] = 0;
    th->page_hint[at] &= ~(0xffu << 8);                                  /* pinned_residency_id = INVALID */
    th->page_word[at] &= ~W_PINNED;
    th->pinned_pages[(size_t)slot * 16 + (page >> 5)] &= ~(1u << (page & 31));
    (void)blk;                                                           /* pinned_pages.count is never decremented by the reference */
}

static const thrash_tables_t *g_th;                                      /* hint tables of the call in progress (NULL: none) */
static uint32_t g_slot;

static void reset_page(thrash_block_t *blk, uint64_t *word, uint32_t *procs, uint32_t *mask, uint32_t page, uint32_t *counters)
{
    if (g_th && g_th->page_throttle) {                                   /* :1057-1061 */
        const size_t at = (size_t)g_slot * 512 + page;
        throttling_reset(g_th, at);
        if (*word & W_PINNED)
            unpin_page(g_th, blk, g_slot, page);
    }
    /* :1063-1070 */
    *word &= W_PINNED;
    *procs = 0;
    if (mask[page >> 5] & (1u << (page & 31))) {
        mask[page >> 5] &= ~(1u << (page & 31));
        blk->num_thrashing_pages--;
    }
    counters[1]++;
}

static void note_events(uint32_t n_events, const thrash_event_t *ev, uint64_t lapse_ns, uint32_t threshold,
                        thrash_block_t *blocks, uint64_t *page_word, uint32_t *page_procs,
                        uint32_t *thrashing_pages, uint32_t *counters, const thrash_tables_t *th)
{
    g_th = th;
    for (uint32_t i = 0; i < n_events; i++) {
        const thrash_event_t *e = &ev[i];
        thrash_block_t *blk = &blocks[e->slot];
        g_slot = e->slot;
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
        if (e->kind == 0 && !(e->flags & 2) && th && th->page_throttle && blk->tracked) {
            /* is_migration_pinned_pages_update :1198-1226: a fault / access-counter migration whose whole region is
             * pinned is the pin being carried out, not an event (flag 2 = cause PREFETCH, which is not in that list) */
            const uint32_t *pinned = th->pinned_pages + (size_t)e->slot * 16;
            int full = 1;
            for (uint32_t p = first; p < outer; p++)
                if (!(pinned[p >> 5] & (1u << (p & 31)))) { full = 0; break; }
            if (full)
                continue;
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

void oracle_thrashing_note_events(uint32_t n_events, const thrash_event_t *ev, uint64_t lapse_ns, uint32_t threshold,
                                  thrash_block_t *blocks, uint64_t *page_word, uint32_t *page_procs,
                                  uint32_t *thrashing_pages, uint32_t *counters)
{
    note_events(n_events, ev, lapse_ns, threshold, blocks, page_word, page_procs, thrashing_pages, counters, NULL);
}

/* The same with the hint half's tables next to the detector's. */
void oracle_thrashing_note_events_tables(uint32_t n_events, const thrash_event_t *ev, const thrash_cfg_t *cfg, const thrash_tables_t *th)
{
    note_events(n_events, ev, cfg->lapse_ns, cfg->threshold, th->blocks, th->page_word, th->page_procs, th->thrashing_pages,
                th->counters, th);
}

/* ---- the hint ---- */
static uint32_t smc_range(uint32_t gpu_id) { return 0xffu << (1u + 8u * ((gpu_id - 1u) / 8u)); }   /* ids sharing the parent GPU */
static int first_gpu(uint32_t mask) { mask &= ~1u; return mask ? __builtin_ctz(mask) : -1; }

/* uvm_processor_mask_find_closest_id, uvm_va_space.c:1879-1934 */
static uint32_t closest_id(const thrash_cfg_t *cfg, uint32_t candidates, uint32_t src)
{
    uint32_t m;
    int g;
    if (candidates & (1u << src))
        return src;
    if (src != 0 && (m = candidates & smc_range(src)) != 0)
        return (uint32_t)first_gpu(m);
    if ((m = candidates & cfg->has_fast_link[src]) != 0 || (m = candidates & cfg->can_access[src]) != 0) {
        g = first_gpu(m);
        return g >= 0 ? (uint32_t)g : 0u;
    }
    return candidates ? (uint32_t)__builtin_ctz(candidates) : ID_NONE;
}

/* get_hint_for_migration_thrashing :1455-1595; returns the type, *residency for a pin */
static uint32_t hint_for_migration(const thrash_cfg_t *cfg, uint32_t procs, uint32_t pinned, uint32_t pinned_residency, uint32_t dnt,
                                   uint32_t throttling_count, uint32_t requester, uint32_t preferred, uint32_t resident,
                                   uint32_t *residency)
{
    uint32_t type = 0, res = ID_NONE;
    const uint32_t closest = resident ? closest_id(cfg, resident, requester) : ID_NONE;
    const int can_access_pref = preferred != ID_NONE && (procs & ~cfg->accessible_from[preferred]) == 0;        /* :1382-1391 */
    int fast = 0;
    if (closest != ID_NONE) {                                                                                    /* :1393-1422 */
        uint32_t fast_to = cfg->has_fast_link[closest] & cfg->has_native_atomics[closest];
        fast_to |= closest == 0 ? 1u : smc_range(closest);
        fast = (procs & ~fast_to) == 0;
    }
    if (can_access_pref) {
        type = 1; res = preferred;
    }
    else if (!(preferred != ID_NONE && (procs & (1u << preferred))) && fast) {
        if (closest == 0) {
            if (requester != 0) { type = 1; res = requester; }
        }
        else { type = 1; res = closest; }
    }
    else if (requester == preferred) {
        if (pinned) {
            if (preferred == pinned_residency || preferred == dnt) { type = 1; res = preferred; }
            else type = 2;
        }
        else if (preferred != dnt) type = 2;
        else if (throttling_count >= cfg->pin_threshold) { type = 1; res = preferred; }
    }
    else if (pinned) {
        if (requester == dnt) {
            type = 1;
            if ((procs & ~cfg->accessible_from[requester]) == 0)
                res = requester;
            else {
                uint32_t common = 0xffffffffu, rest = procs;                                                       /* :1424-1444 */
                while (rest) { common &= cfg->can_access[__builtin_ctz(rest)]; rest &= rest - 1; }
                res = common ? closest_id(cfg, common, requester) : requester;
            }
        }
        else if (cfg->accessible_from[pinned_residency] & (1u << requester)) { type = 1; res = pinned_residency; }
        else type = 2;
    }
    else if (requester != dnt) type = 2;
    else if (throttling_count >= cfg->pin_threshold) { type = 1; res = requester; }
    if (type == 1 && !(cfg->has_memory & (1u << res)))
        res = 0;                                                                                                 /* UVM_ID_CPU */
    *residency = res;
    return type;
}

/* uvm_perf_thrashing_get_hint for every query, in order. */
void oracle_thrashing_get_hints(uint32_t n, const thrash_query_t *q, const thrash_cfg_t *cfg, const thrash_tables_t *th, thrash_hint_t *out)
{
    for (uint32_t i = 0; i < n; i++) {
        const uint32_t slot = q[i].slot, page = q[i].page, req = q[i].requester;
        const uint64_t ts = q[i].time_stamp;
        thrash_block_t *blk = &th->blocks[slot];
        const size_t at = (size_t)slot * 512 + page;
        uint32_t type = 0, residency = ID_NONE;
        int have_page = 0;
        memset(&out[i], 0, sizeof(out[i]));
        out[i].pin_residency = 0xff;
        if (!blk->tracked)
            continue;
        if (blk->last_thrashing_time_stamp != 0 && ts - blk->last_thrashing_time_stamp > cfg->epoch_ns && blk->pinned_count == 0 &&
            blk->reset_count < cfg->max_resets) {                                                                /* :1644-1674 */
            blk->reset_count++;
            for (uint32_t p = 0; p < 512; p++) {
                const size_t a = (size_t)slot * 512 + p;
                th->page_word[a] = 0; th->page_procs[a] = 0; th->page_throttle[a] = 0; th->page_hint[a] = 0;
                th->page_throttled[a] = 0; th->page_deadline[a] = 0;
            }
            blk->tracked = 0; blk->num_thrashing_pages = 0; blk->last_processor = (uint16_t)ID_NONE;
            blk->last_time_stamp = 0; blk->last_thrashing_time_stamp = 0;
            memset(th->thrashing_pages + (size_t)slot * 16, 0, 64);
            continue;                                                                                            /* hint NONE, no page */
        }
        have_page = 1;
        {
            uint64_t w = th->page_word[at];
            uint32_t h = th->page_hint[at];
            const uint32_t num = (uint32_t)(w >> W_NUM_SHIFT) & 7u;
            uint32_t dnt = ((h >> 16) & 0xffu) ? ((h >> 16) & 0xffu) - 1u : ID_NONE;
            const uint32_t pinned_res = ((h >> 8) & 0xffu) ? ((h >> 8) & 0xffu) - 1u : ID_NONE;
            const int pinned = (w & W_PINNED) != 0;
            if (num < cfg->threshold)
                goto done;
            if (th->page_throttled[at] & (1u << req)) {                                                          /* :1684-1694 */
                const uint64_t end = th->page_throttle[at] << 6;
                if (ts < end && req != dnt) { type = 2; goto done; }
                th->page_throttled[at] &= ~(1u << req);                                                          /* :841-861 */
                if (!th->page_throttled[at]) th->page_throttle[at] = 0;
            }
            if (ts - ((w & TS_MASK) << 6) > cfg->lapse_ns && !pinned)                                            /* :1702-1705 */
                goto done;
            th->page_procs[at] |= 1u << req;
            if (ts > (th->page_throttle[at] << 6)) {                                                             /* thrashing_throttle_update :772-795 */
                th->page_throttle[at] = ((ts + cfg->nap_ns) >> 6) & TS_MASK;
                dnt = dnt == req ? ID_NONE : req;
            }
            else if (dnt == ID_NONE)
                dnt = req;
            h = (h & ~(0xffu << 16)) | ((dnt == ID_NONE ? 0u : dnt + 1u) << 16);
            th->page_hint[at] = h;
            if (pinned && (w & W_REV) && req != dnt)
                type = 2;
            else
                type = hint_for_migration(cfg, th->page_procs[at], (uint32_t)pinned, pinned_res, dnt, h & 0xffu, req,
                                          q[i].preferred_location == 0xffff ? ID_NONE : q[i].preferred_location,
                                          q[i].resident_processors, &residency);
        }
done:
        if (type == 1) {                                                                                         /* thrashing_pin_page :902-967 */
            uint64_t w = th->page_word[at];
            if (!(w & W_PINNED) || !(q[i].resident_processors & (1u << residency)))
                throttling_reset(th, at);
            if (!(w & W_PINNED)) {
                if (cfg->pin_ns > 0)
                    th->page_deadline[at] = ts + cfg->pin_ns;
                th->page_hint[at] &= ~0xffu;                                                                     /* throttling_count = 0 */
                th->page_word[at] = w | W_PINNED;
                if (blk->pinned_count != 0xffff) blk->pinned_count++;
                th->pinned_pages[(size_t)slot * 16 + (page >> 5)] |= 1u << (page & 31);
            }
            th->page_hint[at] = (th->page_hint[at] & ~(0xffu << 8)) | ((residency + 1u) << 8);
            th->counters[residency == req ? 4 : 5]++;
            out[i].pin_residency = (uint8_t)residency;
            out[i].pin_processors = th->page_procs[at];
        }
        if (type == 2 && have_page) {                                                                            /* thrashing_throttle_processor :805-832 */
            if (!(th->page_throttled[at] & (1u << req))) {
                th->page_throttled[at] |= 1u << req;
                if (!(th->page_word[at] & W_PINNED) && (th->page_hint[at] & 0xffu) != 0xffu)
                    th->page_hint[at]++;
                if (blk->throttling_count != 0xffffffffu) blk->throttling_count++;
            }
            th->counters[3]++;
            out[i].throttle_end = th->page_throttle[at] << 6;
        }
        out[i].type = (uint8_t)type;
    }
}

/* thrashing_unpin_pages :1822-1906 at time `now`: every pin whose deadline has passed (20 us of timer slack) ends, and
 * the page's tracking state is cleared (thrashing_reset_page).  Deadlines grow with time, so "the head of the list while
 * it is due" is "every page that is due". */
void oracle_thrashing_unpin_expired(uint32_t n_slots, uint64_t now, const thrash_tables_t *th)
{
    g_th = th;
    for (uint32_t s = 0; s < n_slots; s++) {
        thrash_block_t *blk = &th->blocks[s];
        if (!blk->tracked)
            continue;
        g_slot = s;
        for (uint32_t p = 0; p < 512; p++) {
            const size_t at = (size_t)s * 512 + p;
            if ((th->page_word[at] & W_PINNED) && th->page_deadline[at] != 0 && th->page_deadline[at] <= now + 20000ull)
                reset_page(blk, &th->page_word[at], &th->page_procs[at], th->thrashing_pages + (size_t)s * 16, p, th->counters);
        }
    }
}
