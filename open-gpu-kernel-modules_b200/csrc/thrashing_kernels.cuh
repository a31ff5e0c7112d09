// thrashing_kernels.cuh -- the detection half of the thrashing heuristic (UVM/uvm_perf_thrashing.c:1230-1380) on flat
// per-page / per-va_block tables.  The reference runs its event callback under the va_block lock, one event at a
// time; va_blocks are independent, so here a warp takes all events of one va_block (they arrive grouped by va_block,
// in time order inside it) and the lanes take the pages of an event's region, 32 at a time.  Bound: HBM -- 12 B read
// and written per page an event covers, 24 B per event, 96 B per va_block touched.
#pragma once
#include "service_kernels.cuh"   // kMaxSlots
#include "thrashing_hint.cuh"

namespace uvmtier_k {

constexpr int kThrashThreads = 256;
constexpr uint32_t kThrashFullWarp = 0xffffffffu;
constexpr uint64_t kThrashTsMask = (1ull << 58) - 1;       // PAGE_THRASHING_LAST_TIME_STAMP_BITS (:38)
constexpr uint64_t kThrashMig = 1ull << 58, kThrashRev = 1ull << 59, kThrashPinned = 1ull << 63;
constexpr int kThrashNumShift = 60;                        // 3-bit counter (:39,58)

__device__ __forceinline__ bool thrash_event_ok(const uvmtier_perf_event_t &e, uint32_t n_slots)
{
    return e.slot < n_slots && (uint32_t)e.first_page + e.num_pages <= 512u && e.processor < 32u && e.kind <= 1u;
}

// scratch: [0] number of segments, [4 ..] first event of every segment, then one word per slot (the lowest head index
// of the slot's groups; 0xffffffff = none, set by the host before the call).
// counters: [0] detections, [1] page resets, [2] events dropped (malformed, or their va_block showed up in a second,
// separate group of the same call: the group that starts first is kept, whatever order the threads run in)
UVMTIER_KERNEL void __launch_bounds__(kThrashThreads)
k_thrash_heads_min(uint32_t n_events, const uvmtier_perf_event_t *__restrict__ ev, uint32_t n_slots, uint32_t *__restrict__ scratch)
{
    uint32_t *slot_head = scratch + 4 + n_slots;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_events; i += gridDim.x * blockDim.x) {
        const uint32_t slot = ev[i].slot;
        if ((i && ev[i - 1].slot == slot) || slot >= n_slots)
            continue;
        atomicMin(&slot_head[slot], i);
    }
}

UVMTIER_KERNEL void __launch_bounds__(kThrashThreads)
k_thrash_heads(uint32_t n_events, const uvmtier_perf_event_t *__restrict__ ev, uint32_t n_slots, uint32_t *__restrict__ scratch,
               uint32_t *__restrict__ counters)
{
    uint32_t *seg = scratch + 4;
    const uint32_t *slot_head = scratch + 4 + n_slots;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_events; i += gridDim.x * blockDim.x) {
        const uint32_t slot = ev[i].slot;
        if (i && ev[i - 1].slot == slot)
            continue;
        // no such va_block, or a later group of a va_block that already has one in this call: dropped, and reported
        if (slot >= n_slots || slot_head[slot] != i) {
            uint32_t n = 1;
            while (i + n < n_events && ev[i + n].slot == slot) n++;
            atomicAdd(&counters[2], n);
            continue;
        }
        seg[atomicAdd(&scratch[0], 1u)] = i;
    }
}

// th.page_throttle != NULL: the hint half is on (thrashing_hint.cuh) -- a reset also ends the page's throttling and pin, and
// a fault-driven migration whose whole region is pinned is not an event.
UVMTIER_KERNEL void __launch_bounds__(kThrashThreads)
k_thrash_events(uint32_t n_events, const uvmtier_perf_event_t *__restrict__ ev, uint32_t n_slots, unsigned long long lapse_ns,
                uint32_t threshold, uvmtier_thrashing_block_t *__restrict__ blocks, unsigned long long *__restrict__ page_word,
                uint32_t *__restrict__ page_procs, uint32_t *__restrict__ masks, const uint32_t *__restrict__ scratch,
                uint32_t *__restrict__ counters, thrash_hint_tables th)
{
    const bool hints_on = th.page_throttle != NULL;
    const uint32_t *seg = scratch + 4;
    const uint32_t n_seg = scratch[0];
    const int lane = threadIdx.x & 31;
    const uint32_t warps = gridDim.x * (kThrashThreads / 32);
    uint32_t detections = 0, resets = 0, dropped = 0;
    for (uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; s < n_seg; s += warps) {
        uint32_t i = seg[s];
        const uint32_t slot = ev[i].slot;
        uvmtier_thrashing_block_t blk = blocks[slot];
        uint32_t mask = lane < 16 ? masks[(size_t)slot * 16 + lane] : 0u;
        uint32_t pinned_mask = (hints_on && lane < 16) ? th.pinned_pages[(size_t)slot * 16 + lane] : 0u;
        unsigned long long *words = page_word + (size_t)slot * 512;
        uint32_t *procs = page_procs + (size_t)slot * 512;
        uvmtier_perf_event_t e_next = ev[i];
        for (; i < n_events; i++) {
            const uvmtier_perf_event_t e = e_next;
            if (e.slot != slot)
                break;
            if (i + 1 < n_events)
                e_next = ev[i + 1];                        // in flight while this event's pages are worked on
            if (!thrash_event_ok(e, n_slots)) {
                dropped += lane == 0;
                continue;
            }
            const uint32_t first = e.first_page, outer = first + e.num_pages;
            const unsigned long long ts = e.time_stamp;
            if (e.kind == UVMTIER_PERF_EVENT_MIGRATION && (e.flags & UVMTIER_PERF_EVENT_RESETS)) {
                // a migration the heuristic must not second-guess (:1272-1285): thrashing pages of the region forget
                // their history (thrashing_reset_pages_in_region :1077-1093); the va_block's stamp is left alone
                if (blk.tracked) {
                    for (uint32_t base = first & ~31u; base < outer; base += 32) {
                        const uint32_t p = base + lane;
                        const uint32_t mw = __shfl_sync(kThrashFullWarp, mask, base >> 5);
                        const bool hit = p >= first && p < outer && ((mw >> lane) & 1u);
                        if (hit) {
                            if (hints_on) {                // :1057-1061: the throttling and the pin end with the history
                                hint_reset_page_extras(th, (size_t)slot * 512 + p);
                                words[p] = 0;
                            }
                            else
                                words[p] &= kThrashPinned; // thrashing_reset_page :1063-1067
                            procs[p] = 0;
                            resets++;
                        }
                        const uint32_t b = __ballot_sync(kThrashFullWarp, hit);
                        if (lane == (int)(base >> 5)) { mask &= ~b; pinned_mask &= ~b; }
                        blk.num_thrashing_pages = (uint16_t)(blk.num_thrashing_pages - __popc(b));
                    }
                }
                continue;
            }
            if (hints_on && blk.tracked && e.kind == UVMTIER_PERF_EVENT_MIGRATION && !(e.flags & UVMTIER_PERF_EVENT_PREFETCH)) {
                // is_migration_pinned_pages_update :1198-1226: the pin being carried out, not an event
                bool full = true;
                for (uint32_t base = first & ~31u; base < outer; base += 32) {
                    const uint32_t pw = __shfl_sync(kThrashFullWarp, pinned_mask, base >> 5);
                    const uint32_t lo = first > base ? first - base : 0u, hi = outer - base < 32u ? outer - base : 32u;
                    const uint32_t want = (hi >= 32u ? 0xffffffffu : ((1u << hi) - 1u)) & ~((1u << lo) - 1u);
                    full = full && (pw & want) == want;
                }
                if (full)
                    continue;
            }
            bool skip = false;
            if (!blk.tracked) {
                // per-page tracking starts once two processors have touched the va_block inside the window (:1316-1337)
                if (blk.last_time_stamp == 0 || blk.last_processor == e.processor || ts - blk.last_time_stamp > lapse_ns)
                    skip = true;
                else
                    blk.tracked = 1;
            }
            if (!skip) {
                bool thrashing_now = false;
                for (uint32_t base = first & ~31u; base < outer; base += 32) {     // :1342-1375
                    const uint32_t p = base + lane;
                    const bool in = p >= first && p < outer;
                    const uint32_t mw = __shfl_sync(kThrashFullWarp, mask, base >> 5);
                    const bool in_mask = (mw >> lane) & 1u;
                    bool set = false, clr = false;
                    if (in) {
                        unsigned long long w = words[p];
                        const unsigned long long last = (w & kThrashTsMask) << 6;
                        uint32_t num = (uint32_t)(w >> kThrashNumShift) & 7u, pr = procs[p] | (1u << e.processor);
                        w = (w & ~kThrashTsMask) | (ts >> 6);
                        if (last != 0) {
                            if (ts - last <= lapse_ns) {
                                if (num < 7u) num++;
                                w = (w & ~(7ull << kThrashNumShift)) | ((unsigned long long)num << kThrashNumShift);
                                if (num == threshold) {    // thrashing_detected :1023-1040
                                    set = !in_mask;
                                    detections++;
                                }
                                thrashing_now |= num >= threshold;
                                w |= e.kind == UVMTIER_PERF_EVENT_MIGRATION ? kThrashMig : kThrashRev;
                            }
                            else if (num >= threshold && !(w & kThrashPinned)) {
                                if (hints_on)
                                    hint_throttling_reset(th, (size_t)slot * 512 + p);     // :1057 (not pinned: nothing to un-pin)
                                w &= kThrashPinned;        // thrashing_reset_page :1045-1073
                                pr = 0;
                                clr = in_mask;
                                resets++;
                            }
                        }
                        words[p] = w;
                        procs[p] = pr;
                    }
                    const uint32_t bs = __ballot_sync(kThrashFullWarp, set), bc = __ballot_sync(kThrashFullWarp, clr);
                    if (lane == (int)(base >> 5)) mask = (mask | bs) & ~bc;
                    blk.num_thrashing_pages = (uint16_t)(blk.num_thrashing_pages + __popc(bs) - __popc(bc));
                }
                if (__any_sync(kThrashFullWarp, thrashing_now))
                    blk.last_thrashing_time_stamp = ts;
            }
            blk.last_time_stamp = ts;                      // done: (:1377-1379)
            blk.last_processor = e.processor;
        }
        if (lane == 0) blocks[slot] = blk;
        if (lane < 16) masks[(size_t)slot * 16 + lane] = mask;
        if (hints_on && lane < 16) th.pinned_pages[(size_t)slot * 16 + lane] = pinned_mask;
    }
    detections = __reduce_add_sync(kThrashFullWarp, detections);
    resets = __reduce_add_sync(kThrashFullWarp, resets);
    if (lane == 0) {
        if (detections) atomicAdd(&counters[0], detections);
        if (resets) atomicAdd(&counters[1], resets);
        if (dropped) atomicAdd(&counters[2], dropped);
    }
}

} // namespace uvmtier_k
