// thrashing_hint_dev.cuh -- the hint half of the thrashing heuristic (UVM/uvm_perf_thrashing.c:1382-1780 and the throttle / pin
// helpers :772-1021) on flat per-page tables, next to the detector's (thrashing_kernels.cuh).
//
// uvm_perf_thrashing_get_hint is a decision per faulting page that reads AND advances that page's throttling / pinning
// state, so hints for one page are a chain; pages of different va_blocks are independent (a va_block's epoch reset wipes
// all of its pages, so the va_block is the unit).  Queries arrive grouped by va_block, in time order inside a group; one
// thread walks a group.  This is the cold side of the fault path -- only pages the detector has flagged get this far -- so
// the code favours being the reference's logic line for line over width.
//
// Per-page state, all zero = the reference's freshly allocated page (ids are stored + 1):
//   page_throttle   throttling_end_time_stamp in 64 ns units (:66, :386-394)
//   page_hint       bits 0-7 throttling_count (:71), 8-15 pinned_residency_id + 1 (:81), 16-23 do_not_throttle_processor_id + 1 (:84)
//   page_throttled  throttled_processors (:78)
//   page_deadline   deadline of the page's pinned_page_t (:126-140), 0 = none
//   pinned_pages    block_thrashing_info_t.pinned_pages.mask, 16 words per va_block
#pragma once

#include <stdint.h>

#include "uvmtier.h"
#include "devio.cuh"

namespace uvmtier_k {

constexpr uint64_t kHintTsMask = (1ull << 58) - 1;
constexpr uint64_t kHintRev = 1ull << 59, kHintPinned = 1ull << 63;
constexpr int kHintNumShift = 60;
constexpr uint32_t kHintNoId = 0xffffu;

struct thrash_hint_tables {
    uvmtier_thrashing_block_t *blocks;
    unsigned long long *page_word;
    uint32_t *page_procs;
    uint32_t *thrashing_pages;
    unsigned long long *page_throttle;         // NULL: the hint half is off, the detector runs alone
    uint32_t *page_hint;
    uint32_t *page_throttled;
    unsigned long long *page_deadline;
    uint32_t *pinned_pages;
};

// thrashing_throttling_reset_page :865-881: every throttled processor stops being throttled; the end time stamp goes
// with the last of them (:852-853), and stays when nobody was throttled.
__device__ __forceinline__ void hint_throttling_reset(const thrash_hint_tables &th, size_t at)
{
    if (th.page_throttled[at]) {
        th.page_throttled[at] = 0;
        th.page_throttle[at] = 0;
    }
}

// What thrashing_reset_page (:1045-1073) undoes besides the detector's fields: the throttling and the pin (:1057-1061).
// The caller clears the detector's word (the pinned bit with it) and the page's bit in pinned_pages.
__device__ __forceinline__ void hint_reset_page_extras(const thrash_hint_tables &th, size_t at)
{
    hint_throttling_reset(th, at);
    th.page_deadline[at] = 0;                                             // thrashing_unpin_page :974-1021
    th.page_hint[at] &= ~(0xffu << 8);
}

__device__ __forceinline__ uint32_t hint_smc_range(uint32_t gpu_id) { return 0xffu << (1u + 8u * ((gpu_id - 1u) / 8u)); }
__device__ __forceinline__ int hint_first_gpu(uint32_t mask) { mask &= ~1u; return mask ? __ffs((int)mask) - 1 : -1; }

// uvm_processor_mask_find_closest_id, UVM/uvm_va_space.c:1879-1934: the processor itself, an SMC sibling, a fast-link
// peer (GPUs before the CPU), a peer it can reach at all, whoever comes first.
__device__ __forceinline__ uint32_t hint_closest_id(const uvmtier_thrashing_hint_config_t &cfg, uint32_t candidates, uint32_t src)
{
    uint32_t m;
    if (candidates & (1u << src))
        return src;
    if (src != 0 && (m = candidates & hint_smc_range(src)) != 0)
        return (uint32_t)hint_first_gpu(m);
    if ((m = candidates & cfg.has_fast_link[src]) != 0 || (m = candidates & cfg.can_access[src]) != 0) {
        const int g = hint_first_gpu(m);
        return g >= 0 ? (uint32_t)g : 0u;
    }
    return candidates ? (uint32_t)(__ffs((int)candidates) - 1) : kHintNoId;
}

// get_hint_for_migration_thrashing :1455-1595
__device__ __forceinline__ uint32_t hint_for_migration(const uvmtier_thrashing_hint_config_t &cfg, uint32_t procs, bool pinned,
                                                       uint32_t pinned_residency, uint32_t dnt, uint32_t throttling_count,
                                                       uint32_t requester, uint32_t preferred, uint32_t resident, uint32_t *residency)
{
    uint32_t type = UVMTIER_THRASHING_HINT_NONE, res = kHintNoId;
    const uint32_t closest = resident ? hint_closest_id(cfg, resident, requester) : kHintNoId;
    const bool can_access_pref = preferred != kHintNoId && (procs & ~cfg.accessible_from[preferred]) == 0;          // :1382-1391
    bool fast = false;
    if (closest != kHintNoId) {                                                                                      // :1393-1422
        uint32_t fast_to = cfg.has_fast_link[closest] & cfg.has_native_atomics[closest];
        fast_to |= closest == 0 ? 1u : hint_smc_range(closest);
        fast = (procs & ~fast_to) == 0;
    }
    if (can_access_pref) {
        type = UVMTIER_THRASHING_HINT_PIN; res = preferred;
    }
    else if (!(preferred != kHintNoId && (procs & (1u << preferred))) && fast) {
        if (closest == 0) {
            if (requester != 0) { type = UVMTIER_THRASHING_HINT_PIN; res = requester; }
        }
        else { type = UVMTIER_THRASHING_HINT_PIN; res = closest; }
    }
    else if (requester == preferred) {
        if (pinned) {
            if (preferred == pinned_residency || preferred == dnt) { type = UVMTIER_THRASHING_HINT_PIN; res = preferred; }
            else type = UVMTIER_THRASHING_HINT_THROTTLE;
        }
        else if (preferred != dnt) type = UVMTIER_THRASHING_HINT_THROTTLE;
        else if (throttling_count >= cfg.pin_threshold) { type = UVMTIER_THRASHING_HINT_PIN; res = preferred; }
    }
    else if (pinned) {
        if (requester == dnt) {
            type = UVMTIER_THRASHING_HINT_PIN;
            if ((procs & ~cfg.accessible_from[requester]) == 0)
                res = requester;
            else {
                uint32_t common = 0xffffffffu, rest = procs;                                                         // :1424-1444
                while (rest) { common &= cfg.can_access[__ffs((int)rest) - 1]; rest &= rest - 1; }
                res = common ? hint_closest_id(cfg, common, requester) : requester;
            }
        }
        else if (pinned_residency != kHintNoId && (cfg.accessible_from[pinned_residency] & (1u << requester))) {
            type = UVMTIER_THRASHING_HINT_PIN; res = pinned_residency;
        }
        else type = UVMTIER_THRASHING_HINT_THROTTLE;
    }
    else if (requester != dnt) type = UVMTIER_THRASHING_HINT_THROTTLE;
    else if (throttling_count >= cfg.pin_threshold) { type = UVMTIER_THRASHING_HINT_PIN; res = requester; }
    if (type == UVMTIER_THRASHING_HINT_PIN && !(cfg.has_memory & (1u << res)))
        res = 0;                                                                                                     // UVM_ID_CPU
    *residency = res;
    return type;
}

// uvm_perf_thrashing_get_hint :1615-1780 for one query.  tally[0..2]: throttles, pins local, pins remote.
__device__ __forceinline__ uvmtier_thrashing_hint_t hint_get(const uvmtier_thrashing_hint_config_t &cfg, const thrash_hint_tables &th,
                                                             const uvmtier_thrashing_query_t &q, uvmtier_thrashing_block_t &blk,
                                                             uint32_t *tally)
{
    const uint32_t slot = q.slot, page = q.page, req = q.requester;
    const unsigned long long ts = q.time_stamp;
    const size_t at = (size_t)slot * 512 + page;
    uvmtier_thrashing_hint_t out;
    out.throttle_end_time_stamp = 0; out.pin_processors = 0; out.type = UVMTIER_THRASHING_HINT_NONE; out.pin_residency = 0xff; out.reserved = 0;
    if (!blk.tracked)
        return out;
    if (blk.last_thrashing_time_stamp != 0 && ts - blk.last_thrashing_time_stamp > cfg.epoch_ns && blk.pinned_count == 0 &&
        blk.thrashing_reset_count < cfg.max_resets) {
        // :1644-1674: a quiet epoch -- the va_block forgets everything and goes back to untracked
        blk.thrashing_reset_count++;
        for (uint32_t p = 0; p < 512; p++) {
            const size_t a = (size_t)slot * 512 + p;
            th.page_word[a] = 0; th.page_procs[a] = 0; th.page_throttle[a] = 0; th.page_hint[a] = 0; th.page_throttled[a] = 0;
            th.page_deadline[a] = 0;
        }
        for (int w = 0; w < 16; w++)
            th.thrashing_pages[(size_t)slot * 16 + w] = 0;
        blk.tracked = 0; blk.num_thrashing_pages = 0; blk.last_processor = (uint16_t)kHintNoId;
        blk.last_time_stamp = 0; blk.last_thrashing_time_stamp = 0;
        return out;
    }
    uint32_t type = UVMTIER_THRASHING_HINT_NONE, residency = kHintNoId;
    {
        const unsigned long long w = th.page_word[at];
        uint32_t h = th.page_hint[at];
        const uint32_t num = (uint32_t)(w >> kHintNumShift) & 7u;
        uint32_t dnt = ((h >> 16) & 0xffu) ? ((h >> 16) & 0xffu) - 1u : kHintNoId;
        const uint32_t pinned_res = ((h >> 8) & 0xffu) ? ((h >> 8) & 0xffu) - 1u : kHintNoId;
        const bool pinned = (w & kHintPinned) != 0;
        bool decided = num < cfg.threshold;                                       // not enough thrashing events yet (:1679-1680)
        if (!decided && (th.page_throttled[at] & (1u << req))) {                  // :1684-1694
            const unsigned long long end = th.page_throttle[at] << 6;
            if (ts < end && req != dnt) {
                type = UVMTIER_THRASHING_HINT_THROTTLE;
                decided = true;
            }
            else {
                th.page_throttled[at] &= ~(1u << req);                            // thrashing_throttle_end_processor :841-861
                if (!th.page_throttled[at]) th.page_throttle[at] = 0;
            }
        }
        if (!decided && ts - ((w & kHintTsMask) << 6) > cfg.lapse_ns && !pinned)  // :1702-1705
            decided = true;
        if (!decided) {
            th.page_procs[at] |= 1u << req;
            if (ts > (th.page_throttle[at] << 6)) {                               // thrashing_throttle_update :772-795
                th.page_throttle[at] = ((ts + cfg.nap_ns) >> 6) & kHintTsMask;
                dnt = dnt == req ? kHintNoId : req;
            }
            else if (dnt == kHintNoId)
                dnt = req;
            h = (h & ~(0xffu << 16)) | ((dnt == kHintNoId ? 0u : dnt + 1u) << 16);
            th.page_hint[at] = h;
            if (pinned && (w & kHintRev) && req != dnt)
                type = UVMTIER_THRASHING_HINT_THROTTLE;                           // revocation thrashing on a pinned page (:1717-1727)
            else
                type = hint_for_migration(cfg, th.page_procs[at], pinned, pinned_res, dnt, h & 0xffu, req,
                                          q.preferred_location == 0xffffu ? kHintNoId : (uint32_t)q.preferred_location,
                                          q.resident_processors, &residency);
        }
    }
    if (type == UVMTIER_THRASHING_HINT_PIN) {                                     // thrashing_pin_page :902-967
        const unsigned long long w = th.page_word[at];
        if (!(w & kHintPinned) || !(q.resident_processors & (1u << residency)))
            hint_throttling_reset(th, at);
        if (!(w & kHintPinned)) {
            if (cfg.pin_ns > 0)
                th.page_deadline[at] = ts + cfg.pin_ns;
            th.page_hint[at] &= ~0xffu;
            th.page_word[at] = w | kHintPinned;
            if (blk.pinned_count != 0xffffu) blk.pinned_count++;
            th.pinned_pages[(size_t)slot * 16 + (page >> 5)] |= 1u << (page & 31);
        }
        th.page_hint[at] = (th.page_hint[at] & ~(0xffu << 8)) | ((residency + 1u) << 8);
        tally[residency == req ? 1 : 2]++;
        out.pin_residency = (uint8_t)residency;
        out.pin_processors = th.page_procs[at];
    }
    else if (type == UVMTIER_THRASHING_HINT_THROTTLE) {                           // thrashing_throttle_processor :805-832
        if (!(th.page_throttled[at] & (1u << req))) {
            th.page_throttled[at] |= 1u << req;
            if (!(th.page_word[at] & kHintPinned) && (th.page_hint[at] & 0xffu) != 0xffu)
                th.page_hint[at]++;
            if (blk.throttling_count != 0xffffffffu) blk.throttling_count++;
        }
        tally[0]++;
        out.throttle_end_time_stamp = th.page_throttle[at] << 6;
    }
    out.type = (uint8_t)type;
    return out;
}

__device__ __forceinline__ bool hint_query_ok(const uvmtier_thrashing_query_t &q, uint32_t n_slots)
{
    return q.slot < n_slots && q.page < 512u && q.requester < 32u && (q.preferred_location < 32u || q.preferred_location == 0xffffu);
}


} // namespace uvmtier_k
