// thrashing_hint.cuh -- kernels of the hint half of the thrashing heuristic; the decision itself (hint_get and what it calls)
// is in thrashing_hint_dev.cuh, which the servicing pass uses too.
#pragma once

#include "service_kernels.cuh"
#include "thrashing_hint_dev.cuh"

namespace uvmtier_k {

// seg[]: first query of every va_block's group (k_thrash_heads, which only reads .slot -- the query has it where the
// event has it).  One thread per group.  counters: [2] dropped, [3] throttles, [4] pins local, [5] pins remote.
UVMTIER_KERNEL void __launch_bounds__(128)
k_thrash_hints(uint32_t n_queries, const uvmtier_thrashing_query_t *__restrict__ q, uint32_t n_slots,
               const uvmtier_thrashing_hint_config_t *__restrict__ cfg_dev, thrash_hint_tables th, const uint32_t *__restrict__ scratch,
               uvmtier_thrashing_hint_t *__restrict__ hints, uint32_t *__restrict__ counters)
{
    const uint32_t *seg = scratch + 4;
    const uint32_t n_seg = scratch[0];
    uint32_t tally[3] = {0, 0, 0}, dropped = 0;
    for (uint32_t s = blockIdx.x * blockDim.x + threadIdx.x; s < n_seg; s += gridDim.x * blockDim.x) {
        const uvmtier_thrashing_hint_config_t &cfg = *cfg_dev;
        uint32_t i = seg[s];
        const uint32_t slot = q[i].slot;
        uvmtier_thrashing_block_t blk = th.blocks[slot];
        for (; i < n_queries && q[i].slot == slot; i++) {
            const uvmtier_thrashing_query_t qi = q[i];
            if (!hint_query_ok(qi, n_slots)) {
                uvmtier_thrashing_hint_t none;
                none.throttle_end_time_stamp = 0; none.pin_processors = 0; none.type = UVMTIER_THRASHING_HINT_NONE; none.pin_residency = 0xff;
                none.reserved = 0;
                hints[i] = none;
                dropped++;
                continue;
            }
            hints[i] = hint_get(cfg, th, qi, blk, tally);
        }
        th.blocks[slot] = blk;
    }
    if (dropped) atomicAdd(&counters[2], dropped);
    if (tally[0]) atomicAdd(&counters[3], tally[0]);
    if (tally[1]) atomicAdd(&counters[4], tally[1]);
    if (tally[2]) atomicAdd(&counters[5], tally[2]);
}

// thrashing_unpin_pages :1822-1906 at time `now`: every pin whose deadline has passed (20 us of timer slack, :1822,1849)
// ends and the page's tracking state is cleared (thrashing_reset_page).  A warp per va_block, lanes over its pages.
// counters[1] += pages reset.
UVMTIER_KERNEL void __launch_bounds__(128)
k_thrash_unpin_expired(uint32_t n_slots, unsigned long long now, thrash_hint_tables th, uint32_t *__restrict__ counters)
{
    const int lane = threadIdx.x & 31;
    const uint32_t warps = gridDim.x * (blockDim.x / 32);
    uint32_t resets = 0;
    for (uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; s < n_slots; s += warps) {
        const uint32_t pinned_w = lane < 16 ? th.pinned_pages[(size_t)s * 16 + lane] : 0u;
        if (!__any_sync(0xffffffffu, pinned_w != 0u))
            continue;
        uint32_t gone = 0;
        for (uint32_t base = 0; base < 512; base += 32) {
            const uint32_t p = base + lane;
            const size_t at = (size_t)s * 512 + p;
            const uint32_t pw = __shfl_sync(0xffffffffu, pinned_w, base >> 5);
            bool hit = false;
            if ((pw >> lane) & 1u) {
                const unsigned long long dl = th.page_deadline[at];
                hit = dl != 0 && dl <= now + 20000ull;
                if (hit) {
                    hint_reset_page_extras(th, at);
                    th.page_word[at] = 0;
                    th.page_procs[at] = 0;
                    resets++;
                }
            }
            const uint32_t b = __ballot_sync(0xffffffffu, hit);
            if (lane == (int)(base >> 5) && b) {
                th.pinned_pages[(size_t)s * 16 + lane] = pinned_w & ~b;
                const uint32_t tw = th.thrashing_pages[(size_t)s * 16 + lane];
                th.thrashing_pages[(size_t)s * 16 + lane] = tw & ~b;
                gone += __popc(tw & b);
            }
        }
        gone = __reduce_add_sync(0xffffffffu, gone);
        if (lane == 0 && gone)
            th.blocks[s].num_thrashing_pages = (uint16_t)(th.blocks[s].num_thrashing_pages - gone);
    }
    resets = __reduce_add_sync(0xffffffffu, resets);
    if (lane == 0 && resets)
        atomicAdd(&counters[1], resets);
}

} // namespace uvmtier_k
