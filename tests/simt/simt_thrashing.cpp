// simt_thrashing.cpp -- runs the thrashing kernels (thrashing_kernels.cuh, thrashing_hint.cuh) under the CPU SIMT emulator.
// TEST INFRASTRUCTURE ONLY.
#include "simt.h"

#include <algorithm>
#include <vector>

#include "thrashing_kernels.cuh"

using namespace uvmtier_k;

// tables: blocks, page_word, page_procs, thrashing_pages, page_throttle, page_hint, page_throttled, page_deadline,
// pinned_pages, counters (the order of tests/thrashlib.py's HintState.NAMES); page_throttle NULL = detector only
static thrash_hint_tables tables_of(void *const *tb)
{
    thrash_hint_tables th;
    th.blocks = (uvmtier_thrashing_block_t *)tb[0];
    th.page_word = (unsigned long long *)tb[1];
    th.page_procs = (uint32_t *)tb[2];
    th.thrashing_pages = (uint32_t *)tb[3];
    th.page_throttle = (unsigned long long *)tb[4];
    th.page_hint = (uint32_t *)tb[5];
    th.page_throttled = (uint32_t *)tb[6];
    th.page_deadline = (unsigned long long *)tb[7];
    th.pinned_pages = (uint32_t *)tb[8];
    return th;
}

extern "C" int simt_thrash_events(uint32_t n_slots, uint32_t n_events, const uvmtier_perf_event_t *ev, unsigned long long lapse_ns,
                                  uint32_t threshold, void *const *tb)
{
    const thrash_hint_tables th = tables_of(tb);
    uint32_t *counters = (uint32_t *)tb[9];
    std::vector<uint32_t> scratch(4 + 2 * (size_t)n_slots, 0u);
    std::fill(scratch.begin() + 4 + n_slots, scratch.end(), 0xffffffffu);
    if (!n_events)
        return 0;
    simt::launch(dim3(2), dim3(kThrashThreads), [&] { k_thrash_heads_min(n_events, ev, n_slots, scratch.data()); });
    simt::launch(dim3(2), dim3(kThrashThreads), [&] { k_thrash_heads(n_events, ev, n_slots, scratch.data(), counters); });
    simt::launch(dim3(2), dim3(kThrashThreads), [&] {
        k_thrash_events(n_events, ev, n_slots, lapse_ns, threshold, th.blocks, th.page_word, th.page_procs, th.thrashing_pages,
                        scratch.data(), counters, th);
    });
    return 0;
}

extern "C" int simt_thrash_hints(uint32_t n_slots, uint32_t n, const uvmtier_thrashing_query_t *q, const uvmtier_thrashing_hint_config_t *cfg,
                                 void *const *tb, uvmtier_thrashing_hint_t *hints)
{
    const thrash_hint_tables th = tables_of(tb);
    uint32_t *counters = (uint32_t *)tb[9];
    std::vector<uint32_t> scratch(4 + 2 * (size_t)n_slots, 0u);
    std::fill(scratch.begin() + 4 + n_slots, scratch.end(), 0xffffffffu);
    if (!n)
        return 0;
    simt::launch(dim3(2), dim3(kThrashThreads), [&] { k_thrash_heads_min(n, (const uvmtier_perf_event_t *)q, n_slots, scratch.data()); });
    simt::launch(dim3(2), dim3(kThrashThreads), [&] { k_thrash_heads(n, (const uvmtier_perf_event_t *)q, n_slots, scratch.data(), counters); });
    simt::launch(dim3(2), dim3(128), [&] { k_thrash_hints(n, q, n_slots, cfg, th, scratch.data(), hints, counters); });
    return 0;
}

extern "C" int simt_thrash_unpin_expired(uint32_t n_slots, unsigned long long now, void *const *tb)
{
    const thrash_hint_tables th = tables_of(tb);
    simt::launch(dim3(2), dim3(128), [&] { k_thrash_unpin_expired(n_slots, now, th, (uint32_t *)tb[9]); });
    return 0;
}
