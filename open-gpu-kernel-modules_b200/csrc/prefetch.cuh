// prefetch.cuh -- device-side core of the uvm_perf_prefetch bitmap-tree hint, one va_block per
// half-warp (see mask512.cuh).  Reference: UVM/uvm_perf_prefetch.c (cited per function below).
//
// Formulation.  The reference walks leaf->root for every faulted page and keeps the largest
// ancestor whose occupancy passes the threshold (compute_prefetch_region, :102-146).  For blocks
// whose start is big-page aligned (tree offset 0) the union over faulted pages of those regions
// equals the union of ALL tree nodes that (a) pass the threshold and (b) contain at least one
// faulted page: every such node is inside the largest passing ancestor of the faulted page it
// contains, and every largest passing ancestor is such a node.  (Unaligned blocks: see the quirk
// handling in pf_prefetch_mask.)  (The reference's early-out at :316 never changes the union, because any later faulted
// page already lies inside the region that reached the block end.)  So all 1023 nodes of all 10
// levels are evaluated at once: levels with <=32 leaves per node by SWAR field arithmetic inside
// each 32-bit word, levels with 64..512 leaves by a shuffle butterfly over per-word popcounts.
#pragma once

#include "mask512.cuh"
#include "uvmtier.h"

// Union (in tree coordinates, one word per lane) of all nodes that pass `thr` and contain a bit
// of `fw`.  `w` must have no bits at or beyond leaf_count.
//
// offset > 0 (block start not big-page aligned, :275-285): nodes that start below `offset` but
// reach past it get the reference's clamp quirk (:124-127: first = offset rather than
// first - offset).  Those nodes all contain leaf `offset`, so all live in lane 0's view; they are
// left out of the returned union and reported as *bq = the largest outer bound among them
// (0 if none), valid in every lane of the group.
__device__ __forceinline__ uint32_t pf_tree_select(uint32_t w, uint32_t fw, int leaf_count, int offset,
                                                   uint32_t thr, int l, int *bq)
{
    // Minimum population for a full node of 1,2,4,8,16,32 leaves: cnt*100 > size*thr (:118).
    const uint32_t m0 = thr / 100u + 1u;
    const uint32_t m1 = (2u * thr) / 100u + 1u;
    const uint32_t m2 = (4u * thr) / 100u + 1u;
    const uint32_t m3 = (8u * thr) / 100u + 1u;
    const uint32_t m4 = (16u * thr) / 100u + 1u;
    const uint32_t m5 = (32u * thr) / 100u + 1u;

    // per-field populations
    const uint32_t c2 = w - ((w >> 1) & 0x55555555u);
    const uint32_t c2e = c2 & 0x33333333u, c2o = (c2 >> 2) & 0x33333333u;
    const uint32_t c4 = c2e + c2o;
    const uint32_t c8 = (c4 + (c4 >> 4)) & 0x0f0f0f0fu;
    const uint32_t c16 = (c8 + (c8 >> 8)) & 0x00ff00ffu;
    const uint32_t c32 = (c16 + (c16 >> 16)) & 0x3fu;
    // "field holds a faulted page" flags, at the field's lowest bit
    const uint32_t a2 = (fw | (fw >> 1)) & 0x55555555u;
    const uint32_t a4 = (a2 | (a2 >> 2)) & 0x11111111u;
    const uint32_t a8 = (a4 | (a4 >> 4)) & 0x01010101u;
    const uint32_t a16 = (a8 | (a8 >> 8)) & 0x00010001u;
    // selected-node flags at the node's lowest bit (biased add: the field's top bit = "count >= m")
    uint32_t s1 = m0 <= 1u ? (w & fw) : 0u;
    const uint32_t d2e = ((c2e + (8u - m1) * 0x11111111u) >> 3) & 0x11111111u;
    const uint32_t d2o = ((c2o + (8u - m1) * 0x11111111u) >> 3) & 0x11111111u;
    uint32_t s2 = (d2e & a2) | ((d2o & (a2 >> 2)) << 2);
    uint32_t s4 = (((c4 + (8u - m2) * 0x11111111u) >> 3) & 0x11111111u) & a4;
    uint32_t s8 = (((c8 + (128u - m3) * 0x01010101u) >> 7) & 0x01010101u) & a8;
    uint32_t s16 = (((c16 + (0x8000u - m4) * 0x00010001u) >> 15) & 0x00010001u) & a16;
    uint32_t s32 = (c32 >= m5 && fw != 0u) ? 1u : 0u;

    int my_bq = 0;
    const bool quirk_lane = offset > 0 && l == 0;
    if (quirk_lane) {
        // the node of each level that holds leaf `offset` and starts below it
        const uint32_t o = (uint32_t)offset;
        uint32_t b;
        if (o & 1u)  { b = 1u << (o & ~1u);  if (s2 & b)  { s2 &= ~b;  my_bq = max(my_bq, min((int)(o & ~1u) + 2, leaf_count)); } }
        if (o & 3u)  { b = 1u << (o & ~3u);  if (s4 & b)  { s4 &= ~b;  my_bq = max(my_bq, min((int)(o & ~3u) + 4, leaf_count)); } }
        if (o & 7u)  { b = 1u << (o & ~7u);  if (s8 & b)  { s8 &= ~b;  my_bq = max(my_bq, min((int)(o & ~7u) + 8, leaf_count)); } }
        if (o & 15u) { b = 1u << (o & ~15u); if (s16 & b) { s16 &= ~b; my_bq = max(my_bq, min((int)(o & ~15u) + 16, leaf_count)); } }
        if (s32) { s32 = 0u; my_bq = max(my_bq, min(32, leaf_count)); }
    }

    uint32_t fill = s1 | ((s2 << 2) - s2) | ((s4 << 4) - s4) | ((s8 << 8) - s8) | ((s16 << 16) - s16) |
                    (s32 ? 0xffffffffu : 0u);

    // The one node per level that is cut short by leaf_count is easier to pass: its size is the
    // clamped one (:94-100 with uvm_perf_utils.h:170-178).  Levels of <=32 leaves: it sits in the
    // last word.
    const int rem = leaf_count & 31;
    if (rem && l == (leaf_count >> 5)) {
        const int lastbit = rem - 1;
#pragma unroll
        for (int k = 1; k <= 5; k++) {
            const int s = 1 << k;
            const int ns = lastbit & ~(s - 1);
            const int cover = rem - ns;
            if (cover < s) {
                const uint32_t fm = hw_ones(s) << ns;
                const uint32_t cnt = __popc(w & fm);
                if ((fw & fm) && cnt * 100u > (uint32_t)cover * thr) {
                    if (quirk_lane && ns < offset)
                        my_bq = max(my_bq, leaf_count);
                    else
                        fill |= fm;
                }
            }
        }
    }

    // Levels of 64..512 leaves: butterfly over the group's words; the general test also covers
    // the clamped node of these levels.
    uint32_t cnt = c32;
    uint32_t any = fw != 0u;
#pragma unroll
    for (int d = 1; d <= 8; d <<= 1) {
        cnt += __shfl_xor_sync(HW_FULL, cnt, d);
        any |= __shfl_xor_sync(HW_FULL, any, d);
        const int ns = (l & ~(2 * d - 1)) * 32;
        const int cover = min(64 * d, leaf_count - ns);
        if (any && cover > 0 && cnt * 100u > (uint32_t)cover * thr) {
            if (offset > 0 && ns == 0)
                my_bq = max(my_bq, cover);
            else
                fill = 0xffffffffu;
        }
    }
    *bq = __shfl_sync(HW_FULL, my_bq, 0, 16);
    return fill;
}

// compute_prefetch_mask (:299-319) on a tree laid out like init_bitmap_tree_from_region (:222-238)
// [+ the big-page offset of update_bitmap_tree_from_va_block (:275-285)].
//   pages_t : tree bitmap in tree coordinates (leaf = page - max_first + offset)
//   fwork   : faulted pages to walk from, in PAGE coordinates, already cut to faulted_region
// Returns the prefetch mask in page coordinates, clamped to [max_first, max_outer).
__device__ __forceinline__ uint32_t pf_prefetch_mask(uint32_t pages_t, uint32_t fwork, int max_first, int max_outer,
                                                     int offset, uint32_t thr, int l)
{
    const int leaf_count = max_outer - max_first + offset;
    // The two va_blocks of a warp may differ: branch on warp-uniform predicates only, the shifts
    // shuffle with the full mask (a shift by 0 is the identity).
    const bool any_mf = __any_sync(HW_FULL, max_first != 0);
    const bool any_off = __any_sync(HW_FULL, offset != 0);
    uint32_t fw_t = fwork;
    if (any_mf)
        fw_t = hw_shr(fw_t, max_first, l);
    if (any_off)
        fw_t = hw_shl(fw_t, offset, l);
    pages_t &= hw_range(0, leaf_count, l);
    fw_t &= hw_range(0, leaf_count, l);

    int bq;
    uint32_t fill = pf_tree_select(pages_t, fw_t, leaf_count, offset, thr, l, &bq);
    if (any_off) {
        // A faulted page inside a passing quirk node is answered by that node ALONE (the walk keeps
        // only the largest passing ancestor, and every ancestor of a quirk node is a quirk node),
        // so smaller passing nodes around it must not add their own, unclamped, regions.  Quirk
        // nodes are nested; the largest one is [.., bq).  Redo the selection without the faults
        // it captured: no quirk node can be selected any more, the rest is unchanged.
        int bq2;
        fw_t &= ~hw_range(0, bq, l);
        fill = pf_tree_select(pages_t, fw_t, leaf_count, offset, thr, l, &bq2);
    }

    // back to page coordinates + the clamps of :123-143
    if (any_off)
        fill = hw_shr(fill, offset, l);
    if (any_mf)
        fill = hw_shl(fill, max_first, l);
    if (bq - offset > offset)                       // quirk nodes: [offset, outer - offset)
        fill |= hw_range(offset, bq - offset, l);
    return fill & hw_range(max_first, max_outer, l);
}

// uvm_perf_prefetch_prenotify_fault_migrations (:327-413) + the residency decision of
// uvm_perf_prefetch_get_hint_va_block (:477-500) for one managed va_block per half-warp.
//   f, res, thr, mapped : this lane's word of faulted / resident(dst) / thrashing / cpu read-mapped masks
//   flags               : UVMTIER_PF_* ; big_pages: pages per big page of the destination (16 or 32)
//   *pi                 : va_block->prefetch_info, updated;  *hint: hint residency (UVMTIER_ID_INVALID = none)
// Returns this lane's word of prefetch_pages_mask.  Warp-synchronous: all 32 lanes must call it.
__device__ __forceinline__ uint32_t pf_get_hint_block(uint32_t f, uint32_t res, uint32_t thr, uint32_t mapped, bool have_thr,
                                                      uint64_t start, int num_pages, int ff, int fo, uint32_t dst,
                                                      uint32_t pref, uint32_t flags, uint32_t big_pages, uint32_t threshold,
                                                      uint32_t min_faults, uvmtier_prefetch_info_t *pi, uint32_t *hint, int lane)
{
    const int l = lane & 15;
    if (pi->last_migration_proc_id != dst) {                               // :340-343
        pi->last_migration_proc_id = (uint16_t)dst;
        pi->fault_migrations_to_last_proc = 0;
    }
    const bool first_touch = (flags & UVMTIER_PF_BLOCK_RESIDENT_EMPTY) && dst == pref;   // :363-366

    // ---- tree: init (:222-238), big-page alignment (:275-285), grow (:164-220) ----
    // Block start inside its big page, in pages.  big_pages is a power of two (16 or 32).
    const int spb = (int)((start / UVMTIER_PAGE_SIZE) & (big_pages - 1));
    const int bf = spb ? (int)big_pages - spb : 0;                         // first whole big page
    const int bo = num_pages - ((spb + num_pages) & ((int)big_pages - 1)); // end of the last whole one
    const bool has_big = bo > bf;
    const int offset = (has_big && bf > 0) ? spb : 0;
    const int leaf_count = num_pages + offset;

    const bool any_off = __any_sync(HW_FULL, offset != 0);
    uint32_t pages_t = res | f, f_t = f, thr_t = thr;
    if (any_off) {
        pages_t = hw_shl(pages_t, offset, l);
        f_t = hw_shl(f_t, offset, l);
        thr_t = hw_shl(thr_t, offset, l);
    }
    const uint32_t valid_t = hw_range(offset, leaf_count, l);
    uint32_t grow;
    if (big_pages == 16) {
        const uint32_t lo = ((f_t & 0xffffu) && !(thr_t & 0xffffu)) ? 0xffffu : 0u;
        const uint32_t hi = ((f_t >> 16) && !(thr_t >> 16)) ? 0xffff0000u : 0u;
        grow = lo | hi;
    }
    else {
        grow = (f_t && !thr_t) ? 0xffffffffu : 0u;
    }
    // no whole big page in the block: the block is one region (:173-181)
    const bool any_f = hw_any(f, lane), any_thr = hw_any(thr, lane);
    if (!has_big)
        grow = (any_f && !any_thr) ? 0xffffffffu : 0u;
    pages_t |= grow & valid_t;

    // ---- walk (:378-390) ----
    const uint32_t fwork = f & ~thr & hw_range(ff, fo, l);
    const uint32_t walked = pf_prefetch_mask(pages_t, fwork, 0, num_pages, offset, threshold, l);
    uint32_t o = first_touch ? hw_range(0, num_pages, l) : walked;

    // ---- post filters (:392-408) ----
    o &= ~f;
    if (dst == UVMTIER_ID_CPU)
        o &= ~(res & mapped);
    if (have_thr)
        o &= ~thr;
    const uint32_t weight = hw_weight(f & hw_range(ff, fo, l));
    pi->fault_migrations_to_last_proc = (uint16_t)(pi->fault_migrations_to_last_proc + weight);   // :410
    const uint32_t pending = hw_weight(o);
    *hint = (pi->fault_migrations_to_last_proc >= min_faults && pending > 0) ? pi->last_migration_proc_id   // :477-500
                                                                               : (uint32_t)UVMTIER_ID_INVALID;
    return o;
}
