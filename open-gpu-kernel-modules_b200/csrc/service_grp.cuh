// service_grp.cuh -- the servicing pass with a GROUP OF G LANES per va_block (G = 1, 2, 4, 8 or 16; W = 16 / G
// mask words per lane).  Same semantics and reference citations as k_service_blocks in service_kernels.cuh.
//
// Why a group: a va_block's runs are a sequential chain (each run sees the state the previous one left), so the
// parallelism of the pass is (blocks touched) x (lanes per block).  One thread per block (G = 1) keeps the six
// 512-bit masks in 96 registers: 254 registers per thread, 7 warps per SM and a fully unrolled body that misses the
// instruction cache (ncu r1_f: issue slots 27 % busy, top stalls "wait" and "no instruction").  One warp per block
// repeats the scalar bookkeeping 32 times.  G = 4 keeps 4 words per lane (mask operations cost 4 instructions,
// cross-lane steps are two shuffles), quadruples the warps in flight and shrinks the code fourfold.
//
// Why lockstep: the 32 / G groups of a warp walk through their blocks' runs TOGETHER -- every loop bound is the
// maximum over the warp, a group that has nothing to do in a step computes on zeros and its stores are predicated
// off.  Under SIMT a group that skipped a step would only wait for the others anyway; but with sub-warp member
// masks every shuffle released its group separately, the groups never met again, and everything after the first
// shuffle ran with 4 of 32 lanes (ncu r1_g: 14 active threads per instruction, 2.6 in the copy-run loop).  With
// full-mask, width-G shuffles the warp stays converged and claims run-array space with one atomic per step.
#pragma once

#include "service_kernels.cuh"

namespace uvmtier_k {

constexpr int kGrpThreads = 128;
constexpr int kGrpLanes = 4;                    // lanes per va_block the library runs with (W = 4 mask words per lane)
constexpr int kGrpMinCtas = 4;                  // CTAs per SM the register allocation leaves room for
constexpr uint32_t kFullWarp = 0xffffffffu;

// ---- reductions over the G lanes of each group; every lane of the warp takes part (xor butterflies stay inside
//      the aligned group of G lanes) ----
template <int G> __device__ __forceinline__ uint32_t seg_add(uint32_t v)
{
#pragma unroll
    for (int d = 1; d < G; d <<= 1) v += __shfl_xor_sync(kFullWarp, v, d);
    return v;
}
template <int G> __device__ __forceinline__ uint32_t seg_or(uint32_t v)
{
#pragma unroll
    for (int d = 1; d < G; d <<= 1) v |= __shfl_xor_sync(kFullWarp, v, d);
    return v;
}
template <int G> __device__ __forceinline__ uint32_t seg_min(uint32_t v)
{
#pragma unroll
    for (int d = 1; d < G; d <<= 1) v = min(v, __shfl_xor_sync(kFullWarp, v, d));
    return v;
}
template <int G> __device__ __forceinline__ uint32_t seg_max(uint32_t v)
{
#pragma unroll
    for (int d = 1; d < G; d <<= 1) v = max(v, __shfl_xor_sync(kFullWarp, v, d));
    return v;
}

template <int W> __device__ __forceinline__ void ld_words(const uint32_t *p, uint32_t (&v)[W])
{
    if constexpr (W % 4 == 0) {
#pragma unroll
        for (int q = 0; q < W / 4; q++) {
            const uint4 x = ((const uint4 *)p)[q];
            v[4 * q] = x.x; v[4 * q + 1] = x.y; v[4 * q + 2] = x.z; v[4 * q + 3] = x.w;
        }
    }
    else if constexpr (W == 2) {
        const uint2 x = *(const uint2 *)p;
        v[0] = x.x; v[1] = x.y;
    }
    else {
        v[0] = *p;
    }
}
template <int W> __device__ __forceinline__ void st_words(uint32_t *p, const uint32_t (&v)[W])
{
    if constexpr (W % 4 == 0) {
#pragma unroll
        for (int q = 0; q < W / 4; q++) ((uint4 *)p)[q] = make_uint4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
    }
    else if constexpr (W == 2) {
        *(uint2 *)p = make_uint2(v[0], v[1]);
    }
    else {
        *p = v[0];
    }
}

// Prefetch hint of a whole aligned va_block held by a group (uvm_perf_prefetch_prenotify_fault_migrations
// :327-413 + :477-500, GPU destination, no thrashing).  Called by all 32 lanes; `has` = this group's run has
// serviced faults (otherwise f is zero, nothing is selected and *pi stays as it is).  f = faulted (new_residency)
// words of this lane, gres = resident on the GPU.  Tree nodes of up to 32 leaves: decided per word from its
// population alone (the filled big pages pass every level below);
// up to 32 W leaves: sums of this lane's per-word populations; above: an xor butterfly over the group (at step d
// the sum covers the aligned 2d-lane node this lane belongs to).  Writes the prefetch mask to out[],
// *n_prefetch = its population over the group, updates *pi (kept identically by every lane of the group), returns
// the hint residency.
template <int G>
__device__ __forceinline__ uint32_t pf_get_hint_grp(bool has, const uint32_t (&f)[16 / G], const uint32_t (&gres)[16 / G],
                                                    bool block_empty, uint32_t gpu_id, uint32_t pref, uint32_t thr,
                                                    uint32_t min_faults, uvmtier_prefetch_info_t *pi, uint32_t (&out)[16 / G],
                                                    uint32_t *n_prefetch)
{
    constexpr int W = 16 / G;
    const bool first_touch = has && block_empty && gpu_id == pref;          // :363-366
    uint32_t pages[W], cnt[W];
    uint32_t anyw = 0, weight = 0, lane_cnt = 0;
#pragma unroll
    for (int w = 0; w < W; w++) {
        // tree = resident | faulted, every 64 KB big page holding a fault is filled (:164-220, 64 KB = 16 pages)
        pages[w] = gres[w] | f[w];
        pages[w] |= (f[w] & 0xffffu) ? 0xffffu : 0u;
        pages[w] |= (f[w] >> 16) ? 0xffff0000u : 0u;
        out[w] = 0;
        cnt[w] = __popc(pages[w]);
        lane_cnt += cnt[w];
        anyw |= (f[w] != 0u) ? (1u << w) : 0u;
        weight += __popc(f[w]);
    }
    // nodes of 2 .. 32 leaves.  Every big page that holds a fault is filled above, so each node of up to 16 leaves
    // around a fault is full and passes (count * 100 > size * thr) for every threshold below 100, and none does at
    // 100; what is left to decide inside the word is its node of 32 leaves.
    if (thr < 100u) {
#pragma unroll
        for (int w = 0; w < W; w++) {
            const uint32_t fill = ((f[w] & 0xffffu) ? 0xffffu : 0u) | ((f[w] >> 16) ? 0xffff0000u : 0u);
            out[w] = fill | ((f[w] != 0u && cnt[w] * 100u > 32u * thr) ? 0xffffffffu : 0u);
        }
    }
    // nodes of 64 .. 32 W leaves: this lane's words
#pragma unroll
    for (int span = 2; span <= W; span <<= 1) {
#pragma unroll
        for (int j = 0; j < W; j += span) {
            uint32_t c = 0;
#pragma unroll
            for (int k = 0; k < span; k++) c += cnt[j + k];
            const uint32_t wm = ((1u << span) - 1u) << j;
            if ((anyw & wm) && c * 100u > (uint32_t)(32 * span) * thr) {
#pragma unroll
                for (int k = 0; k < span; k++) out[j + k] = 0xffffffffu;
            }
        }
    }
    // nodes of 64 W .. 512 leaves: low half = pages in the node, high half = lanes of the node holding a fault
    if constexpr (G > 1) {
        uint32_t v = lane_cnt | (anyw ? 0x10000u : 0u);
        bool full = false;
#pragma unroll
        for (int d = 1; d < G; d <<= 1) {
            v += __shfl_xor_sync(kFullWarp, v, d);
            if ((v >> 16) && (v & 0xffffu) * 100u > (uint32_t)(64 * W * d) * thr)
                full = true;
        }
        if (full) {
#pragma unroll
            for (int w = 0; w < W; w++) out[w] = 0xffffffffu;
        }
    }
    uint32_t pending = 0;
#pragma unroll
    for (int w = 0; w < W; w++) {
        if (first_touch) out[w] = 0xffffffffu;
        out[w] &= ~f[w];                                                    // :392
        pending += __popc(out[w]);
    }
    const uint32_t both = seg_add<G>(pending | (weight << 16));             // each <= 512
    pending = both & 0xffffu;
    weight = both >> 16;
    *n_prefetch = pending;
    if (!has)
        return UVMTIER_ID_INVALID;
    if (pi->last_migration_proc_id != gpu_id) {                            // :340-343
        pi->last_migration_proc_id = (uint16_t)gpu_id;
        pi->fault_migrations_to_last_proc = 0;
    }
    pi->fault_migrations_to_last_proc = (uint16_t)(pi->fault_migrations_to_last_proc + weight);   // :410
    return (pi->fault_migrations_to_last_proc >= min_faults && pending > 0) ? pi->last_migration_proc_id
                                                                            : (uint32_t)UVMTIER_ID_INVALID;
}

// Claims n consecutive elements of the run array for every lane of the warp (n = 0: none) with ONE atomic: a
// single counter would otherwise serialise a million of them.  The exclusive prefix over the lanes is summed
// bit-plane by bit-plane with ballots (the values are small).  In two halves, so that the atomic's round trip to
// L2 overlaps whatever the caller puts between them: _issue returns lane 0's pending answer and this lane's
// prefix, _collect turns them into the calling lane's first element.
__device__ __forceinline__ uint32_t svc_claim_runs_issue(uint32_t *counter, uint32_t n, int lane, uint32_t *before_out)
{
    const uint32_t below = (1u << lane) - 1u;
    uint32_t before = 0, total = 0;
    for (int j = 0; __any_sync(kFullWarp, (n >> j) != 0u); j++) {
        const uint32_t plane = __ballot_sync(kFullWarp, (n >> j) & 1u);
        before += (uint32_t)__popc(plane & below) << j;
        total += (uint32_t)__popc(plane) << j;
    }
    *before_out = before;
    uint32_t base = 0;
    if (lane == 0)
        base = atomicAdd(counter, total);
    return base;
}
__device__ __forceinline__ uint32_t svc_claim_runs_collect(uint32_t pending_base, uint32_t before)
{
    return __shfl_sync(kFullWarp, pending_base, 0) + before;
}

// ------------------------------------------------------------------------------------------------
// G lanes per state slot; the slot's run list is sorted.  Slots whose runs hold heavy_min or more entries in
// total are left to the entry-parallel kernel (k_service_blocks).
// M = CTAs per SM the register allocation has to leave room for (occupancy against spills; measured, see service.cu).
template <int G, int M>
__global__ void __launch_bounds__(kGrpThreads, M)
k_service_blocks_grp(service_kparams prm, const uvmtier_ordered_fault_t *__restrict__ ordered,
                     const uvmtier_fault_group_t *__restrict__ groups, uvmtier_fault_batch_t *__restrict__ batches,
                     block_state_ptrs st, const uint2 *__restrict__ cw, const uint32_t *__restrict__ offset,
                     const uint32_t *__restrict__ list, const uint32_t *__restrict__ light,
                     const uint32_t *__restrict__ n_light, uint16_t *__restrict__ entry_flags,
                     uvmtier_service_group_t *__restrict__ service, uvmtier_copy_run_t *__restrict__ runs,
                     uint32_t *__restrict__ counters, uint32_t *__restrict__ work)
{
    constexpr int W = 16 / G;
    const int lane = threadIdx.x & 31, gl = lane & (G - 1), gb = lane & ~(G - 1);
    const uint32_t n_slots_light = *n_light;
    // The grid is one wave of resident CTAs; every warp keeps taking the next 32 / G slots of light[] from a counter
    // until the list is used up.  light[] holds the slots this kernel services, longest run list first: the groups of
    // a warp then have about the same number of steps (a warp takes as many as its longest list), and a warp that
    // finishes early takes up the next, shorter, lists -- no second wave of CTAs on a half-empty GPU.
    for (;;) {
    uint32_t item = 0;
    if (lane == 0)
        item = atomicAdd(work, 1u);
    item = __shfl_sync(kFullWarp, item, 0);
    if (item * (32u / G) >= n_slots_light)
        break;                                                            // the whole warp
    const uint32_t gid = item * (32u / G) + (uint32_t)lane / G;
    uint32_t slot = 0, cnt = 0;
    if (gid < n_slots_light) {
        slot = light[gid];
        cnt = cw[slot].x;
    }
    const uint32_t cnt_max = __reduce_max_sync(kFullWarp, cnt);
    if (!cnt_max)
        continue;
    const bool owns = cnt != 0;
    const uint32_t word0 = (uint32_t)gl * W;                              // first mask word of this lane
    const size_t wbase = (size_t)slot * 16 + word0;
    uint32_t gres[W], c0[W], c1[W], p0[W], p1[W], p2[W];
#pragma unroll
    for (int w = 0; w < W; w++) gres[w] = c0[w] = c1[w] = p0[w] = p1[w] = p2[w] = 0;
    uint32_t logical = PROT_RWA, pref = 0, off = 0;
    uvmtier_prefetch_info_t pi;
    pi.last_migration_proc_id = 0;
    pi.fault_migrations_to_last_proc = 0;
    // Three runs in flight: the loads of run k+1's first entry, run k+2's group and run k+3's id are issued before
    // run k is worked on, so the list -> group -> entry chain is off the critical path.
    const uint4 *graw = (const uint4 *)groups, *oraw = (const uint4 *)ordered;
    uint32_t t0 = 0, t1 = 0, t2 = 0, b0 = 0;
    uint4 g0 = make_uint4(0, 0, 0, 0), g1 = g0, e0 = g0;
    if (owns) {
        off = offset[slot];
        ld_words<W>(st.gpu_resident + wbase, gres);
        ld_words<W>(st.cpu_resident[0] + wbase, c0);
        ld_words<W>(st.cpu_resident[1] + wbase, c1);
        ld_words<W>(st.gpu_pte[0] + wbase, p0);
        ld_words<W>(st.gpu_pte[1] + wbase, p1);
        ld_words<W>(st.gpu_pte[2] + wbase, p2);
        logical = st.logical_prot[slot];
        pref = st.preferred_location[slot];
        pi = st.prefetch_info[slot];
        t0 = list[off];
        t1 = cnt > 1 ? list[off + 1] : t0;
        t2 = cnt > 2 ? list[off + 2] : t0;
        g0 = graw[t0];
        g1 = graw[t1];
        b0 = svc_batch_of(prm, t0);
        e0 = oraw[b0 * prm.B + (g0.z >> 16)];
    }

    for (uint32_t k = 0; k < cnt_max; k++) {
        const bool live = k < cnt;                                        // this group still has a run to service
        uint32_t t3 = t0, b1 = b0;
        uint4 g2 = g1, e1 = e0;
        if (live) {
            // (plain loads: forcing their issue point with volatile inline-asm loads measured 4 % slower)
            t3 = k + 3 < cnt ? list[off + k + 3] : t0;
            g2 = graw[t2];
            b1 = svc_batch_of(prm, t1);
            e1 = oraw[b1 * prm.B + (g1.z >> 16)];
        }
        const uint32_t t = t0, b = b0, base = b * prm.B;
        const uint32_t first = g0.z >> 16, n = live ? (g0.w & 0xffffu) : 0u;
        const uint32_t n_max = __reduce_max_sync(kFullWarp, n);

        // ---- service_fault_batch_block_locked :1433-1580 (every lane of the group, on the same entry) ----
        uint32_t newres[W];
#pragma unroll
        for (int w = 0; w < W; w++) newres[w] = 0;
        uint32_t page_faults = 0, dup_sum = 0, ip_sum = 0, prev_page = 0xffffffffu, state = 0;
        uint4 e = e0;
        for (uint32_t i = 0; i < n_max; i++) {
            const bool on = i < n;
            uint4 e_next = e;
            if (i + 1 < n)
                e_next = oraw[base + first + i + 1];
            const uint32_t acc = (e.w >> 24) & 7u, num_instances = e.w & 0xffffu, access_type_mask = (e.w >> 16) & 0xffu;
            const uint32_t page = on ? (e.x - g0.x) >> 12 : 0u;       // fault_address - block_start < 2 MB
            const bool dup = on && i > 0 && page == prev_page;
            // is the page already mapped for READ / WRITE / ATOMIC?  Its mask words sit on lane `owner`.
            const uint32_t wi = page >> 5, sh = page & 31u;
            const int owner = (int)(wi / W), lw = (int)(wi % W);
            uint32_t a0 = p0[0], a1 = p1[0], a2 = p2[0];
#pragma unroll
            for (int w = 1; w < W; w++)
                if (w == lw) { a0 = p0[w]; a1 = p1[w]; a2 = p2[w]; }
            uint32_t auth = ((a0 >> sh) & 1u) | (((a1 >> sh) & 1u) << 1) | (((a2 >> sh) & 1u) << 2);
            if (G > 1)
                auth = __shfl_sync(kFullWarp, auth, gb + owner);
            if (prm.non_replayable)
                auth = 0;
            const svc_entry_eval ev = svc_eval_entry(acc, num_instances, access_type_mask, dup, state, logical, auth);
            if (on) {
                dup_sum += dup ? num_instances : num_instances - 1u;  // :1200-1203
                ip_sum += ev.ip_add;
                if (ev.serviced) {
                    const uint32_t mine = gl == owner ? (1u << sh) : 0u;
#pragma unroll
                    for (int w = 0; w < W; w++) newres[w] |= w == lw ? mine : 0u;
                    page_faults++;
                }
                if (gl == 0)
                    entry_flags[base + first + i] = (uint16_t)(prm.non_replayable ? ev.flags & ~(1u << UVMTIER_SF_CANCEL_SHIFT) : ev.flags);
                state = ev.state;
                prev_page = page;
            }
            e = e_next;
        }
        if (live && gl == 0) {
            if (dup_sum) atomicAdd(&batches[b].num_duplicate_faults, dup_sum);
            if (ip_sum) atomicAdd(&batches[b].num_invalid_prefetch_faults, ip_sum);
        }

        // ---- uvm_va_block_service_locked ----
        const bool has = page_faults != 0;                                // the same on every lane of a group
        uint32_t hint = UVMTIER_ID_INVALID, n_pf = 0, n0 = 0, n1 = 0, nt = 0, run_off = 0, run_cnt = 0;
        uint32_t rf = 0, ro = 0;
        if (__any_sync(kFullWarp, has)) {
            uint32_t pf[W];
            uint32_t occupied = 0;
#pragma unroll
            for (int w = 0; w < W; w++) occupied |= gres[w] | c0[w] | c1[w];
            occupied = seg_or<G>(occupied);
            if (prm.pf_enable)
                hint = pf_get_hint_grp<G>(has, newres, gres, occupied == 0u, prm.gpu_id, pref, prm.pf_threshold, prm.pf_min_faults,
                                          &pi, pf, &n_pf);
            if (hint != UVMTIER_ID_INVALID) {                             // uvm_va_block_get_prefetch_hint :11856-11876
#pragma unroll
                for (int w = 0; w < W; w++) newres[w] |= pf[w];
            }
            else {
                n_pf = 0;
#pragma unroll
                for (int w = 0; w < W; w++) pf[w] = 0;
            }
            // region = span of new_residency (with or without the hint it is first..last set bit)
            uint32_t lf = 512u, lo = 0u;
#pragma unroll
            for (int w = W - 1; w >= 0; w--)
                if (newres[w]) lf = 32u * (word0 + w) + (uint32_t)__ffs((int)newres[w]) - 1u;
#pragma unroll
            for (int w = 0; w < W; w++)
                if (newres[w]) lo = 32u * (word0 + w) + 32u - (uint32_t)__clz((int)newres[w]);
            rf = seg_min<G>(lf);
            ro = seg_max<G>(lo);
            if (!has)
                rf = 0;

            // block_copy_resident_pages :4660-4847: per page the source (CPU node 0, node 1, nowhere = zero fill);
            // a copy run = maximal sequence of consecutive pages with the same source and the same cause
            // (fault / prefetch), uvm_va_block.c:4242-4335.  start[] marks the first page of every run.
            uint32_t todo[W], from0[W], from1[W], start[W];
            uint32_t npk = 0, rpk = 0;                                    // three 10-bit fields: node 0, node 1, zero fill
#pragma unroll
            for (int w = 0; w < W; w++) {
                todo[w] = newres[w] & ~gres[w];
                from0[w] = todo[w] & c0[w];
                from1[w] = todo[w] & ~from0[w] & c1[w];
            }
            uint32_t carry = (todo[W - 1] >> 31) | ((from0[W - 1] >> 31) << 1) | ((from1[W - 1] >> 31) << 2) | ((pf[W - 1] >> 31) << 3);
            if (G > 1)
                carry = __shfl_up_sync(kFullWarp, carry, 1, G);
            if (gl == 0)
                carry = 0;
            uint32_t ct = carry & 1u, cf0 = (carry >> 1) & 1u, cf1 = (carry >> 2) & 1u, cp = (carry >> 3) & 1u;
#pragma unroll
            for (int w = 0; w < W; w++) {
                const uint32_t tp = (todo[w] << 1) | ct, f0p = (from0[w] << 1) | cf0, f1p = (from1[w] << 1) | cf1, pp = (pf[w] << 1) | cp;
                start[w] = todo[w] & (~tp | (from0[w] ^ f0p) | (from1[w] ^ f1p) | (pf[w] ^ pp));
                ct = todo[w] >> 31; cf0 = from0[w] >> 31; cf1 = from1[w] >> 31; cp = pf[w] >> 31;
                const uint32_t touch = todo[w] & ~from0[w] & ~from1[w];
                npk += (uint32_t)__popc(from0[w]) | ((uint32_t)__popc(from1[w]) << 10) | ((uint32_t)__popc(touch) << 20);
                rpk += (uint32_t)__popc(start[w] & from0[w]) | ((uint32_t)__popc(start[w] & from1[w]) << 10) |
                       ((uint32_t)__popc(start[w] & touch) << 20);
            }
            npk = seg_add<G>(npk);
            n0 = npk & 1023u; n1 = (npk >> 10) & 1023u; nt = npk >> 20;
            uint32_t incl = rpk;                                          // inclusive prefix over the lanes of the group
            if constexpr (G > 1) {
#pragma unroll
                for (int d = 1; d < G; d <<= 1) {
                    const uint32_t y = __shfl_up_sync(kFullWarp, incl, d, G);
                    if (gl >= d) incl += y;
                }
            }
            const uint32_t tot = G > 1 ? __shfl_sync(kFullWarp, incl, gb + G - 1) : incl;
            const uint32_t r0 = tot & 1023u, r1 = (tot >> 10) & 1023u, r2 = tot >> 20;
            run_cnt = r0 + r1 + r2;
            // run-array space for the whole warp: the atomic is issued here, its answer is picked up after the state
            // update and the stop positions below (its round trip to L2 was 13 % of the kernel's stall samples)
            const bool any_runs = __any_sync(kFullWarp, run_cnt != 0u);
            uint32_t claim_pending = 0, claim_before = 0;
            if (any_runs) {
                claim_pending = svc_claim_runs_issue(&counters[0], gl == 0 ? run_cnt : 0u, lane, &claim_before);
            }
            // uvm_va_block_make_resident_finish :5042-5084 and the mapping step (:11019-11066, :12063-12080)
#pragma unroll
            for (int w = 0; w < W; w++) {
                // make_resident_copy first unmaps the pages that are not yet resident here (:4919-4926) ...
                p0[w] &= ~todo[w]; p1[w] &= ~todo[w]; p2[w] &= ~todo[w];
                // ... then residency moves and the faulting GPU is mapped
                gres[w] |= newres[w];
                c0[w] &= ~newres[w];
                c1[w] &= ~newres[w];
                p0[w] |= newres[w];
                if (logical == PROT_RWA) {
                    p1[w] |= newres[w];
                    p2[w] |= newres[w];
                }
            }
            if (any_runs) {
                // where a run that is still open at the end of this lane's words stops: first stop position in the
                // following lanes
                uint32_t stop[W], next_abs[W];
                uint32_t mine = 512u;
#pragma unroll
                for (int w = W - 1; w >= 0; w--) {
                    stop[w] = ~todo[w] | start[w];
                    if (stop[w]) mine = 32u * (word0 + w) + (uint32_t)__ffs((int)stop[w]) - 1u;
                }
                uint32_t nxt = 512u;
                if constexpr (G > 1) {
                    nxt = __shfl_down_sync(kFullWarp, mine, 1, G);
                    if (gl == G - 1) nxt = 512u;
#pragma unroll
                    for (int d = 1; d < G; d <<= 1) {
                        const uint32_t y = __shfl_down_sync(kFullWarp, nxt, d, G);
                        if (gl + d <= G - 1) nxt = min(nxt, y);
                    }
                }
#pragma unroll
                for (int w = W - 1; w >= 0; w--) {
                    next_abs[w] = nxt;
                    if (stop[w]) nxt = 32u * (word0 + w) + (uint32_t)__ffs((int)stop[w]) - 1u;
                }
                run_off = svc_claim_runs_collect(claim_pending, claim_before);
                if (G > 1)
                    run_off = __shfl_sync(kFullWarp, run_off, gb);
                const uint32_t ex = incl - rpk;
                uint32_t k0 = run_off + (ex & 1023u), k1 = run_off + r0 + ((ex >> 10) & 1023u), k2 = run_off + r0 + r1 + (ex >> 20);
                // Every lane works on ITS first word that still holds run starts (picked with selects: the words
                // are registers).  Walking the words in a fixed order instead made the warp pass through all W
                // words with the handful of lanes whose runs sit in that word (ncu r1_m: 30 % of the kernel's
                // instructions at 4 active lanes).
                uint32_t words_left = 0;
#pragma unroll
                for (int w = 0; w < W; w++) words_left |= start[w] ? (1u << w) : 0u;
                while (words_left) {
                    const int j = __ffs((int)words_left) - 1;
                    words_left &= words_left - 1u;
                    uint32_t s_bits = start[0], stop_j = stop[0], next_j = next_abs[0], f0 = from0[0], f1 = from1[0], pf_j = pf[0];
#pragma unroll
                    for (int w = 1; w < W; w++)
                        if (j == w) { s_bits = start[w]; stop_j = stop[w]; next_j = next_abs[w]; f0 = from0[w]; f1 = from1[w]; pf_j = pf[w]; }
                    const uint32_t page0 = 32u * (word0 + (uint32_t)j);
                    while (s_bits) {
                        const int s = __ffs((int)s_bits) - 1;
                        s_bits &= s_bits - 1;
                        const uint32_t rest = s == 31 ? 0u : (stop_j >> (s + 1));
                        const uint32_t first_page = page0 + (uint32_t)s;
                        const uint32_t end = rest ? first_page + (uint32_t)__ffs((int)rest) : next_j;
                        const uint32_t is0 = (f0 >> s) & 1u, is1 = (f1 >> s) & 1u;
                        const uint32_t pos = is0 ? k0 : (is1 ? k1 : k2);
                        k0 += is0; k1 += is1; k2 += 1u - is0 - is1;
                        if (pos < prm.run_cap) {
                            // uvmtier_copy_run_t: slot | first_page:16, num_pages:10, src:3, prefetch:1, dst:2 (0 = HBM)
                            const uint32_t src = is0 ? 0u : (is1 ? 1u : (uint32_t)UVMTIER_RUN_SRC_ZERO);
                            ((uint2 *)runs)[pos] = make_uint2(slot, first_page | ((end - first_page) << 16) | (src << 26) |
                                                                        (((pf_j >> s) & 1u) << 29));
                        }
                    }
                }
            }
            if (!run_cnt)
                run_off = 0;
        }
        // ---- record: 32-byte header by the first lane, every lane its words of new_residency ----
        if (live) {
            uint32_t *rec = (uint32_t *)&service[t];
            if (gl == 0) {
                ((uint4 *)rec)[0] = make_uint4(slot, rf | (ro << 16), page_faults | (n_pf << 16), n0 | (n1 << 16));
                ((uint4 *)rec)[1] = make_uint4(nt | (hint << 16), run_off, run_cnt | ((b + prm.batch_base) << 16), t - base);
            }
            st_words<W>(rec + 8 + word0, newres);
        }
        t0 = t1; t1 = t2; t2 = t3;
        g0 = g1; g1 = g2;
        e0 = e1;
        b0 = b1;
    }
    if (owns) {
        st_words<W>(st.gpu_resident + wbase, gres);
        st_words<W>(st.cpu_resident[0] + wbase, c0);
        st_words<W>(st.cpu_resident[1] + wbase, c1);
        st_words<W>(st.gpu_pte[0] + wbase, p0);
        st_words<W>(st.gpu_pte[1] + wbase, p1);
        st_words<W>(st.gpu_pte[2] + wbase, p2);
        if (st.gpu_evicted) {
            // pages that an eviction had moved off the GPU and that came back (block_make_resident_clear_evicted,
            // uvm_va_block.c:4978-4988): evicted pages are never GPU resident, so whatever is resident now came back
            uint32_t ev[W];
            ld_words<W>(st.gpu_evicted + wbase, ev);
#pragma unroll
            for (int w = 0; w < W; w++) ev[w] &= ~gres[w];
            st_words<W>(st.gpu_evicted + wbase, ev);
        }
        if (st.read_duplicated) {
            // break_read_duplication_in_region (:4538-4579): the pages handed to make_resident are the ones that lost
            // their CPU copy; a read-duplicated page of a one-GPU block always has one (:5279-5310)
            uint32_t rd[W];
            ld_words<W>(st.read_duplicated + wbase, rd);
#pragma unroll
            for (int w = 0; w < W; w++) rd[w] &= c0[w] | c1[w];
            st_words<W>(st.read_duplicated + wbase, rd);
        }
        if (gl == 0)
            st.prefetch_info[slot] = pi;
    }
    }   // next work item
}

} // namespace uvmtier_k
