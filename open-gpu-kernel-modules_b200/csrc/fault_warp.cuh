// fault_warp.cuh -- K0/K1 for batches of up to 256 entries: ONE WARP per fault batch, E = 1..8 ring entries per lane.
// Same phases, same results and same reference citations as k_fault_fetch_preprocess in fault_kernels.cuh (one CTA
// per batch, one entry per thread), which stays in use for max_batch_size above 256.
//
// Why a warp: with one entry per thread the batch spans eight warps, and every phase that looks at other entries
// (same-uTLB predecessor, the classification rounds, the representative election, the sort, the run boundaries)
// ends in a CTA barrier -- some thirty of them per batch -- while the sort pays two shuffles and a 64-bit compare
// per key and stage.  Here entry i = r * 32 + lane sits in row r of its lane's registers:
//   * no CTA barrier at all after the instance-pointer table is in shared memory (warps of a CTA work on different
//     batches and never meet again); phases are separated by __syncwarp;
//   * ballots are warp-uniform, so "the latest earlier NEW entry" of the classification rounds comes from eight
//     ballot words every lane already holds instead of a per-warp table in shared memory;
//   * the bitonic network runs on index lane * E + q: 21 of its 36 stages for 256 keys compare registers of the same
//     lane and need no shuffle;
//   * eight independent packets per lane give the decode and the table lookups instruction-level parallelism.
#pragma once

#include "devio.cuh"
#include "fault_decode.cuh"

namespace uvmtier_k {

constexpr int kFaultWarpsPerCta = 4;

// Bytes of dynamic shared memory per warp (one batch of 32 * E slots).
__host__ __device__ inline size_t fault_warp_smem_bytes(uint32_t E, uint32_t utlb_count)
{
    const size_t slots = 32u * E;
    size_t per_entry = 8 + 8 + 8              /* inst, page, sort */
                     + 4 * 5                  /* cntT, cntG, mask, best, t1 */
                     + 2 * 5                  /* jt, head, ctx, num, gstart */
                     + 4;                     /* acc, st, mask8, flags */
    size_t b = slots * per_entry + (((size_t)utlb_count * 2 + 15) & ~(size_t)15);
    return (b + 15) & ~(size_t)15;
}
__host__ __device__ inline size_t fault_warp_cta_smem_bytes(uint32_t E, uint32_t utlb_count, uint32_t tbl_n)
{
    return (((size_t)tbl_n * 16 + 15) & ~(size_t)15) + kFaultWarpsPerCta * fault_warp_smem_bytes(E, utlb_count) + 16;
}

__device__ __forceinline__ uint64_t fw_shfl_xor_u64(uint64_t v, int m)
{
    const uint32_t lo = __shfl_xor_sync(0xffffffffu, (uint32_t)v, m);
    const uint32_t hi = __shfl_xor_sync(0xffffffffu, (uint32_t)(v >> 32), m);
    return ((uint64_t)hi << 32) | lo;
}

// Bitonic sort of 32 * E keys held E per lane; key q of lane l has index l * E + q; ascending by index.
template <int E> __device__ __forceinline__ void fault_warp_sort(uint64_t (&key)[E], int lane)
{
    constexpr int N = 32 * E;
#pragma unroll
    for (int k = 2; k <= N; k <<= 1) {
#pragma unroll
        for (int j = k >> 1; j > 0; j >>= 1) {
            if (j < E) {
#pragma unroll
                for (int q = 0; q < E; q++) {
                    if ((q & j) == 0) {
                        const int p = lane * E + q;
                        const bool asc = (p & k) == 0;
                        const uint64_t a = key[q], b = key[q | j];
                        const bool sw = (a > b) == asc;
                        key[q] = sw ? b : a;
                        key[q | j] = sw ? a : b;
                    }
                }
            }
            else {
#pragma unroll
                for (int q = 0; q < E; q++) {
                    const int p = lane * E + q;
                    const uint64_t other = fw_shfl_xor_u64(key[q], j / E);
                    const bool keep_min = ((p & j) == 0) == ((p & k) == 0);
                    key[q] = (keep_min == (key[q] < other)) ? key[q] : other;
                }
            }
        }
    }
}

template <int E>
__global__ void __launch_bounds__(kFaultWarpsPerCta * 32)
k_fault_fetch_preprocess_warp(fault_kparams prm, uvmtier_fault_packet_t *__restrict__ ring,
                              const uvmtier_instance_ptr_entry_t *__restrict__ tbl, uint32_t *__restrict__ disp,
                              uvmtier_ordered_fault_t *__restrict__ ordered, uvmtier_fault_group_t *__restrict__ groups,
                              uvmtier_fault_batch_t *__restrict__ batches, uvmtier_fault_buffer_entry_t *__restrict__ entries,
                              uint16_t *__restrict__ utlb_pending)
{
    constexpr int BW = 32 * E;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t U = prm.utlb_count;
    const uint32_t lane_lt = (1u << lane) - 1u;

    // ---- carve shared memory: the table rows (shared by the CTA), then this warp's slice ----
    UVMTIER_DYN_SMEM(smem);
    uvmtier_instance_ptr_entry_t *s_rows = (uvmtier_instance_ptr_entry_t *)smem;
    unsigned char *sp = smem + (((size_t)prm.tbl_n * 16 + 15) & ~(size_t)15) + (size_t)warp * fault_warp_smem_bytes(E, U);
    uint64_t *s_inst = (uint64_t *)sp;                        sp += (size_t)BW * 8;
    uint64_t *s_page = (uint64_t *)sp;                        sp += (size_t)BW * 8;
    uint64_t *s_sort = (uint64_t *)sp;                        sp += (size_t)BW * 8;
    uint32_t *s_cntT = (uint32_t *)sp;                        sp += (size_t)BW * 4;
    uint32_t *s_cntG = (uint32_t *)sp;                        sp += (size_t)BW * 4;
    uint32_t *s_mask = (uint32_t *)sp;                        sp += (size_t)BW * 4;
    uint32_t *s_best = (uint32_t *)sp;                        sp += (size_t)BW * 4;
    uint32_t *s_t1 = (uint32_t *)sp;                          sp += (size_t)BW * 4;
    int16_t *s_jt = (int16_t *)sp;                            sp += (size_t)BW * 2;
    int16_t *s_head = (int16_t *)sp;                          sp += (size_t)BW * 2;
    uint16_t *s_ctx = (uint16_t *)sp;                         sp += (size_t)BW * 2;
    uint16_t *s_num = (uint16_t *)sp;                         sp += (size_t)BW * 2;
    uint16_t *s_gstart = (uint16_t *)sp;                      sp += (size_t)BW * 2;
    uint8_t *s_acc = (uint8_t *)sp;                           sp += (size_t)BW;
    uint8_t *s_st = (uint8_t *)sp;                            sp += (size_t)BW;
    uint8_t *s_mask8 = (uint8_t *)sp;                         sp += (size_t)BW;
    uint8_t *s_flags = (uint8_t *)sp;                         sp += (size_t)BW;
    uint16_t *s_tbl = (uint16_t *)sp;

    for (uint32_t k = threadIdx.x; k < prm.tbl_n; k += blockDim.x)
        s_rows[k] = tbl[k];
    __syncthreads();
    const uint32_t n_batches = (prm.n_entries + prm.B - 1) / prm.B;
    const uint32_t batch = blockIdx.x * (blockDim.x >> 5) + (uint32_t)warp;
    if (batch >= n_batches)
        return;                                                           // the whole warp; no CTA barrier follows
    const uint32_t base = batch * prm.B;
    const uint32_t nb = min(prm.B, prm.n_entries - base);

    // ---- phase 1: load + decode (F1), E packets per lane ----
    uint64_t myinst[E], mypage[E];
    uint32_t small[E];                                                    // utlb_id | access type << 16 | flags << 24
    uint32_t w7[E];
    uint32_t invalid_rows = 0;                                            // bit r: row r holds an entry that is not valid yet
    uint32_t first_invalid = nb;
#pragma unroll
    for (int r = 0; r < E; r++) {
        const uint32_t i = (uint32_t)(r * 32 + lane);
        const bool inwin = i < nb;
        fault_fields f;
        f.valid = 0; f.flags = 0; f.utlb_id = 0; f.access_type = 0; f.fault_address = 0; f.inst_address = 0; f.inst_aperture = 0; f.ve_id = 0;
        f.timestamp = 0; f.client_id = 0; f.engine_id = 0; f.fault_type = 0; f.client_type = 0; f.engine_type = 0; f.gpc_id = 0;
        w7[r] = 0;
        if (inwin) {
            // get < ring_entries and base + i < n_entries < ring_entries: at most one wrap (:975)
            const uint64_t at = (uint64_t)prm.get + base + i;
            const uint32_t ring_idx = (uint32_t)(at >= prm.ring_entries ? at - prm.ring_entries : at);
            const uint4 a = ld_stream_v4((const uint4 *)&ring[ring_idx]);
            const uint4 b = ld_stream_v4((const uint4 *)&ring[ring_idx] + 1);
            w7[r] = b.w;
            fd_decode(a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w, prm.utlb_per_gpc, prm.va_bits, &f);
        }
        const uint32_t bad = __ballot_sync(0xffffffffu, inwin && !f.valid);
        if (bad && first_invalid == nb)                                   // BATCH_READY: stop at the first entry not yet valid (:899-903)
            first_invalid = (uint32_t)(r * 32) + (uint32_t)__ffs((int)bad) - 1u;
        invalid_rows |= bad ? 1u << r : 0u;
        myinst[r] = ((uint64_t)f.inst_aperture << 60) | ((f.inst_address >> 12) << 8) | f.ve_id;   // cmp_fault_instance_ptr order
        mypage[r] = f.fault_address >> 12;
        small[r] = f.utlb_id | (f.access_type << 16) | (f.flags << 24);
        if (entries && inwin) {
            // the parity / debug view; num_instances, access_type_mask and the filtered flag follow in phase 4
            uvmtier_fault_buffer_entry_t e;
            e.fault_address = f.fault_address;
            e.timestamp = f.timestamp;
            e.instance_ptr_address = f.inst_address;
            e.client_id = (uint16_t)f.client_id;
            e.mmu_engine_id = (uint16_t)f.engine_id;
            e.utlb_id = (uint16_t)f.utlb_id;
            e.num_instances = 0;
            e.instance_ptr_aperture = (uint8_t)f.inst_aperture;
            e.fault_type = (uint8_t)f.fault_type;
            e.fault_access_type = (uint8_t)f.access_type;
            e.client_type = (uint8_t)f.client_type;
            e.mmu_engine_type = (uint8_t)f.engine_type;
            e.gpc_id = (uint8_t)f.gpc_id;
            e.ve_id = (uint8_t)f.ve_id;
            e.flags = (uint8_t)f.flags;
            e.access_type_mask = 0;
            for (int k = 0; k < 7; k++) e.pad[k] = 0;
            s_flags[i] = 1;                                               // marks "entry written" for phase 4 (overwritten there)
            entries[base + i] = e;
        }
    }
    (void)invalid_rows;
    const uint32_t n = first_invalid;                                     // num_cached_faults
    uint32_t max_utlb = 0;
#pragma unroll
    for (int r = 0; r < E; r++) {
        const uint32_t i = (uint32_t)(r * 32 + lane);
        const bool act = i < n;
        if (act && prm.clear_valid) {
            const uint64_t at = (uint64_t)prm.get + base + i;
            const uint32_t ring_idx = (uint32_t)(at >= prm.ring_entries ? at - prm.ring_entries : at);
            ring[ring_idx].w[7] = w7[r] & 0x7fffffffu;                    // entry_clear_valid (uvm_pascal_fault_buffer.c:304-318)
        }
        s_inst[i] = myinst[r];
        s_page[i] = mypage[r];
        s_acc[i] = (uint8_t)((small[r] >> 16) & 0xffu);
        if (act)
            max_utlb = max(max_utlb, small[r] & 0xffffu);
    }
    max_utlb = __reduce_max_sync(0xffffffffu, max_utlb);
    for (uint32_t k = lane; k < U; k += 32)
        s_tbl[k] = 0xffffu;
    __syncwarp();

    // ---- phase 2: p(i) = latest earlier entry from the same uTLB ----
    int p[E];
#pragma unroll
    for (int r = 0; r < E; r++) {
        const uint32_t i = (uint32_t)(r * 32 + lane);
        const bool act = i < n;
        const uint32_t myutlb = small[r] & 0xffffu;
        const uint32_t mkey = act ? myutlb : (0x10000u + (uint32_t)lane);
        const uint32_t mm = __match_any_sync(0xffffffffu, mkey);
        p[r] = -1;
        if (act) {
            const uint32_t lower = mm & lane_lt;
            if (lower) {
                p[r] = r * 32 + 31 - __clz((int)lower);
            }
            else {
                const uint32_t v = s_tbl[myutlb];
                if (v != 0xffffu)
                    p[r] = (int)v;
            }
        }
        __syncwarp();
        if (act && (mm >> lane) == 1u)                                    // highest lane of its uTLB in this row
            s_tbl[myutlb] = (uint16_t)i;
        __syncwarp();
    }

    // ---- phase 3: classification to the fixed point (F2, :782-820) ----
    uint32_t st[E];
    int jt[E], head[E], gt[E];
#pragma unroll
    for (int r = 0; r < E; r++) {
        const uint32_t i = (uint32_t)(r * 32 + lane);
        st[r] = ST_NEW; jt[r] = p[r]; head[r] = -1; gt[r] = -1;
        s_st[i] = (uint8_t)ST_NEW;
        s_jt[i] = (int16_t)p[r];
        s_head[i] = (int16_t)-1;
    }
    __syncwarp();
    uint32_t bn[E];                                                       // ballots of "is NEW" per row: the same in every lane
    for (;;) {
        int last_new_before[E];                                           // latest NEW entry in the rows before row r
        {
            int run = -1;
#pragma unroll
            for (int r = 0; r < E; r++) {
                const uint32_t i = (uint32_t)(r * 32 + lane);
                bn[r] = __ballot_sync(0xffffffffu, i < n && st[r] == ST_NEW);
                last_new_before[r] = run;
                if (bn[r])
                    run = r * 32 + 31 - __clz((int)bn[r]);
            }
        }
        uint32_t st_new[E];
        int jt_new[E], head_new[E], gt_new[E];
        bool changed = false;
#pragma unroll
        for (int r = 0; r < E; r++) {
            const uint32_t i = (uint32_t)(r * 32 + lane);
            const bool act = i < n;
            const bool fatal = ((small[r] >> 24) & UVMTIER_FE_FATAL) != 0;
            const bool mergeable = act && i > 0 && prm.coalesce && !fatal;
            const uint32_t myacc = (small[r] >> 16) & 0xffu;
            st_new[r] = ST_NEW; jt_new[r] = -1; head_new[r] = -1; gt_new[r] = -1;
            if (act) {
                if (p[r] >= 0)
                    jt_new[r] = (s_st[p[r]] != ST_G) ? p[r] : (int)s_jt[p[r]];    // latest earlier same-uTLB entry that is not G-merged
                if (mergeable) {
                    const uint32_t lower_new = bn[r] & lane_lt;          // batch_context->last_fault: latest earlier NEW
                    const int g = lower_new ? r * 32 + 31 - __clz((int)lower_new) : last_new_before[r];
                    const int j = jt_new[r];
                    const bool is_t = j >= 0 && s_inst[j] == myinst[r] && s_page[j] == mypage[r];
                    const bool is_g = !is_t && g >= 0 && s_inst[g] == myinst[r] && s_page[g] == mypage[r] && myacc <= s_acc[g];
                    if (is_t) {
                        st_new[r] = ST_T;
                        head_new[r] = (s_st[j] == ST_NEW) ? j : (int)s_head[j];
                    }
                    else if (is_g) {
                        st_new[r] = ST_G;
                        gt_new[r] = g;
                    }
                }
            }
            changed |= st_new[r] != st[r] || jt_new[r] != jt[r] || head_new[r] != head[r] || gt_new[r] != gt[r];
        }
        __syncwarp();                                                     // everybody has read the previous round
#pragma unroll
        for (int r = 0; r < E; r++) {
            const uint32_t i = (uint32_t)(r * 32 + lane);
            st[r] = st_new[r]; jt[r] = jt_new[r]; head[r] = head_new[r]; gt[r] = gt_new[r];
            s_st[i] = (uint8_t)st[r];
            s_jt[i] = (int16_t)jt[r];
            s_head[i] = (int16_t)head[r];
        }
        __syncwarp();
        if (!__any_sync(0xffffffffu, changed))
            break;
    }
    // bn[] of the last round is final (nothing changed in it)

    // ---- phase 4: representatives, num_instances, access_type_mask (:753-780) ----
    // Per segment (a NEW entry h and the entries T-merged behind it): the representative is the first entry that
    // reaches the segment's highest access type; it inherits every T-merge of the segment and the G-merges that hit h
    // before h was first displaced (index t1) -- later ones land on the already filtered h and are lost, exactly as in
    // the reference (:804-817).
#pragma unroll
    for (int r = 0; r < E; r++) {
        const uint32_t i = (uint32_t)(r * 32 + lane);
        const uint32_t myacc = (small[r] >> 16) & 0xffu;
        s_cntT[i] = 0;
        s_cntG[i] = 0;
        s_mask[i] = 0;
        s_t1[i] = 0xffffu;
        s_best[i] = (i < n && st[r] == ST_NEW) ? ((myacc << 16) | (0xffffu - i)) : 0u;
    }
    __syncwarp();
#pragma unroll
    for (int r = 0; r < E; r++) {
        const uint32_t i = (uint32_t)(r * 32 + lane);
        const uint32_t myacc = (small[r] >> 16) & 0xffu;
        if (i < n && st[r] == ST_T) {
            const int h = head[r];
            atomicAdd(&s_cntT[h], 1u);
            atomicOr(&s_mask[h], 1u << myacc);
            atomicMax(&s_best[h], (myacc << 16) | (0xffffu - i));
            if (myacc > s_acc[h])
                atomicMin(&s_t1[h], i);
        }
    }
    __syncwarp();
#pragma unroll
    for (int r = 0; r < E; r++) {
        const uint32_t i = (uint32_t)(r * 32 + lane);
        const uint32_t myacc = (small[r] >> 16) & 0xffu;
        if (i < n && st[r] == ST_G && i < s_t1[gt[r]]) {
            atomicAdd(&s_cntG[gt[r]], 1u);
            atomicOr(&s_mask[gt[r]], 1u << myacc);
        }
    }
    __syncwarp();
    uint32_t live_rows[E];                                                // live = fetched and not filtered
    int merged_into[E];
    bool breaks_single = false;
    uint32_t num_coalesced = 0;
    {
        int run = -1;                                                     // latest NEW entry before the current row
#pragma unroll
        for (int r = 0; r < E; r++) {
            const uint32_t i = (uint32_t)(r * 32 + lane);
            const bool act = i < n;
            bool filtered = true;
            uint32_t num = 0, amask = 0;
            merged_into[r] = (int)i;
            if (act) {
                const int h = st[r] == ST_NEW ? (int)i : (st[r] == ST_T ? head[r] : -1);
                if (h >= 0) {
                    const int rep = (int)(0xffffu - (s_best[h] & 0xffffu));
                    filtered = rep != (int)i;
                    merged_into[r] = rep;
                    if (!filtered) {
                        num = 1u + s_cntT[h] + s_cntG[h];
                        amask = (1u << s_acc[h]) | s_mask[h];
                    }
                }
                else {
                    merged_into[r] = gt[r];
                }
            }
            live_rows[r] = (act && !filtered) ? 1u : 0u;
            s_num[i] = (uint16_t)num;
            s_mask8[i] = (uint8_t)amask;
            const uint32_t fl = ((small[r] >> 24) & 0xffu) | (filtered ? UVMTIER_FE_FILTERED : 0u);
            const bool entry_written = entries && i < nb;
            s_flags[i] = (uint8_t)fl;
            if (entry_written && act) {
                uvmtier_fault_buffer_entry_t *e = &entries[base + i];
                e->num_instances = (uint16_t)num;
                e->flags = (uint8_t)fl;
                e->access_type_mask = (uint8_t)amask;
            }
            // is_single_instance_ptr (:959-960): every new representative has the instance pointer of the one before it
            if (act && st[r] == ST_NEW && i > 0) {
                const uint32_t lower_new = bn[r] & lane_lt;
                const int g = lower_new ? r * 32 + 31 - __clz((int)lower_new) : run;
                breaks_single |= g >= 0 && s_inst[g] != myinst[r];
            }
            if (bn[r])
                run = r * 32 + 31 - __clz((int)bn[r]);
            num_coalesced += (uint32_t)__popc(bn[r]);
        }
    }
    const bool not_single = __any_sync(0xffffffffu, breaks_single);
    // utlbs[].num_pending_faults (:881, :966): NEW entries per uTLB
    if (utlb_pending) {
        uint16_t *up = utlb_pending + (size_t)batch * U;
        __syncwarp();
        for (uint32_t k = lane; k < U; k += 32)
            s_tbl[k] = 0;
        __syncwarp();
#pragma unroll
        for (int r = 0; r < E; r++) {
            const uint32_t i = (uint32_t)(r * 32 + lane);
            if (i < n && st[r] == ST_NEW) {
                // one shared-memory atomic per distinct uTLB of the row
                const uint32_t myutlb = small[r] & 0xffffu;
                atomicAdd((uint32_t *)(s_tbl + (myutlb & ~1u)), (myutlb & 1u) ? 0x10000u : 1u);
            }
        }
        __syncwarp();
        for (uint32_t k = lane; k < U; k += 32)
            up[k] = s_tbl[k];
    }

    // ---- phase 5: instance pointer -> (va_space, gpu) and the packed sort key (F3) ----
    const uint32_t ib = prm.idx_bits;
    bool xl_fail = false;
    uint64_t key[E];
    uint32_t rank[E];
#pragma unroll
    for (int r = 0; r < E; r++) {
        const uint32_t i = (uint32_t)(r * 32 + lane);
        int ctx = -1;
        rank[r] = 0;
        if (live_rows[r]) {
            const uint64_t want = myinst[r] & ~0xffull;
            const uint32_t ve = (uint32_t)(myinst[r] & 0xffu);
            for (uint32_t k = 0; k < prm.tbl_n; k++) {
                const uvmtier_instance_ptr_entry_t row = s_rows[k];
                const uint64_t have = ((uint64_t)row.aperture << 60) | ((row.instance_ptr >> 12) << 8);
                if (have == want && (row.instance_ptr & 0xfffull) == 0 && (row.ve_id == UVMTIER_VE_ID_ANY || row.ve_id == ve)) {
                    ctx = (int)k;
                    rank[r] = row.sort_rank;
                    break;
                }
            }
            xl_fail |= ctx < 0;
        }
        s_ctx[i] = (uint16_t)ctx;
    }
    const bool xlat_fail = __any_sync(0xffffffffu, xl_fail);
    // key = rank | page | (7 - access_type) | fault_cache index  (ascending = :1000-1020 with the canonical tie-break by
    // index; access type descending, :739)
#pragma unroll
    for (int r = 0; r < E; r++) {
        const uint32_t i = (uint32_t)(r * 32 + lane);
        const uint32_t myacc = (small[r] >> 16) & 0xffu;
        uint64_t kv = ~0ull;
        if (live_rows[r] && !xlat_fail)
            kv = ((uint64_t)rank[r] << (ib + 3 + kPageBits)) | ((mypage[r] & ((1ull << kPageBits) - 1)) << (ib + 3)) |
                 ((uint64_t)(7u - myacc) << ib) | (uint64_t)i;
        s_sort[i] = kv;
    }
    __syncwarp();
    // the network runs on index lane * E + q: lane l takes E consecutive keys
#pragma unroll
    for (int q = 0; q < E; q++) key[q] = s_sort[lane * E + q];
    __syncwarp();
    fault_warp_sort<E>(key, lane);
#pragma unroll
    for (int q = 0; q < E; q++) s_sort[lane * E + q] = key[q];
    __syncwarp();

    // ---- phase 6: outputs + va_block runs ----
    const uint32_t m = xlat_fail ? 0u : num_coalesced;
    const uint32_t idxmask = (1u << ib) - 1u;
    uint32_t ngroups = 0;
    uint32_t bnd[E];                                                      // boundary ballots per row
#pragma unroll
    for (int r = 0; r < E; r++) {
        const uint32_t pos = (uint32_t)(r * 32 + lane);
        const bool have = pos < m;
        bool boundary = false;
        if (have) {
            const uint64_t kv = s_sort[pos];
            const uint32_t idx = (uint32_t)kv & idxmask;
            const uint32_t sacc = s_acc[idx];
            const uint64_t saddr = s_page[idx] << 12;
            const bool sfatal = (s_flags[idx] & UVMTIER_FE_FATAL) != 0;
            uvmtier_ordered_fault_t o;
            o.fault_address = saddr;
            o.fault_index = (uint16_t)idx;
            o.ctx = s_ctx[idx];
            o.num_instances = s_num[idx];
            o.access_type_mask = s_mask8[idx];
            o.fault_access_type = (uint8_t)(sacc | (sfatal ? UVMTIER_OF_FATAL : 0u));
            ordered[base + pos] = o;
            // a run breaks when (va_space/gpu rank, va_block) changes, and around fatal entries (:2294, :1433-1438)
            if (pos == 0) {
                boundary = true;
            }
            else {
                const uint64_t prev = s_sort[pos - 1];
                const uint32_t pidx = (uint32_t)prev & idxmask;
                const bool pfatal = (s_flags[pidx] & UVMTIER_FE_FATAL) != 0;
                boundary = (prev >> (ib + 3 + 9)) != (kv >> (ib + 3 + 9)) || sfatal || pfatal;
            }
        }
        bnd[r] = __ballot_sync(0xffffffffu, boundary);
        if (boundary)
            s_gstart[ngroups + (uint32_t)__popc(bnd[r] & lane_lt)] = (uint16_t)pos;
        ngroups += (uint32_t)__popc(bnd[r]);
    }
    __syncwarp();
    {
        uint32_t before = 0;
#pragma unroll
        for (int r = 0; r < E; r++) {
            const uint32_t pos = (uint32_t)(r * 32 + lane);
            if ((bnd[r] >> lane) & 1u) {
                const uint32_t gidx = before + (uint32_t)__popc(bnd[r] & lane_lt);
                const uint32_t next = gidx + 1 < ngroups ? s_gstart[gidx + 1] : m;
                const uint32_t idx = (uint32_t)s_sort[pos] & idxmask;
                uvmtier_fault_group_t g;
                g.block_start = (s_page[idx] << 12) & ~((uint64_t)UVMTIER_VA_BLOCK_SIZE - 1);
                g.ctx = s_ctx[idx];
                g.first_fault_index = (uint16_t)pos;
                g.block_faults = (uint16_t)(next - pos);
                g.flags = (s_flags[idx] & UVMTIER_FE_FATAL) ? 1u : 0u;
                groups[base + gidx] = g;
            }
            before += (uint32_t)__popc(bnd[r]);
        }
    }
#pragma unroll
    for (int r = 0; r < E; r++) {
        const uint32_t i = (uint32_t)(r * 32 + lane);
        if (i < nb) {
            uint32_t d = 0;
            if (i < n) {
                const bool fatal = ((small[r] >> 24) & UVMTIER_FE_FATAL) != 0;
                d = ((uint32_t)merged_into[r] & UVMTIER_DISP_INDEX_MASK) | (live_rows[r] ? 0u : UVMTIER_DISP_FILTERED) |
                    (fatal ? UVMTIER_DISP_FATAL : 0u) | UVMTIER_DISP_VALID | (st[r] << UVMTIER_DISP_RULE_SHIFT);
            }
            disp[base + i] = d;
        }
    }
    if (lane == 0) {
        uvmtier_fault_batch_t bt;
        bt.num_cached_faults = n;
        bt.num_coalesced_faults = num_coalesced;
        bt.is_single_instance_ptr = not_single ? 0u : 1u;
        bt.max_utlb_id = max_utlb;
        bt.status = xlat_fail ? NV_WARN_MORE_PROCESSING_REQUIRED : NV_OK;
        bt.num_duplicate_faults = 0;
        bt.num_invalid_prefetch_faults = 0;
        bt.num_block_groups = ngroups;
        batches[batch] = bt;
    }
}

} // namespace uvmtier_k
