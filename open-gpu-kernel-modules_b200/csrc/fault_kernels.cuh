// fault_kernels.cuh -- K0/K1: fault packet decode, fetch-time coalescing, sort and va_block grouping,
// one CTA per fault batch, one thread per ring entry.  Device code only (host side: faults.cu); kept
// in a header so tests/simt can run the same source on the CPU emulator.
//
// Reference (UVM = kernel-open/nvidia-uvm): fetch_fault_buffer_entries (uvm_gpu_replayable_faults.c:844-986),
// fetch_fault_buffer_try_merge_entry/_merge_entry (:753-820), preprocess_fault_batch (:1134-1178), comparators
// (:695-740, :990-1020), and the block-run scan of service_fault_batch / _block_locked (:2252-2349, :1433-1438).
//
// The reference's coalescing is a sequential scan with per-uTLB and batch-wide "last fault" pointers.
// Here every entry classifies itself as NEW / merged-by-uTLB-rule (T) / merged-by-batch-rule (G) from
// the classification of the entries before it, and the CTA iterates that to the fixed point.  The
// fixed point is unique and equals the sequential result (entry i is right as soon as all j < i are),
// and is reached in a handful of rounds on real traces (worst case: batch-size rounds).
#pragma once

#include "devio.cuh"
#include "fault_decode.cuh"

namespace uvmtier_k {

constexpr uint32_t kFaultMaxThreads = 1024;
constexpr uint32_t kFaultMaxTableRows = 256;
constexpr uint32_t kPageBits = 45;          // page-number bits kept in the sort key (57-bit VA)

enum : uint32_t { ST_NEW = 0, ST_T = 1, ST_G = 2 };

struct fault_kparams {
    uint32_t B;              // max_batch_size
    uint32_t utlb_per_gpc;
    uint32_t utlb_count;
    uint32_t va_bits;
    uint32_t coalesce;
    uint32_t ring_entries;
    uint32_t get;
    uint32_t n_entries;      // (put - get) mod ring_entries
    uint32_t tbl_n;
    uint32_t idx_bits;       // log2(blockDim.x)
    uint32_t clear_valid;
};

struct fault_smem_scalars {
    uint32_t first_invalid;
    uint32_t max_utlb;
    int warp_last_new[32];
    uint32_t warp_count[32];
};

// Bytes of dynamic shared memory the kernel needs for T threads.
__host__ __device__ inline size_t fault_smem_bytes(uint32_t T, uint32_t utlb_count, uint32_t tbl_n)
{
    size_t per_entry = 8 + 8 + 8 + 8      /* inst, page, sort ping-pong */
                     + 4 * 5              /* cntT, cntG, mask, best, t1 */
                     + 2 * 6              /* p, jt, head, ctx, num, gstart */
                     + 4;                 /* acc, st, mask8, flags */
    size_t tbl = (size_t)(T / 32) * utlb_count * 2;
    tbl = (tbl + 15) & ~(size_t)15;
    return sizeof(fault_smem_scalars) + (size_t)T * per_entry + tbl + (size_t)tbl_n * 16 + 64;
}

__device__ __forceinline__ uint64_t shfl_xor_u64(uint64_t v, int m)
{
    uint32_t lo = __shfl_xor_sync(0xffffffffu, (uint32_t)v, m);
    uint32_t hi = __shfl_xor_sync(0xffffffffu, (uint32_t)(v >> 32), m);
    return ((uint64_t)hi << 32) | lo;
}

// Bitonic sort of T keys, one per thread of the CTA (ascending by thread id).  s0 / s1: T keys each.
template <int T>
__device__ __forceinline__ uint64_t fault_sort_keys(uint64_t key, int tid, uint64_t *s0, uint64_t *s1)
{
    int buf = 0;
#pragma unroll
    for (int k = 2; k <= T; k <<= 1) {
#pragma unroll
        for (int j = k >> 1; j > 0; j >>= 1) {
            uint64_t other;
            if (j < 32) {
                other = shfl_xor_u64(key, j);
            }
            else {
                uint64_t *s = buf ? s1 : s0;
                s[tid] = key;
                __syncthreads();
                other = s[tid ^ j];
                buf ^= 1;
            }
            const bool keep_min = ((tid & j) == 0) == ((tid & k) == 0);
            key = (keep_min == (key < other)) ? key : other;
        }
    }
    return key;
}

// ------------------------------------------------------------------------------------------------
// F1 only: one thread per packet.
__global__ void __launch_bounds__(256)
k_fault_parse(const uvmtier_fault_packet_t *__restrict__ packets, size_t n, uint32_t utlb_per_gpc, uint32_t va_bits,
              uvmtier_fault_buffer_entry_t *__restrict__ out)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const uint4 a = ld_stream_v4((const uint4 *)&packets[i]);
        const uint4 b = ld_stream_v4((const uint4 *)&packets[i] + 1);
        fault_fields f;
        fd_decode(a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w, utlb_per_gpc, va_bits, &f);
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
        out[i] = e;
    }
}

// ------------------------------------------------------------------------------------------------
// Non-replayable faults, front end (UVM/uvm_gpu_non_replayable_faults.c): fetch_non_replayable_fault_buffer_entries
// (:179-236) parses the shadow-buffer copy -- the same 32-byte packet, no valid bit to clear, one instance per entry, no
// coalescing -- and service_fault_once (:606-712) looks every entry's channel up and services the entries one after
// the other, in buffer order.  Here: one thread per packet; the entry becomes a va_block run of its own, in the layout
// the servicing pass takes (run g of "batch" b is entry b * B + g, so the pass applies the runs of a va_block in
// buffer order).  A fatal fault type or an instance pointer that does not translate (:623-632: the channel is gone,
// the entry is dropped) makes a run the pass skips (flags bit 0).
__global__ void __launch_bounds__(256)
k_nonreplayable_prepare(const uvmtier_fault_packet_t *__restrict__ packets, uint32_t n, uint32_t B, uint32_t utlb_per_gpc,
                        uint32_t va_bits, const uvmtier_instance_ptr_entry_t *__restrict__ tbl, uint32_t tbl_n,
                        uvmtier_ordered_fault_t *__restrict__ ordered, uvmtier_fault_group_t *__restrict__ groups,
                        uvmtier_fault_batch_t *__restrict__ batches, uint16_t *__restrict__ entry_flags,
                        uvmtier_fault_buffer_entry_t *__restrict__ entries)
{
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint4 a = ld_stream_v4((const uint4 *)&packets[i]);
        const uint4 b = ld_stream_v4((const uint4 *)&packets[i] + 1);
        fault_fields f;
        fd_decode(a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w, utlb_per_gpc, va_bits, &f);
        const bool fatal = (f.flags & UVMTIER_FE_FATAL) != 0;
        int ctx = -1;
        for (uint32_t k = 0; k < tbl_n; k++) {
            const uvmtier_instance_ptr_entry_t r = tbl[k];
            if (r.instance_ptr == f.inst_address && r.aperture == f.inst_aperture && (r.ve_id == UVMTIER_VE_ID_ANY || r.ve_id == f.ve_id)) {
                ctx = (int)k;
                break;
            }
        }
        uvmtier_ordered_fault_t o;
        o.fault_address = f.fault_address;
        o.fault_index = (uint16_t)(i % B);
        o.ctx = (uint16_t)(ctx < 0 ? 0 : ctx);
        o.num_instances = 1;                                                        // :215
        o.access_type_mask = (uint8_t)(1u << f.access_type);                        // :216
        o.fault_access_type = (uint8_t)(f.access_type | (fatal ? UVMTIER_OF_FATAL : 0u));
        ordered[i] = o;
        uvmtier_fault_group_t g;
        g.block_start = f.fault_address & ~((uint64_t)UVMTIER_VA_BLOCK_SIZE - 1);
        g.ctx = o.ctx;
        g.first_fault_index = o.fault_index;
        g.block_faults = 1;
        g.flags = (fatal || ctx < 0) ? 1u : 0u;
        groups[i] = g;
        // a fatal fault type is reported as such (fatal_reason = UvmEventFatalReasonInvalidFaultType, :221-224) -- if its
        // channel still exists; a dropped entry reports nothing
        if (fatal && ctx >= 0)
            entry_flags[i] = (uint16_t)(UVMTIER_SF_FATAL | (3u << UVMTIER_SF_REASON_SHIFT));
        if (i % B == 0) {
            uvmtier_fault_batch_t bt;
            const uint32_t left = n - i;
            bt.num_cached_faults = bt.num_coalesced_faults = left < B ? left : B;
            bt.is_single_instance_ptr = 0;
            bt.max_utlb_id = 0;
            bt.status = NV_OK;
            bt.num_duplicate_faults = 0;
            bt.num_invalid_prefetch_faults = 0;
            bt.num_block_groups = bt.num_cached_faults;
            batches[i / B] = bt;
        }
        if (entries) {
            uvmtier_fault_buffer_entry_t e;
            e.fault_address = f.fault_address;
            e.timestamp = f.timestamp;
            e.instance_ptr_address = f.inst_address;
            e.client_id = (uint16_t)f.client_id;
            e.mmu_engine_id = (uint16_t)f.engine_id;
            e.utlb_id = (uint16_t)f.utlb_id;
            e.num_instances = 1;
            e.instance_ptr_aperture = (uint8_t)f.inst_aperture;
            e.fault_type = (uint8_t)f.fault_type;
            e.fault_access_type = (uint8_t)f.access_type;
            e.client_type = (uint8_t)f.client_type;
            e.mmu_engine_type = (uint8_t)f.engine_type;
            e.gpc_id = (uint8_t)f.gpc_id;
            e.ve_id = (uint8_t)f.ve_id;
            e.flags = (uint8_t)f.flags;
            e.access_type_mask = (uint8_t)(1u << f.access_type);
            for (int k = 0; k < 7; k++) e.pad[k] = 0;
            entries[i] = e;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// One CTA per batch.  blockDim.x = T = power of two >= B, 32 <= T <= 1024.
__global__ void __launch_bounds__(kFaultMaxThreads)
k_fault_fetch_preprocess(fault_kparams prm, uvmtier_fault_packet_t *__restrict__ ring,
                         const uvmtier_instance_ptr_entry_t *__restrict__ tbl, uint32_t *__restrict__ disp,
                         uvmtier_ordered_fault_t *__restrict__ ordered, uvmtier_fault_group_t *__restrict__ groups,
                         uvmtier_fault_batch_t *__restrict__ batches, uvmtier_fault_buffer_entry_t *__restrict__ entries,
                         uint16_t *__restrict__ utlb_pending)
{
    const int T = (int)blockDim.x, tid = (int)threadIdx.x, lane = tid & 31, warp = tid >> 5, W = T >> 5;
    const uint32_t U = prm.utlb_count;
    const uint32_t base = blockIdx.x * prm.B;
    const uint32_t nb = min(prm.B, prm.n_entries - base);
    const uint32_t lane_lt = (1u << lane) - 1u;

    // ---- carve shared memory ----
    UVMTIER_DYN_SMEM(smem);
    unsigned char *sp = smem;
    fault_smem_scalars *sc = (fault_smem_scalars *)sp;        sp += (sizeof(fault_smem_scalars) + 15) & ~15u;
    uint64_t *s_inst = (uint64_t *)sp;                        sp += (size_t)T * 8;
    uint64_t *s_page = (uint64_t *)sp;                        sp += (size_t)T * 8;
    uint64_t *s_sort0 = (uint64_t *)sp;                       sp += (size_t)T * 8;
    uint64_t *s_sort1 = (uint64_t *)sp;                       sp += (size_t)T * 8;
    uint32_t *s_cntT = (uint32_t *)sp;                        sp += (size_t)T * 4;
    uint32_t *s_cntG = (uint32_t *)sp;                        sp += (size_t)T * 4;
    uint32_t *s_mask = (uint32_t *)sp;                        sp += (size_t)T * 4;
    uint32_t *s_best = (uint32_t *)sp;                        sp += (size_t)T * 4;
    uint32_t *s_t1 = (uint32_t *)sp;                          sp += (size_t)T * 4;
    int16_t *s_jt = (int16_t *)sp;                            sp += (size_t)T * 2;
    int16_t *s_head = (int16_t *)sp;                          sp += (size_t)T * 2;
    uint16_t *s_ctx = (uint16_t *)sp;                         sp += (size_t)T * 2;
    uint16_t *s_num = (uint16_t *)sp;                         sp += (size_t)T * 2;
    uint16_t *s_gstart = (uint16_t *)sp;                      sp += (size_t)T * 2;
    sp += (size_t)T * 2;                                      // (spare, keeps fault_smem_bytes' count)
    uint8_t *s_acc = (uint8_t *)sp;                           sp += (size_t)T;
    uint8_t *s_st = (uint8_t *)sp;                            sp += (size_t)T;
    uint8_t *s_mask8 = (uint8_t *)sp;                         sp += (size_t)T;
    uint8_t *s_flags = (uint8_t *)sp;                         sp += (size_t)T;
    sp = (unsigned char *)(((uintptr_t)sp + 15) & ~(uintptr_t)15);
    uint16_t *s_tbl = (uint16_t *)sp;                         sp += (((size_t)W * U * 2) + 15) & ~(size_t)15;
    uvmtier_instance_ptr_entry_t *s_rows = (uvmtier_instance_ptr_entry_t *)sp;

    // ---- phase 1: load + decode (F1) ----
    const bool inwin = (uint32_t)tid < nb;
    uint32_t ring_idx = 0;
    fault_fields f;
    f.valid = 0; f.flags = 0; f.utlb_id = 0; f.access_type = 0; f.fault_address = 0; f.inst_address = 0; f.inst_aperture = 0; f.ve_id = 0;
    uint32_t w7 = 0;
    if (inwin) {
        // get < ring_entries and base + tid < n_entries < ring_entries: at most one wrap (:975)
        const uint64_t at = (uint64_t)prm.get + base + (uint32_t)tid;
        ring_idx = (uint32_t)(at >= prm.ring_entries ? at - prm.ring_entries : at);
        const uint4 a = ld_stream_v4((const uint4 *)&ring[ring_idx]);
        const uint4 b = ld_stream_v4((const uint4 *)&ring[ring_idx] + 1);
        w7 = b.w;
        fd_decode(a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w, prm.utlb_per_gpc, prm.va_bits, &f);
    }
    if (tid == 0) {
        sc->first_invalid = nb;
        sc->max_utlb = 0;
    }
    for (uint32_t k = tid; k < (uint32_t)W * U; k += T)
        s_tbl[k] = 0xffffu;
    for (uint32_t k = tid; k < prm.tbl_n; k += T)
        s_rows[k] = tbl[k];
    __syncthreads();
    if (inwin && !f.valid)
        atomicMin(&sc->first_invalid, (uint32_t)tid);      // BATCH_READY: stop at the first entry not yet valid (:899-903)
    __syncthreads();
    const uint32_t n = sc->first_invalid;                   // num_cached_faults
    const bool act = (uint32_t)tid < n;
    if (act && prm.clear_valid)
        ring[ring_idx].w[7] = w7 & 0x7fffffffu;             // entry_clear_valid (uvm_pascal_fault_buffer.c:304-318)

    const uint64_t myinst = ((uint64_t)f.inst_aperture << 60) | ((f.inst_address >> 12) << 8) | f.ve_id;   // cmp_fault_instance_ptr order
    const uint64_t mypage = f.fault_address >> 12;
    const uint32_t myacc = f.access_type;
    const uint32_t myutlb = f.utlb_id;
    const bool fatal = (f.flags & UVMTIER_FE_FATAL) != 0;
    s_inst[tid] = myinst;
    s_page[tid] = mypage;
    s_acc[tid] = (uint8_t)myacc;
    if (act)
        atomicMax(&sc->max_utlb, myutlb);

    // ---- phase 2: p(i) = latest earlier entry from the same uTLB ----
    const uint32_t mkey = act ? myutlb : (0x10000u + (uint32_t)lane);
    const uint32_t mm = __match_any_sync(0xffffffffu, mkey);
    int p = -1;
    if (act) {
        const uint32_t lower = mm & lane_lt;
        if (lower)
            p = (tid & ~31) + 31 - __clz((int)lower);
        if ((mm >> lane) == 1u)                              // highest lane of its uTLB in this warp
            s_tbl[(uint32_t)warp * U + myutlb] = (uint16_t)tid;
    }
    __syncthreads();
    if (act && p < 0) {
        for (int ww = warp - 1; ww >= 0; --ww) {
            const uint32_t v = s_tbl[(uint32_t)ww * U + myutlb];
            if (v != 0xffffu) { p = (int)v; break; }
        }
    }

    // ---- phase 3: classification to the fixed point (F2, :782-820) ----
    const bool mergeable = act && tid > 0 && prm.coalesce && !fatal;
    uint32_t st = ST_NEW;
    int jt = p, head = -1, gt = -1;
    s_st[tid] = (uint8_t)st;
    s_jt[tid] = (int16_t)jt;
    s_head[tid] = (int16_t)head;
    for (;;) {
        const uint32_t bn = __ballot_sync(0xffffffffu, act && st == ST_NEW);
        if (lane == 0)
            sc->warp_last_new[warp] = bn ? (tid + 31 - __clz((int)bn)) : -1;
        __syncthreads();
        uint32_t st_new = ST_NEW;
        int jt_new = -1, head_new = -1, gt_new = -1;
        if (act) {
            if (p >= 0)
                jt_new = (s_st[p] != ST_G) ? p : (int)s_jt[p];        // latest earlier same-uTLB entry that is not G-merged
            if (mergeable) {
                int g = -1;                                           // batch_context->last_fault: latest earlier NEW
                const uint32_t lower_new = bn & lane_lt;
                if (lower_new)
                    g = (tid & ~31) + 31 - __clz((int)lower_new);
                else
                    for (int ww = warp - 1; ww >= 0; --ww) {
                        const int v = sc->warp_last_new[ww];
                        if (v >= 0) { g = v; break; }
                    }
                const bool is_t = jt_new >= 0 && s_inst[jt_new] == myinst && s_page[jt_new] == mypage;
                const bool is_g = !is_t && g >= 0 && s_inst[g] == myinst && s_page[g] == mypage && myacc <= s_acc[g];
                if (is_t) {
                    st_new = ST_T;
                    head_new = (s_st[jt_new] == ST_NEW) ? jt_new : (int)s_head[jt_new];
                }
                else if (is_g) {
                    st_new = ST_G;
                    gt_new = g;
                }
            }
        }
        const bool changed = st_new != st || jt_new != jt || head_new != head || gt_new != gt;
        __syncthreads();                                              // everybody has read the previous round
        st = st_new; jt = jt_new; head = head_new; gt = gt_new;
        s_st[tid] = (uint8_t)st;
        s_jt[tid] = (int16_t)jt;
        s_head[tid] = (int16_t)head;
        if (!__syncthreads_or(changed))
            break;
    }

    // ---- phase 4: representatives, num_instances, access_type_mask (:753-780) ----
    // Per segment (a NEW entry h and the entries T-merged behind it): the representative is the first
    // entry that reaches the segment's highest access type; it inherits every T-merge of the segment
    // and the G-merges that hit h before h was first displaced (index t1) -- later ones land on the
    // already filtered h and are lost, exactly as in the reference (:804-817).
    s_cntT[tid] = 0;
    s_cntG[tid] = 0;
    s_mask[tid] = 0;
    s_t1[tid] = 0xffffu;
    s_best[tid] = (act && st == ST_NEW) ? ((myacc << 16) | (0xffffu - (uint32_t)tid)) : 0u;
    __syncthreads();
    if (act && st == ST_T) {
        atomicAdd(&s_cntT[head], 1u);
        atomicOr(&s_mask[head], 1u << myacc);
        atomicMax(&s_best[head], (myacc << 16) | (0xffffu - (uint32_t)tid));
        if (myacc > s_acc[head])
            atomicMin(&s_t1[head], (uint32_t)tid);
    }
    __syncthreads();
    if (act && st == ST_G && (uint32_t)tid < s_t1[gt]) {
        atomicAdd(&s_cntG[gt], 1u);
        atomicOr(&s_mask[gt], 1u << myacc);
    }
    __syncthreads();
    bool filtered = true;
    int merged_into = tid;
    uint32_t num = 0, amask = 0;
    if (act) {
        const int h = st == ST_NEW ? tid : (st == ST_T ? head : -1);
        if (h >= 0) {
            const int rep = (int)(0xffffu - (s_best[h] & 0xffffu));
            filtered = rep != tid;
            merged_into = rep;
            if (!filtered) {
                num = 1u + s_cntT[h] + s_cntG[h];
                amask = (1u << s_acc[h]) | s_mask[h];
            }
        }
        else {
            merged_into = gt;
        }
    }
    const bool live = act && !filtered;
    s_num[tid] = (uint16_t)num;
    s_mask8[tid] = (uint8_t)amask;
    s_flags[tid] = (uint8_t)(f.flags | (filtered ? UVMTIER_FE_FILTERED : 0u));

    // is_single_instance_ptr (:959-960): every new representative has the instance pointer of the one before it
    bool breaks_single = false;
    {
        // warp_last_new[] of the last round is final (nothing changed in it)
        const uint32_t bnf = __ballot_sync(0xffffffffu, act && st == ST_NEW);
        if (act && st == ST_NEW && tid > 0) {
            int g = -1;
            const uint32_t lower_new = bnf & lane_lt;
            if (lower_new)
                g = (tid & ~31) + 31 - __clz((int)lower_new);
            else
                for (int ww = warp - 1; ww >= 0; --ww) {
                    const int v = sc->warp_last_new[ww];
                    if (v >= 0) { g = v; break; }
                }
            breaks_single = g >= 0 && s_inst[g] != myinst;
        }
    }
    const int not_single = __syncthreads_or(breaks_single);
    const uint32_t num_coalesced = (uint32_t)__syncthreads_count(act && st == ST_NEW);

    // utlbs[].num_pending_faults (:881, :966): entries that opened a new representative, per uTLB.  The per-warp
    // uTLB table of phase 2 is free again: its first U slots take the counts.
    if (utlb_pending) {
        for (uint32_t k = tid; k < U; k += T)
            s_tbl[k] = 0;
        __syncthreads();
        if (act && st == ST_NEW)
            atomicAdd((uint32_t *)(s_tbl + (myutlb & ~1u)), (myutlb & 1u) ? 0x10000u : 1u);   // two 16-bit counters per word
        __syncthreads();
        for (uint32_t k = tid; k < U; k += T)
            utlb_pending[(size_t)blockIdx.x * U + k] = s_tbl[k];
    }

    // ---- phase 5: instance pointer -> (va_space, gpu) and the packed sort key (F3) ----
    int ctx = -1;
    uint32_t rank = 0;
    if (live) {
        for (uint32_t k = 0; k < prm.tbl_n; k++) {
            const uvmtier_instance_ptr_entry_t r = s_rows[k];
            if (r.instance_ptr == f.inst_address && r.aperture == f.inst_aperture &&
                (r.ve_id == UVMTIER_VE_ID_ANY || r.ve_id == f.ve_id)) {
                ctx = (int)k;
                rank = r.sort_rank;
                break;
            }
        }
    }
    s_ctx[tid] = (uint16_t)ctx;
    const int xlat_fail = __syncthreads_or(live && ctx < 0);

    // key = rank | page | (7 - access_type) | fault_cache index  (ascending = :1000-1020 with the
    // canonical tie-break by index; access type descending, :739)
    const uint32_t ib = prm.idx_bits;
    uint64_t key = ~0ull;
    if (live && !xlat_fail)
        key = ((uint64_t)rank << (ib + 3 + kPageBits)) | ((mypage & ((1ull << kPageBits) - 1)) << (ib + 3)) |
              ((uint64_t)(7u - myacc) << ib) | (uint64_t)tid;

    // bitonic sort of T keys: partner distances < 32 by shuffle, >= 32 through shared memory (ping-pong); the
    // network is unrolled for the CTA size (the generic loop spent more instructions on its control than on keys)
    switch (T) {
        case 32: key = fault_sort_keys<32>(key, tid, s_sort0, s_sort1); break;
        case 64: key = fault_sort_keys<64>(key, tid, s_sort0, s_sort1); break;
        case 128: key = fault_sort_keys<128>(key, tid, s_sort0, s_sort1); break;
        case 256: key = fault_sort_keys<256>(key, tid, s_sort0, s_sort1); break;
        case 512: key = fault_sort_keys<512>(key, tid, s_sort0, s_sort1); break;
        default: key = fault_sort_keys<1024>(key, tid, s_sort0, s_sort1); break;
    }
    // publish the sorted keys for neighbour lookups
    __syncthreads();
    s_sort0[tid] = key;
    __syncthreads();

    // ---- phase 6: outputs + va_block runs ----
    const uint32_t m = xlat_fail ? 0u : num_coalesced;
    const bool have = (uint32_t)tid < m;
    const uint32_t idxmask = (1u << ib) - 1u;
    uint32_t idx = 0, sacc = 0;
    uint64_t saddr = 0;
    bool sfatal = false, boundary = false;
    if (have) {
        idx = (uint32_t)key & idxmask;
        sacc = s_acc[idx];
        saddr = s_page[idx] << 12;
        sfatal = (s_flags[idx] & UVMTIER_FE_FATAL) != 0;
        uvmtier_ordered_fault_t o;
        o.fault_address = saddr;
        o.fault_index = (uint16_t)idx;
        o.ctx = s_ctx[idx];
        o.num_instances = s_num[idx];
        o.access_type_mask = s_mask8[idx];
        o.fault_access_type = (uint8_t)(sacc | (sfatal ? UVMTIER_OF_FATAL : 0u));
        ordered[base + tid] = o;
        // a run breaks when (va_space/gpu rank, va_block) changes, and around fatal entries (:2294, :1433-1438)
        if (tid == 0) {
            boundary = true;
        }
        else {
            const uint64_t prev = s_sort0[tid - 1];
            const uint32_t pidx = (uint32_t)prev & idxmask;
            const bool pfatal = (s_flags[pidx] & UVMTIER_FE_FATAL) != 0;
            boundary = (prev >> (ib + 3 + 9)) != (key >> (ib + 3 + 9)) || sfatal || pfatal;
        }
    }
    const uint32_t bb = __ballot_sync(0xffffffffu, boundary);
    if (lane == 0)
        sc->warp_count[warp] = __popc(bb);
    __syncthreads();
    uint32_t gidx = __popc(bb & lane_lt), ngroups = 0;
    for (int ww = 0; ww < W; ww++) {
        const uint32_t c = sc->warp_count[ww];
        if (ww < warp) gidx += c;
        ngroups += c;
    }
    if (boundary)
        s_gstart[gidx] = (uint16_t)tid;
    __syncthreads();
    if (boundary) {
        const uint32_t next = gidx + 1 < ngroups ? s_gstart[gidx + 1] : m;
        uvmtier_fault_group_t g;
        g.block_start = saddr & ~((uint64_t)UVMTIER_VA_BLOCK_SIZE - 1);
        g.ctx = s_ctx[idx];
        g.first_fault_index = (uint16_t)tid;
        g.block_faults = (uint16_t)(next - (uint32_t)tid);
        g.flags = sfatal ? 1u : 0u;
        groups[base + gidx] = g;
    }

    if (inwin) {
        uint32_t d = 0;
        if (act)
            d = ((uint32_t)merged_into & UVMTIER_DISP_INDEX_MASK) | (filtered ? UVMTIER_DISP_FILTERED : 0u) |
                (fatal ? UVMTIER_DISP_FATAL : 0u) | UVMTIER_DISP_VALID | (st << UVMTIER_DISP_RULE_SHIFT);
        disp[base + tid] = d;
    }
    if (entries && act) {
        uvmtier_fault_buffer_entry_t e;
        e.fault_address = f.fault_address;
        e.timestamp = f.timestamp;
        e.instance_ptr_address = f.inst_address;
        e.client_id = (uint16_t)f.client_id;
        e.mmu_engine_id = (uint16_t)f.engine_id;
        e.utlb_id = (uint16_t)f.utlb_id;
        e.num_instances = (uint16_t)num;
        e.instance_ptr_aperture = (uint8_t)f.inst_aperture;
        e.fault_type = (uint8_t)f.fault_type;
        e.fault_access_type = (uint8_t)f.access_type;
        e.client_type = (uint8_t)f.client_type;
        e.mmu_engine_type = (uint8_t)f.engine_type;
        e.gpc_id = (uint8_t)f.gpc_id;
        e.ve_id = (uint8_t)f.ve_id;
        e.flags = (uint8_t)(f.flags | (filtered ? UVMTIER_FE_FILTERED : 0u));
        e.access_type_mask = (uint8_t)amask;
        for (int k = 0; k < 7; k++) e.pad[k] = 0;
        entries[base + tid] = e;
    }
    if (tid == 0) {
        uvmtier_fault_batch_t bt;
        bt.num_cached_faults = n;
        bt.num_coalesced_faults = num_coalesced;
        bt.is_single_instance_ptr = not_single ? 0u : 1u;
        bt.max_utlb_id = sc->max_utlb;
        bt.status = xlat_fail ? NV_WARN_MORE_PROCESSING_REQUIRED : NV_OK;
        bt.num_duplicate_faults = 0;
        bt.num_invalid_prefetch_faults = 0;
        bt.num_block_groups = ngroups;
        batches[blockIdx.x] = bt;
    }
}

// ------------------------------------------------------------------------------------------------
// Batches of more than kFaultMaxThreads entries (uvm_perf_fault_batch_count up to the size of the fault buffer,
// UVM/uvm_gpu_replayable_faults.c:153-165): one CTA of 1024 threads per batch, several entries per thread, the per-entry
// arrays in global working space (kBigBytesPerEntry bytes per entry, T = power of two >= B entries per batch).
// Decode, translation, sort and run formation are strided over the threads.  The fetch-time coalescing is the
// reference's own sequential scan (:844-986, :753-820), done by one thread: with thousands of entries the fixed-point
// formulation above would need its arrays in global memory and a barrier per round, and a big batch count means few
// batches, so the scan of one batch overlaps the scans of the others.  Same outputs as k_fault_fetch_preprocess.
constexpr uint32_t kBigMaxBatch = 16384;
constexpr uint32_t kBigThreads = 1024;
constexpr size_t kBigBytesPerEntry = 8 + 8 + 8 + 4 + 2 + 2 + 2 + 2 + 1 + 1 + 1 + 1 + 1;   // inst, page, key, merged, num, utlb, ctx, gstart, acc, st, flags, mask8 (+ alignment slack)

struct fault_big_arrays {
    uint64_t *inst, *page, *key;
    uint32_t *merged;
    uint16_t *num, *utlb, *ctx, *gstart;
    uint8_t *acc, *st, *flags, *mask8;
};
// The arrays the sequential scan walks (28 bytes per entry) sit in shared memory when the batch fits (T <= kBigSmemEntries),
// else in the working space with the rest.
constexpr uint32_t kBigSmemEntries = 2048;   // (at 4096 the 147 KB leave one CTA per SM: measured no faster than global)
constexpr size_t kBigScanBytesPerEntry = 8 + 8 + 4 + 2 + 2 + 1 + 1 + 1 + 1;
__device__ __forceinline__ fault_big_arrays fault_big_carve(unsigned char *base, unsigned char *scan_base, uint32_t T)
{
    fault_big_arrays a;
    unsigned char *p = scan_base;
    a.inst = (uint64_t *)p; p += (size_t)T * 8;
    a.page = (uint64_t *)p; p += (size_t)T * 8;
    a.merged = (uint32_t *)p; p += (size_t)T * 4;
    a.num = (uint16_t *)p; p += (size_t)T * 2;
    a.utlb = (uint16_t *)p; p += (size_t)T * 2;
    a.acc = p; p += T;
    a.st = p; p += T;
    a.flags = p; p += T;
    a.mask8 = p; p += T;
    unsigned char *q = scan_base == base ? p : base;
    q = (unsigned char *)(((uintptr_t)q + 7) & ~(uintptr_t)7);
    a.key = (uint64_t *)q; q += (size_t)T * 8;
    a.ctx = (uint16_t *)q; q += (size_t)T * 2;
    a.gstart = (uint16_t *)q;
    return a;
}

UVMTIER_KERNEL void __launch_bounds__(kBigThreads)
k_fault_fetch_preprocess_big(fault_kparams prm, uint32_t T, unsigned char *__restrict__ work, uvmtier_fault_packet_t *__restrict__ ring,
                             const uvmtier_instance_ptr_entry_t *__restrict__ tbl, uint32_t *__restrict__ disp,
                             uvmtier_ordered_fault_t *__restrict__ ordered, uvmtier_fault_group_t *__restrict__ groups,
                             uvmtier_fault_batch_t *__restrict__ batches, uvmtier_fault_buffer_entry_t *__restrict__ entries,
                             uint16_t *__restrict__ utlb_pending)
{
    __shared__ uint32_t s_first_invalid, s_coalesced, s_max_utlb, s_single, s_xlat_fail, s_ngroups;
    __shared__ uint32_t s_scan[kBigThreads];
    UVMTIER_DYN_SMEM(smem);                                               // per uTLB: last representative (int32) and pending count (uint32)
    int32_t *s_last = (int32_t *)smem;
    uint32_t *s_cnt = (uint32_t *)(smem + (size_t)prm.utlb_count * 4);
    unsigned char *s_scan_arrays = smem + (((size_t)prm.utlb_count * 8 + 15) & ~(size_t)15);
    const uint32_t tid = threadIdx.x, NT = blockDim.x, U = prm.utlb_count;
    const uint32_t base = blockIdx.x * prm.B;
    const uint32_t nb = min(prm.B, prm.n_entries - base);
    unsigned char *my_work = work + (size_t)blockIdx.x * T * kBigBytesPerEntry;
    const fault_big_arrays a = fault_big_carve(my_work, T <= kBigSmemEntries ? s_scan_arrays : my_work, T);

    // ---- phase 1: load + decode ----
    if (tid == 0) { s_first_invalid = nb; s_xlat_fail = 0; }
    for (uint32_t k = tid; k < U; k += NT) { s_last[k] = -1; s_cnt[k] = 0; }
    __syncthreads();
    for (uint32_t i = tid; i < nb; i += NT) {
        const uint64_t at = (uint64_t)prm.get + base + i;
        const uint32_t ring_idx = (uint32_t)(at >= prm.ring_entries ? at - prm.ring_entries : at);
        const uint4 x = ld_stream_v4((const uint4 *)&ring[ring_idx]);
        const uint4 y = ld_stream_v4((const uint4 *)&ring[ring_idx] + 1);
        fault_fields f;
        fd_decode(x.x, x.y, x.z, x.w, y.x, y.y, y.z, y.w, prm.utlb_per_gpc, prm.va_bits, &f);
        if (!f.valid)
            atomicMin(&s_first_invalid, i);
        a.inst[i] = ((uint64_t)f.inst_aperture << 60) | ((f.inst_address >> 12) << 8) | f.ve_id;
        a.page[i] = f.fault_address >> 12;
        a.acc[i] = (uint8_t)f.access_type;
        a.utlb[i] = (uint16_t)f.utlb_id;
        a.flags[i] = (uint8_t)f.flags;
        a.num[i] = 0;
        a.mask8[i] = 0;
    }
    __syncthreads();
    const uint32_t n = s_first_invalid;                                   // num_cached_faults (:899-903)
    if (prm.clear_valid)
        for (uint32_t i = tid; i < n; i += NT) {
            const uint64_t at = (uint64_t)prm.get + base + i;
            const uint32_t ring_idx = (uint32_t)(at >= prm.ring_entries ? at - prm.ring_entries : at);
            ring[ring_idx].w[7] &= 0x7fffffffu;                           // entry_clear_valid
        }

    // ---- phases 2-4: the reference's scan, one thread ----
    if (tid == 0) {
        int32_t last = -1;
        uint32_t coalesced = 0, max_utlb = 0;
        bool single = true;
        for (uint32_t i = 0; i < n; i++) {
            const uint64_t inst = a.inst[i], page = a.page[i];
            const uint32_t acc = a.acc[i], u = a.utlb[i];
            const bool fatal = (a.flags[i] & UVMTIER_FE_FATAL) != 0;
            bool merged = false, same_inst = true;
            if (u > max_utlb) max_utlb = u;
            if (i > 0) {
                same_inst = inst == a.inst[last];                         // :944
                if (prm.coalesce && !fatal) {                             // :947-956 -> :782-820
                    const int32_t lt = s_cnt[u] ? s_last[u] : -1;
                    if (lt >= 0 && inst == a.inst[lt] && page == a.page[lt]) {
                        // per-uTLB rule (:753-780): the entry joins lt's segment; a higher access type takes over
                        a.num[lt]++;
                        a.mask8[lt] |= (uint8_t)(1u << acc);
                        if (acc > a.acc[lt]) {
                            a.mask8[i] = a.mask8[lt];
                            a.num[i] = a.num[lt];
                            a.flags[lt] |= UVMTIER_FE_FILTERED;
                            s_last[u] = (int32_t)i;                       // the batch-wide `last` is NOT moved (:804-809)
                        }
                        else
                            a.flags[i] |= UVMTIER_FE_FILTERED;
                        a.st[i] = ST_T;
                        a.merged[i] = a.st[lt] == ST_NEW ? (uint32_t)lt : a.merged[lt];      // the segment's head, for now
                        merged = true;
                    }
                    else if (same_inst && page == a.page[last] && acc <= a.acc[last]) {
                        a.num[last]++;                                    // lands on `last` even if that was displaced (:804-817)
                        a.mask8[last] |= (uint8_t)(1u << acc);
                        a.flags[i] |= UVMTIER_FE_FILTERED;
                        a.st[i] = ST_G;
                        a.merged[i] = (uint32_t)last;
                        merged = true;
                    }
                }
            }
            if (!merged) {
                if (single && !same_inst) single = false;                 // :959-960
                a.num[i] = 1;
                a.mask8[i] = (uint8_t)(1u << acc);
                s_cnt[u]++;
                s_last[u] = (int32_t)i;
                last = (int32_t)i;
                coalesced++;
                a.st[i] = ST_NEW;
                a.merged[i] = i;
            }
        }
        // second pass: NEW and T entries report their segment's representative (the one entry of it that is not filtered)
        for (uint32_t i = 0; i < n; i++)
            if (a.st[i] != ST_G && !(a.flags[i] & UVMTIER_FE_FILTERED))
                a.key[a.st[i] == ST_NEW ? i : a.merged[i]] = i;           // key[] is free until phase 5: rep of head h at key[h]
        for (uint32_t i = 0; i < n; i++)
            if (a.st[i] != ST_G)
                a.merged[i] = (uint32_t)a.key[a.st[i] == ST_NEW ? i : a.merged[i]];
        s_coalesced = coalesced; s_max_utlb = max_utlb; s_single = single ? 1u : 0u;
    }
    __syncthreads();
    const uint32_t num_coalesced = s_coalesced;
    if (utlb_pending)
        for (uint32_t k = tid; k < U; k += NT)
            utlb_pending[(size_t)blockIdx.x * U + k] = (uint16_t)s_cnt[k];

    // ---- phase 5: translation + sort keys ----
    const uint32_t ib = prm.idx_bits;
    for (uint32_t i = tid; i < T; i += NT) {
        uint64_t key = ~0ull;
        if (i < n && !(a.flags[i] & UVMTIER_FE_FILTERED)) {
            const uint64_t inst = a.inst[i];
            const uint64_t inst_address = ((inst >> 8) & ((1ull << 52) - 1)) << 12;
            const uint32_t aperture = (uint32_t)(inst >> 60), ve_id = (uint32_t)(inst & 0xffu);
            int ctx = -1;
            uint32_t rank = 0;
            for (uint32_t k = 0; k < prm.tbl_n; k++) {
                const uvmtier_instance_ptr_entry_t r = tbl[k];
                if (r.instance_ptr == inst_address && r.aperture == aperture && (r.ve_id == UVMTIER_VE_ID_ANY || r.ve_id == ve_id)) {
                    ctx = (int)k;
                    rank = r.sort_rank;
                    break;
                }
            }
            a.ctx[i] = (uint16_t)ctx;
            if (ctx < 0)
                s_xlat_fail = 1;
            key = ((uint64_t)rank << (ib + 3 + kPageBits)) | ((a.page[i] & ((1ull << kPageBits) - 1)) << (ib + 3)) |
                  ((uint64_t)(7u - a.acc[i]) << ib) | (uint64_t)i;
        }
        a.key[i] = key;
    }
    __syncthreads();
    const bool xlat_fail = s_xlat_fail != 0;
    if (xlat_fail)
        for (uint32_t i = tid; i < T; i += NT) a.key[i] = ~0ull;
    __syncthreads();
    // bitonic sort of T keys in the working space
    for (uint32_t k = 2; k <= T; k <<= 1)
        for (uint32_t j = k >> 1; j > 0; j >>= 1) {
            for (uint32_t t = tid; t < T / 2; t += NT) {
                const uint32_t lo = ((t & ~(j - 1)) << 1) | (t & (j - 1)), hi = lo | j;
                const uint64_t x = a.key[lo], y = a.key[hi];
                const bool up = (lo & k) == 0;
                if ((x > y) == up) { a.key[lo] = y; a.key[hi] = x; }
            }
            __syncthreads();
        }

    // ---- phase 6: outputs + va_block runs ----
    const uint32_t m = xlat_fail ? 0u : num_coalesced;
    const uint32_t idxmask = (1u << ib) - 1u;
    // run boundaries: each thread takes a contiguous stretch of the sorted view
    const uint32_t per = (m + NT - 1) / NT, p0 = min(tid * per, m), p1 = min(p0 + per, m);
    uint32_t mine = 0;
    for (uint32_t p = p0; p < p1; p++) {
        const uint64_t key = a.key[p];
        const uint32_t idx = (uint32_t)key & idxmask;
        const bool sfatal = (a.flags[idx] & UVMTIER_FE_FATAL) != 0;
        bool boundary = p == 0;
        if (p > 0) {
            const uint64_t prev = a.key[p - 1];
            const bool pfatal = (a.flags[(uint32_t)prev & idxmask] & UVMTIER_FE_FATAL) != 0;
            boundary = (prev >> (ib + 3 + 9)) != (key >> (ib + 3 + 9)) || sfatal || pfatal;
        }
        mine += boundary ? 1u : 0u;
        uvmtier_ordered_fault_t o;
        o.fault_address = a.page[idx] << 12;
        o.fault_index = (uint16_t)idx;
        o.ctx = a.ctx[idx];
        o.num_instances = a.num[idx];
        o.access_type_mask = a.mask8[idx];
        o.fault_access_type = (uint8_t)(a.acc[idx] | (sfatal ? UVMTIER_OF_FATAL : 0u));
        ordered[base + p] = o;
    }
    s_scan[tid] = mine;
    __syncthreads();
    if (tid == 0) {
        uint32_t run = 0;
        for (uint32_t t = 0; t < NT; t++) { const uint32_t c = s_scan[t]; s_scan[t] = run; run += c; }
        s_ngroups = run;
    }
    __syncthreads();
    uint32_t gidx = s_scan[tid];
    for (uint32_t p = p0; p < p1; p++) {
        const uint64_t key = a.key[p];
        const uint32_t idx = (uint32_t)key & idxmask;
        const bool sfatal = (a.flags[idx] & UVMTIER_FE_FATAL) != 0;
        bool boundary = p == 0;
        if (p > 0) {
            const uint64_t prev = a.key[p - 1];
            const bool pfatal = (a.flags[(uint32_t)prev & idxmask] & UVMTIER_FE_FATAL) != 0;
            boundary = (prev >> (ib + 3 + 9)) != (key >> (ib + 3 + 9)) || sfatal || pfatal;
        }
        if (boundary)
            a.gstart[gidx++] = (uint16_t)p;
    }
    __syncthreads();
    const uint32_t ngroups = s_ngroups;
    for (uint32_t g = tid; g < ngroups; g += NT) {
        const uint32_t p = a.gstart[g], next = g + 1 < ngroups ? a.gstart[g + 1] : m;
        const uint32_t idx = (uint32_t)a.key[p] & idxmask;
        uvmtier_fault_group_t grp;
        grp.block_start = (a.page[idx] << 12) & ~((uint64_t)UVMTIER_VA_BLOCK_SIZE - 1);
        grp.ctx = a.ctx[idx];
        grp.first_fault_index = (uint16_t)p;
        grp.block_faults = (uint16_t)(next - p);
        grp.flags = (a.flags[idx] & UVMTIER_FE_FATAL) ? 1u : 0u;
        groups[base + g] = grp;
    }
    for (uint32_t i = tid; i < nb; i += NT) {
        uint32_t d = 0;
        if (i < n) {
            const bool filtered = (a.flags[i] & UVMTIER_FE_FILTERED) != 0, fatal = (a.flags[i] & UVMTIER_FE_FATAL) != 0;
            d = (a.merged[i] & UVMTIER_DISP_INDEX_MASK) | (filtered ? UVMTIER_DISP_FILTERED : 0u) | (fatal ? UVMTIER_DISP_FATAL : 0u) |
                UVMTIER_DISP_VALID | ((uint32_t)a.st[i] << UVMTIER_DISP_RULE_SHIFT);
        }
        disp[base + i] = d;
    }
    if (entries)
        for (uint32_t i = tid; i < n; i += NT) {
            // decode again (cheaper than keeping every field): the packet's VALID bit may be cleared by now, nothing else
            const uint64_t at = (uint64_t)prm.get + base + i;
            const uint32_t ring_idx = (uint32_t)(at >= prm.ring_entries ? at - prm.ring_entries : at);
            const uint4 x = *((const uint4 *)&ring[ring_idx]);
            const uint4 y = *((const uint4 *)&ring[ring_idx] + 1);
            fault_fields f;
            fd_decode(x.x, x.y, x.z, x.w, y.x, y.y, y.z, y.w | 0x80000000u, prm.utlb_per_gpc, prm.va_bits, &f);
            const bool live = !(a.flags[i] & UVMTIER_FE_FILTERED);
            uvmtier_fault_buffer_entry_t e;
            e.fault_address = f.fault_address;
            e.timestamp = f.timestamp;
            e.instance_ptr_address = f.inst_address;
            e.client_id = (uint16_t)f.client_id;
            e.mmu_engine_id = (uint16_t)f.engine_id;
            e.utlb_id = (uint16_t)f.utlb_id;
            e.num_instances = live ? a.num[i] : (uint16_t)0;
            e.instance_ptr_aperture = (uint8_t)f.inst_aperture;
            e.fault_type = (uint8_t)f.fault_type;
            e.fault_access_type = (uint8_t)f.access_type;
            e.client_type = (uint8_t)f.client_type;
            e.mmu_engine_type = (uint8_t)f.engine_type;
            e.gpc_id = (uint8_t)f.gpc_id;
            e.ve_id = (uint8_t)f.ve_id;
            e.flags = a.flags[i];
            e.access_type_mask = live ? a.mask8[i] : (uint8_t)0;
            for (int k = 0; k < 7; k++) e.pad[k] = 0;
            entries[base + i] = e;
        }
    if (tid == 0) {
        uvmtier_fault_batch_t bt;
        bt.num_cached_faults = n;
        bt.num_coalesced_faults = num_coalesced;
        bt.is_single_instance_ptr = s_single;
        bt.max_utlb_id = s_max_utlb;
        bt.status = xlat_fail ? NV_WARN_MORE_PROCESSING_REQUIRED : NV_OK;
        bt.num_duplicate_faults = 0;
        bt.num_invalid_prefetch_faults = 0;
        bt.num_block_groups = xlat_fail ? 0u : ngroups;
        batches[blockIdx.x] = bt;
    }
}

} // namespace uvmtier_k
