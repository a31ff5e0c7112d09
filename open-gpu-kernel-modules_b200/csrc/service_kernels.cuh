// service_kernels.cuh -- stage C/D/E: servicing the sorted fault batches per va_block.
// Device code only (host side: service.cu); also compiled by tests/simt for the CPU emulator.
//
// Reference (UVM = kernel-open/nvidia-uvm): service_fault_batch (uvm_gpu_replayable_faults.c:2232-2377),
// service_fault_batch_dispatch (:1946-2035), service_fault_batch_block_locked (:1375-1597),
// check_fault_access_permissions (:1293-1360), uvm_va_block_service_locked (uvm_va_block.c:12307-12359) with
// uvm_va_block_get_prefetch_hint (:11828-11881), block_copy_resident_pages (:4660-4847),
// uvm_va_block_make_resident_finish (:5042-5084) and the mapping step of uvm_va_block_service_finish
// (:12063-12302, compute_new_permission :11019-11066).
//
// The reference services batch after batch on one thread.  Different va_blocks never interact, so here the
// runs (one va_block's faults of one batch) are first regrouped by va_block, then every va_block's runs are
// applied in batch order by a small group of lanes or by a warp:
//   service_lists.cuh   k_service_count / _alloc / _fill / _sort_* -> per state slot the sorted list of its runs
//   service_grp.cuh     k_service_blocks_grp                         -> va_blocks with few entries, G lanes each
//   this file           k_service_blocks                             -> va_blocks with many entries, a warp each
//                       k_service_unmapped -> runs whose address lies in no va_range (fatal / invalid prefetch,
//                                             :2022-2031)
// In k_service_blocks a va_block's masks stay in registers across all its runs: one 32-bit word per lane, lanes
// 16-31 mirror lanes 0-15 (the mask helpers are half-warp wide) so that control flow is uniform across the warp.
#pragma once

#include "devio.cuh"
#include "fault_decode.cuh"
#include "prefetch.cuh"
#include "thrashing_hint_dev.cuh"

namespace uvmtier_k {

constexpr int kSvcThreads = 256;
constexpr uint32_t kMaxSlots = 1u << 26;    // va_block slots of one state table (128 TB of VA)

enum : uint32_t { PROT_NONE = 0, PROT_RO = 1, PROT_RW = 2, PROT_RWA = 3 };
enum : uint32_t { REASON_INVALID_ADDRESS = 1, REASON_INVALID_PERMISSIONS = 2 };   // UvmEventFatalReason
enum : uint32_t { CANCEL_ALL = 0, CANCEL_WRITE_AND_ATOMIC = 1 };                  // uvm_fault_cancel_va_mode_t

struct service_kparams {
    uint64_t va_base;
    uint32_t B, n_batches;
    uint32_t blocks_per_space, n_spaces, gpu_id, n_slots;
    uint32_t pf_enable, pf_threshold, pf_min_faults;
    uint32_t run_cap;
    uint32_t B_shift;                          // log2(B) when B is a power of two, else 0xffffffff
    uint32_t batch_base;                       // batch number of d_batches[0] (a pass over one chunk of a longer trace)
    uint32_t non_replayable;                   // servicing non-replayable faults (UVM/uvm_gpu_non_replayable_faults.c:345-445):
                                               // no "already authorised" shortcut, no cancel mode in the flags
};

__host__ __device__ __forceinline__ uint32_t svc_log2_or_none(uint32_t B)
{
    for (uint32_t s = 0; s < 32; s++)
        if ((1u << s) == B)
            return s;
    return 0xffffffffu;
}

// Batch of ordered-array / group position t (B entries per batch).
__device__ __forceinline__ uint32_t svc_batch_of(const service_kparams &prm, uint32_t t)
{
    return prm.B_shift != 0xffffffffu ? t >> prm.B_shift : t / prm.B;
}

struct block_state_ptrs {
    uint32_t *gpu_resident;
    uint32_t *cpu_resident[2];
    uint32_t *gpu_pte[3];
    uvmtier_prefetch_info_t *prefetch_info;
    uint8_t *logical_prot;
    uint16_t *preferred_location;
    uint32_t *gpu_evicted;                     // may be NULL
    uint32_t *read_duplicated;                 // NULL = read duplication is off for the table
};

// uvm_fault_access_type_to_prot, UVM/uvm_hal_types.h:522-537
__device__ __forceinline__ uint32_t svc_prot_of(uint32_t acc)
{
    // PREFETCH, READ -> RO; WRITE, ATOMIC_WEAK -> RW; ATOMIC_STRONG -> RWA; anything else -> RO: two bits per
    // access type, looked up in a constant
    return (0x57A5u >> (2u * (acc & 7u))) & 3u;
}

// State slot of a run, or 0xFFFFFFFF when its va_block lies in no va_range of its VA space.
__device__ __forceinline__ uint32_t svc_slot(const service_kparams &prm, uint32_t va_space, uint64_t block_start)
{
    const uint64_t off = block_start - prm.va_base;
    if (block_start < prm.va_base || (off >> 21) >= prm.blocks_per_space || va_space >= prm.n_spaces)
        return 0xffffffffu;
    return va_space * prm.blocks_per_space + (uint32_t)(off >> 21);
}

// ------------------------------------------------------------------------------------------------
// Counts (and, with runs != NULL, writes) the copy runs of mask `m`: maximal sequences of consecutive pages
// with the same cause (UVM/uvm_va_block.c:4242-4335).  Warp-synchronous; word per lane, half-warp wide.
// Lanes 16-31 mirror lanes 0-15 in most callers and do not write (both_halves = false); with both_halves the two
// half-warps work on two different va_blocks and both write.
__device__ __forceinline__ uint32_t svc_emit_runs(uint32_t m, uint32_t pf, uint32_t slot, uint32_t src,
                                                  uvmtier_copy_run_t *runs, uint32_t at, uint32_t run_cap, int lane,
                                                  uint32_t dst = UVMTIER_RUN_DST_HBM, bool both_halves = false)
{
    const int l = lane & 15;
    uint32_t pm = __shfl_up_sync(HW_FULL, m >> 31, 1, 16);
    uint32_t pp = __shfl_up_sync(HW_FULL, pf >> 31, 1, 16);
    if (l == 0) { pm = 0; pp = 0; }
    const uint32_t m_prev = (m << 1) | pm, pf_prev = (pf << 1) | pp;
    uint32_t start = m & (~m_prev | (pf ^ pf_prev));
    const uint32_t stop = ~m | start;                       // a run cannot extend through these positions
    const uint32_t nstart = __popc(start);
    // exclusive prefix of nstart over the 16 lanes, and the total
    uint32_t incl = nstart;
    for (int d = 1; d < 16; d <<= 1) {
        const uint32_t y = __shfl_up_sync(HW_FULL, incl, d, 16);
        if (l >= d) incl += y;
    }
    const uint32_t total = __shfl_sync(HW_FULL, incl, 15, 16);
    if (!runs)
        return total;
    // first stop at or after the start of each FOLLOWING word (suffix minimum), 512 if none
    uint32_t mine = stop ? (uint32_t)(32 * l + __ffs((int)stop) - 1) : 512u;
    uint32_t next_abs = __shfl_down_sync(HW_FULL, mine, 1, 16);
    if (l == 15) next_abs = 512u;
    for (int d = 1; d < 16; d <<= 1) {
        const uint32_t y = __shfl_down_sync(HW_FULL, next_abs, d, 16);
        if (l + d <= 15) next_abs = min(next_abs, y);
    }
    uint32_t k = at + incl - nstart;
    while (start) {
        const int s = __ffs((int)start) - 1;
        start &= start - 1;
        const uint32_t rest = s == 31 ? 0u : (stop >> (s + 1));
        const uint32_t first = (uint32_t)(32 * l + s);
        const uint32_t end = rest ? first + (uint32_t)__ffs((int)rest) : next_abs;
        if (k < run_cap && (lane < 16 || both_halves)) {
            uvmtier_copy_run_t r;
            r.slot = slot;
            r.first_page = (uint16_t)first;
            r.num_pages = (uint16_t)(end - first);
            r.src = (uint16_t)src;
            r.prefetch = (uint16_t)((pf >> s) & 1u);
            r.dst = (uint16_t)dst;
            runs[k] = r;
        }
        k++;
    }
    return total;
}

// What one entry of a run does, as a function of what the sorted entry before it did (service_fault_batch_block_locked
// :1433-1580 with fault_entry_duplicate_flags :1257-1272 and check_fault_access_permissions :1293-1360).
// state bit 0 = that entry was fatal, bit 1 = it was an invalid prefetch; `auth` = the page is already mapped with
// READ (bit 0) / WRITE (bit 1) / ATOMIC (bit 2) permission.
struct svc_entry_eval {
    uint32_t flags;        // UVMTIER_SF_*
    uint32_t state;        // handed to the next entry
    uint32_t ip_add;       // added to the batch's num_invalid_prefetch_faults
    bool serviced;         // the page joins new_residency
};
__device__ __forceinline__ svc_entry_eval svc_eval_entry(uint32_t acc, uint32_t num_instances, uint32_t access_type_mask, bool dup,
                                                         uint32_t state_in, uint32_t logical, uint32_t auth)
{
    svc_entry_eval r;
    const bool prev_fatal = state_in & 1u, prev_ip = state_in & 2u;
    uint32_t service_acc = acc, fl = 0, ip_add = 0;
    bool entry_fatal = false, entry_ip = false, done = false, serviced = false;
    if (dup) {
        fl |= UVMTIER_SF_DUPLICATE;
        if (prev_ip) {                                        // fault_entry_duplicate_flags :1257-1272
            entry_ip = true;
            ip_add += num_instances;
        }
        if (!prev_fatal)
            done = true;                                      // page already serviced (:1481-1483)
    }
    if (!done && svc_prot_of(acc) > logical) {                // check_fault_access_permissions :1293-1360
        if (acc == ACC_PREFETCH) {
            entry_ip = true;
            ip_add += num_instances;
            done = true;
        }
        else {
            uint32_t cancel = CANCEL_ALL;
            service_acc = ACC_COUNT;
            if (acc > ACC_READ) {
                cancel = CANCEL_WRITE_AND_ATOMIC;
                if (access_type_mask & (1u << ACC_READ)) {
                    if (svc_prot_of(ACC_READ) <= logical)
                        service_acc = ACC_READ;
                    else
                        cancel = CANCEL_ALL;
                }
            }
            entry_fatal = true;
            fl |= UVMTIER_SF_FATAL | (REASON_INVALID_PERMISSIONS << UVMTIER_SF_REASON_SHIFT) | (cancel << UVMTIER_SF_CANCEL_SHIFT);
            if (service_acc == ACC_COUNT)
                done = true;
        }
    }
    if (!done) {
        // already authorised: nothing to do (:1518-1522, uvm_va_block.c:2576-2615); otherwise the page moves to the
        // faulting GPU (harness policy; block_select_processor_residency falls through to the faulting processor,
        // uvm_va_block.c:11560-11712)
        if ((auth >> (svc_prot_of(service_acc) - 1u)) & 1u) {
            fl |= UVMTIER_SF_AUTHORIZED;
        }
        else {
            serviced = true;
            fl |= UVMTIER_SF_SERVICED | (service_acc << UVMTIER_SF_ACCESS_SHIFT);
        }
    }
    if (entry_ip)
        fl |= UVMTIER_SF_INVALID_PREFETCH;
    r.flags = fl;
    r.state = (entry_fatal ? 1u : 0u) | (entry_ip ? 2u : 0u);
    r.ip_add = ip_add;
    r.serviced = serviced;
    return r;
}

// Thrashing hints inside the fault loop (UVM/uvm_gpu_replayable_faults.c:1524-1545), for the va_blocks the detector tracks
// page by page (uvmtier_service_fault_batches_thrashing_device; th.blocks == NULL: off).  batch_time[b]: NV_GETTIME() when
// batch b is serviced.  tally: [3] throttles [4] pins on this GPU [5] pins elsewhere.
struct svc_thrash_params {
    thrash_hint_tables th;
    const uvmtier_thrashing_hint_config_t *cfg;
    const unsigned long long *batch_time;
    uint32_t *tally;
};

// One warp per slot, for the slots the sorter put on the `heavy` list: runs with many entries.  The entries of a
// run are filtered 32 at a time.  An entry depends on the one before it only when both fault on the same page
// (a duplicate inherits "fatal" / "invalid prefetch"), so every lane evaluates its entry from its neighbour's last
// result and the warp repeats until nothing changes: as many rounds as the longest chain of duplicates, and the
// result of the sequential scan.
UVMTIER_KERNEL void __launch_bounds__(kSvcThreads)
k_service_blocks(service_kparams prm, const uvmtier_ordered_fault_t *__restrict__ ordered,
                 const uvmtier_fault_group_t *__restrict__ groups, uvmtier_fault_batch_t *__restrict__ batches,
                 block_state_ptrs st, const uint2 *__restrict__ cw, const uint32_t *__restrict__ offset,
                 const uint32_t *__restrict__ list, const uint32_t *__restrict__ heavy, uint16_t *__restrict__ entry_flags,
                 uvmtier_service_group_t *__restrict__ service, uvmtier_copy_run_t *__restrict__ runs,
                 uint32_t *__restrict__ counters, svc_thrash_params tp)
{
    __shared__ uint32_t s_newres[kSvcThreads / 32][16];
    const int lane = threadIdx.x & 31, l = lane & 15, warp = threadIdx.x >> 5;
    const uint32_t n_warps = gridDim.x * (kSvcThreads / 32);
    const uint32_t n_heavy = heavy[0];
    const uint32_t flag_mask = prm.non_replayable ? ~(1u << UVMTIER_SF_CANCEL_SHIFT) : ~0u;
    const uint4 *oraw = (const uint4 *)ordered, *graw = (const uint4 *)groups;
    for (uint32_t q = blockIdx.x * (kSvcThreads / 32) + warp; q < n_heavy; q += n_warps) {
        const uint32_t slot = heavy[1u + q];
        const uint2 cwv = cw[slot];
        const uint32_t cnt = cwv.x;
        // one run per lane pays only when runs are short; long runs are better served 32 entries at a time
        const bool short_runs = cwv.y <= 4u * cwv.x;
        const uint32_t off = offset[slot];
        const size_t w = (size_t)slot * 16 + l;
        uint32_t gres = st.gpu_resident[w], c0 = st.cpu_resident[0][w], c1 = st.cpu_resident[1][w];
        uint32_t pte0 = st.gpu_pte[0][w], pte1 = st.gpu_pte[1][w], pte2 = st.gpu_pte[2][w];
        const uint32_t logical = st.logical_prot[slot], pref = st.preferred_location[slot];
        uvmtier_prefetch_info_t pi = st.prefetch_info[slot];
        // a va_block the thrashing detector tracks page by page: every fault that would be serviced asks for a hint first,
        // one entry after the other (the hint advances the page's throttling / pinning state)
        uvmtier_thrashing_block_t tblk;
        tblk.tracked = 0;
        if (tp.th.blocks)
            tblk = tp.th.blocks[slot];
        const bool tracked = tblk.tracked != 0;
        uint32_t tally[3] = {0u, 0u, 0u};

        // Three runs in flight (as in k_service_blocks_grp): run k+1's first 32 entries, run k+2's group and run
        // k+3's id are loaded while run k is worked on -- a hot va_block is one long chain of dependent steps.
        uint32_t t0 = 0, t1 = 0, t2 = 0;
        uint4 g0 = make_uint4(0, 0, 0, 0), g1 = g0, e0 = g0;
        auto prime = [&](uint32_t at) {                                   // (re)fills the pipeline for run `at`
            t0 = list[off + at];
            t1 = at + 1 < cnt ? list[off + at + 1] : t0;
            t2 = at + 2 < cnt ? list[off + at + 2] : t0;
            g0 = graw[t0];
            g1 = graw[t1];
            e0 = make_uint4(0, 0, 0, 0);
            if ((uint32_t)lane < (g0.w & 0xffffu))
                e0 = oraw[svc_batch_of(prm, t0) * prm.B + (g0.z >> 16) + lane];
        };
        prime(0);

        // A run in which no entry needs service leaves the block's state as it was.  Once that happens (typically:
        // the block is fully resident and mapped, and faults on it are stale), the next runs are independent of each
        // other as long as they also need no service: they are evaluated 32 at a time, one run per lane, against the
        // current masks; the runs before the first one that does need service are final, that one goes through the
        // regular path below.  A hot va_block is otherwise one chain of thousands of dependent steps.
        bool streak = false;
        for (uint32_t k = 0; k < cnt;) {
            if (short_runs && streak && cnt - k >= 2 && !tracked) {
                const uint32_t m = min(32u, cnt - k);
                const bool mine = (uint32_t)lane < m;
                uint32_t ts = 0;
                uint4 gs = make_uint4(0, 0, 0, 0);
                if (mine) {
                    ts = list[off + k + lane];
                    gs = graw[ts];
                }
                const uint32_t bs = svc_batch_of(prm, ts), base_s = bs * prm.B, first_s = gs.z >> 16;
                const uint32_t ns = mine ? (gs.w & 0xffffu) : 0u;
                const uint32_t ns_max = __reduce_max_sync(HW_FULL, ns);
                uint32_t state = 0, prev_page = 0xffffffffu, dsum = 0, isum = 0;
                bool needs_service = false;
                for (uint32_t i = 0; i < ns_max; i++) {
                    const bool on = i < ns;
                    uint4 e = make_uint4(0, 0, 0, 0);
                    if (on)
                        e = oraw[base_s + first_s + i];
                    const uint32_t acc = (e.w >> 24) & 7u, num_instances = e.w & 0xffffu, access_type_mask = (e.w >> 16) & 0xffu;
                    const uint32_t page = on ? (e.x - gs.x) >> 12 : 0u;
                    const bool dup = on && i > 0 && page == prev_page;
                    const int wl = (int)((page >> 5) & 15u);
                    const uint32_t w0 = __shfl_sync(HW_FULL, pte0, wl), w1 = __shfl_sync(HW_FULL, pte1, wl), w2 = __shfl_sync(HW_FULL, pte2, wl);
                    const uint32_t sh = page & 31u;
                    const uint32_t auth = prm.non_replayable ? 0u : ((w0 >> sh) & 1u) | (((w1 >> sh) & 1u) << 1) | (((w2 >> sh) & 1u) << 2);
                    const svc_entry_eval ev = svc_eval_entry(acc, num_instances, access_type_mask, dup, state, logical, auth);
                    if (on) {
                        entry_flags[base_s + first_s + i] = (uint16_t)(ev.flags & flag_mask);   // rewritten later if the run is not final
                        dsum += dup ? num_instances : num_instances - 1u;
                        isum += ev.ip_add;
                        needs_service |= ev.serviced;
                        state = ev.state;
                        prev_page = page;
                    }
                }
                const uint32_t bad = __ballot_sync(HW_FULL, mine && needs_service);
                const uint32_t adv = bad ? (uint32_t)__ffs((int)bad) - 1u : m;
                if ((uint32_t)lane < adv) {
                    if (dsum) atomicAdd(&batches[bs].num_duplicate_faults, dsum);
                    if (isum) atomicAdd(&batches[bs].num_invalid_prefetch_faults, isum);
                    uint4 *rec = (uint4 *)&service[ts];
                    rec[0] = make_uint4(slot, 0u, 0u, 0u);
                    rec[1] = make_uint4((uint32_t)UVMTIER_ID_INVALID << 16, 0u, (bs + prm.batch_base) << 16, ts - base_s);
                    rec[2] = rec[3] = rec[4] = rec[5] = make_uint4(0u, 0u, 0u, 0u);
                }
                k += adv;
                streak = adv == m;
                if (k < cnt)
                    prime(k);
                if (streak || k >= cnt)
                    continue;
            }
            const uint32_t t3 = k + 3 < cnt ? list[off + k + 3] : t0;
            const uint4 g2 = graw[t2];
            uint4 e1 = make_uint4(0, 0, 0, 0);
            if ((uint32_t)lane < (g1.w & 0xffffu))
                e1 = oraw[svc_batch_of(prm, t1) * prm.B + (g1.z >> 16) + lane];
            const uint32_t t = t0;                                        // ascending (batch, run): the list is sorted
            const uint32_t b = svc_batch_of(prm, t), base = b * prm.B;
            const uint32_t first = g0.z >> 16, n = g0.w & 0xffffu;

            // ---- service_fault_batch_block_locked :1433-1580, 32 entries at a time ----
            if (lane < 16)
                s_newres[warp][lane] = 0;
            __syncwarp();
            uint32_t page_faults = 0, dup_sum = 0, ip_sum = 0, carry_page = 0xffffffffu, carry_state = 0;
            if (tracked) {
                // with hints: one entry at a time, the same entry on every lane
                uint32_t state = 0, prev_page = 0xffffffffu;
                for (uint32_t i = 0; i < n; i++) {
                    const uint4 e = oraw[base + first + i];
                    const uint32_t acc = (e.w >> 24) & 7u, num_instances = e.w & 0xffffu, access_type_mask = (e.w >> 16) & 0xffu;
                    const uint32_t page = (e.x - g0.x) >> 12;
                    const bool dup = i > 0 && page == prev_page;
                    const int wl = (int)((page >> 5) & 15u);
                    const uint32_t sh = page & 31u;
                    const uint32_t w0 = __shfl_sync(HW_FULL, pte0, wl), w1 = __shfl_sync(HW_FULL, pte1, wl), w2 = __shfl_sync(HW_FULL, pte2, wl);
                    const uint32_t auth = prm.non_replayable ? 0u : ((w0 >> sh) & 1u) | (((w1 >> sh) & 1u) << 1) | (((w2 >> sh) & 1u) << 2);
                    const svc_entry_eval ev = svc_eval_entry(acc, num_instances, access_type_mask, dup, state, logical, auth);
                    uint32_t fl = ev.flags;
                    bool serviced = ev.serviced;
                    if (serviced) {                                       // :1524-1545
                        const uint32_t on_gpu = (__shfl_sync(HW_FULL, gres, wl) >> sh) & 1u;
                        const uint32_t on_cpu = (__shfl_sync(HW_FULL, c0 | c1, wl) >> sh) & 1u;
                        uint32_t packed = 0;
                        if (lane == 0) {
                            uvmtier_thrashing_query_t q;
                            q.time_stamp = tp.batch_time[b + prm.batch_base];
                            q.slot = slot; q.page = (uint16_t)page; q.requester = (uint8_t)prm.gpu_id; q.reserved = 0;
                            q.resident_processors = (on_gpu ? 1u << prm.gpu_id : 0u) | on_cpu;
                            q.preferred_location = pref == UVMTIER_ID_INVALID ? (uint16_t)0xffffu : (uint16_t)pref;
                            q.reserved2 = 0;
                            const uvmtier_thrashing_hint_t h = hint_get(*tp.cfg, tp.th, q, tblk, tally);
                            packed = (uint32_t)h.type | ((uint32_t)h.pin_residency << 8);
                        }
                        packed = __shfl_sync(HW_FULL, packed, 0);
                        const uint32_t type = packed & 0xffu, where = packed >> 8;
                        if (type == UVMTIER_THRASHING_HINT_THROTTLE || (type == UVMTIER_THRASHING_HINT_PIN && where != prm.gpu_id)) {
                            // mark_fault_throttled (:1537) / a pin this pass does not carry out: not serviced
                            fl = (fl & ~(UVMTIER_SF_SERVICED | (7u << UVMTIER_SF_ACCESS_SHIFT))) |
                                 (type == UVMTIER_THRASHING_HINT_THROTTLE ? UVMTIER_SF_THROTTLED : UVMTIER_SF_PIN_REMOTE);
                            serviced = false;
                        }
                        else if (type == UVMTIER_THRASHING_HINT_PIN)
                            fl |= UVMTIER_SF_PINNED;                      // thrashing_pin_mask (:1541-1545)
                    }
                    if (lane == 0) {
                        entry_flags[base + first + i] = (uint16_t)(fl & flag_mask);
                        if (serviced)
                            s_newres[warp][page >> 5] |= 1u << sh;
                    }
                    dup_sum += dup ? num_instances : num_instances - 1u;
                    ip_sum += ev.ip_add;
                    page_faults += serviced ? 1u : 0u;
                    state = ev.state;
                    prev_page = page;
                }
            }
            else
            for (uint32_t c = 0; c < n; c += 32) {
                const uint32_t i = c + lane;
                const bool valid = i < n;
                uint4 e = e0;
                if (c && valid)
                    e = oraw[base + first + i];
                const uint32_t acc = (e.w >> 24) & 7u, num_instances = e.w & 0xffffu, access_type_mask = (e.w >> 16) & 0xffu;
                const uint32_t page = valid ? (e.x - g0.x) >> 12 : 0x80000000u + lane;
                uint32_t prev_page = __shfl_up_sync(HW_FULL, page, 1);
                if (lane == 0)
                    prev_page = carry_page;
                const bool dup = valid && i > 0 && page == prev_page;     // same va_space inside a run (:1180-1191)
                // is the page already mapped for READ / WRITE / ATOMIC?  (the masks do not change inside a run)
                const int wl = (int)((page >> 5) & 15u);
                const uint32_t w0 = __shfl_sync(HW_FULL, pte0, wl), w1 = __shfl_sync(HW_FULL, pte1, wl), w2 = __shfl_sync(HW_FULL, pte2, wl);
                const uint32_t sh = page & 31u;
                const uint32_t auth = prm.non_replayable ? 0u : ((w0 >> sh) & 1u) | (((w1 >> sh) & 1u) << 1) | (((w2 >> sh) & 1u) << 2);
                svc_entry_eval ev;
                uint32_t state = 0;
                for (;;) {
                    uint32_t state_in = __shfl_up_sync(HW_FULL, state, 1);
                    if (lane == 0)
                        state_in = carry_state;
                    ev = svc_eval_entry(acc, num_instances, access_type_mask, dup, dup ? state_in : 0u, logical, auth);
                    const bool changed = valid && ev.state != state;
                    if (valid)
                        state = ev.state;
                    if (!__any_sync(HW_FULL, changed))
                        break;
                }
                if (valid) {
                    entry_flags[base + first + i] = (uint16_t)(ev.flags & flag_mask);
                    if (ev.serviced)
                        atomicOr(&s_newres[warp][page >> 5], 1u << sh);
                }
                const uint32_t dup_add = valid ? (dup ? num_instances : num_instances - 1u) : 0u;   // :1200-1203
                dup_sum += __reduce_add_sync(HW_FULL, dup_add);
                ip_sum += __reduce_add_sync(HW_FULL, valid ? ev.ip_add : 0u);
                page_faults += (uint32_t)__popc(__ballot_sync(HW_FULL, valid && ev.serviced));
                carry_page = __shfl_sync(HW_FULL, page, 31);
                carry_state = __shfl_sync(HW_FULL, state, 31);
            }
            __syncwarp();
            uint32_t newres = s_newres[warp][l];
            __syncwarp();
            if (lane == 0) {
                if (dup_sum) atomicAdd(&batches[b].num_duplicate_faults, dup_sum);
                if (ip_sum) atomicAdd(&batches[b].num_invalid_prefetch_faults, ip_sum);
            }

            // ---- uvm_va_block_service_locked ----
            uint32_t hint = UVMTIER_ID_INVALID, pf = 0, n_pf = 0, n0 = 0, n1 = 0, nt = 0, run_off = 0, run_cnt = 0;
            int rf = 0, ro = 0;
            if (page_faults) {                                            // uniform across the warp
                rf = hw_first_bit(newres, lane);
                ro = hw_last_bit(newres, lane) + 1;
                const bool block_empty = !hw_any(gres | c0 | c1, lane);
                if (prm.pf_enable) {
                    const uint32_t flags = UVMTIER_PF_HAS_RESIDENT_MASK | UVMTIER_PF_GPU_VA_SPACE_REGISTERED |
                                           (block_empty ? UVMTIER_PF_BLOCK_RESIDENT_EMPTY : 0u);
                    const uint64_t block_start = ((uint64_t)g0.y << 32) | g0.x;
                    // thrashing pages seed no prefetch region and are not prefetched (uvm_perf_prefetch.c:378-381, 407)
                    const uint32_t thr_w = tracked ? *((volatile const uint32_t *)&tp.th.thrashing_pages[w]) : 0u;
                    pf = pf_get_hint_block(newres, gres, thr_w, 0u, tracked, block_start, 512, rf, ro, prm.gpu_id, pref, flags,
                                           16u, prm.pf_threshold, prm.pf_min_faults, &pi, &hint, lane);
                }
                if (hint != UVMTIER_ID_INVALID) {                         // uvm_va_block_get_prefetch_hint :11856-11876
                    n_pf = hw_weight(pf);
                    newres |= pf;
                    rf = hw_first_bit(newres, lane);
                    ro = hw_last_bit(newres, lane) + 1;
                }
                else {
                    pf = 0;
                }
                // block_copy_resident_pages :4660-4847: GPU destination, CPU NUMA nodes in ascending order
                uint32_t todo = newres & ~gres;
                const uint32_t from0 = todo & c0;
                todo &= ~from0;
                const uint32_t from1 = todo & c1;
                const uint32_t touch = todo & ~from1;                     // resident nowhere: populated + zeroed
                n0 = hw_weight(from0); n1 = hw_weight(from1); nt = hw_weight(touch);
                const uint32_t r0 = svc_emit_runs(from0, pf, slot, 0u, NULL, 0u, 0u, lane);
                const uint32_t r1 = svc_emit_runs(from1, pf, slot, 1u, NULL, 0u, 0u, lane);
                const uint32_t r2 = svc_emit_runs(touch, pf, slot, UVMTIER_RUN_SRC_ZERO, NULL, 0u, 0u, lane);
                run_cnt = r0 + r1 + r2;
                if (lane == 0 && run_cnt)
                    run_off = atomicAdd(&counters[0], run_cnt);
                run_off = __shfl_sync(HW_FULL, run_off, 0);
                svc_emit_runs(from0, pf, slot, 0u, runs, run_off, prm.run_cap, lane);
                svc_emit_runs(from1, pf, slot, 1u, runs, run_off + r0, prm.run_cap, lane);
                svc_emit_runs(touch, pf, slot, UVMTIER_RUN_SRC_ZERO, runs, run_off + r0 + r1, prm.run_cap, lane);
                // make_resident_copy first unmaps the pages that are not yet resident here (:4919-4926)
                {
                    const uint32_t arriving = newres & ~gres;
                    pte0 &= ~arriving; pte1 &= ~arriving; pte2 &= ~arriving;
                }
                // uvm_va_block_make_resident_finish :5042-5084
                gres |= newres;
                c0 &= ~newres;
                c1 &= ~newres;
                // service_finish: map the faulting GPU.  compute_new_permission (:11019-11066): the pages now live on
                // the faulting GPU, which has native atomics on its own memory, and nobody else holds an atomic
                // mapping of a page that just moved -> READ_WRITE_ATOMIC under a read-write VMA, READ_ONLY otherwise.
                pte0 |= newres;
                if (logical == PROT_RWA) {
                    pte1 |= newres;
                    pte2 |= newres;
                }
            }
            // ---- record ----
            if (lane < 16) {
                uint32_t *rec = (uint32_t *)&service[t];
                if (l < 8) {
                    uint32_t v = 0;
                    switch (l) {
                        case 0: v = slot; break;
                        case 1: v = (uint32_t)rf | ((uint32_t)ro << 16); break;
                        case 2: v = page_faults | (n_pf << 16); break;
                        case 3: v = n0 | (n1 << 16); break;
                        case 4: v = nt | (hint << 16); break;
                        case 5: v = run_off; break;
                        case 6: v = run_cnt | ((b + prm.batch_base) << 16); break;
                        case 7: v = t - base; break;
                    }
                    rec[l] = v;
                }
                rec[8 + l] = page_faults ? newres : 0u;
            }
            t0 = t1; t1 = t2; t2 = t3;
            g0 = g1; g1 = g2;
            e0 = e1;
            streak = page_faults == 0;
            k++;
        }
        if (lane < 16) {
            st.gpu_resident[w] = gres; st.cpu_resident[0][w] = c0; st.cpu_resident[1][w] = c1;
            st.gpu_pte[0][w] = pte0; st.gpu_pte[1][w] = pte1; st.gpu_pte[2][w] = pte2;
            // pages that an eviction had moved off the GPU and that came back (block_make_resident_clear_evicted,
            // uvm_va_block.c:4978-4988): evicted pages are never GPU resident, so whatever is resident now came back
            if (st.gpu_evicted)
                st.gpu_evicted[w] &= ~gres;
            // break_read_duplication_in_region (:4538-4579) clears read_duplicated for every page handed to
            // make_resident; those are exactly the pages that lost their CPU copy (a read-duplicated page of a
            // one-GPU block always has one, :5279-5310), so what still has a CPU copy stays duplicated
            if (st.read_duplicated)
                st.read_duplicated[w] &= c0 | c1;
            if (l == 0)
                st.prefetch_info[slot] = pi;
        }
        if (tracked && lane == 0) {
            tp.th.blocks[slot] = tblk;
            if (tally[0]) atomicAdd(&tp.tally[3], tally[0]);
            if (tally[1]) atomicAdd(&tp.tally[4], tally[1]);
            if (tally[2]) atomicAdd(&tp.tally[5], tally[2]);
        }
    }
}

// Runs whose va_block lies in no va_range: service_fault_batch_fatal_notify, one entry at a time
// (:1656-1698, :2022-2031).  One thread per batch that has such runs; sequential like the reference because an
// entry inherits the invalid-prefetch flag of the sorted entry before it.
UVMTIER_KERNEL void __launch_bounds__(128)
k_service_unmapped(service_kparams prm, const uvmtier_ordered_fault_t *__restrict__ ordered,
                   const uvmtier_fault_group_t *__restrict__ groups, uvmtier_fault_batch_t *__restrict__ batches,
                   const uvmtier_instance_ptr_entry_t *__restrict__ tbl, const uint32_t *__restrict__ batch_unmapped,
                   uint16_t *__restrict__ entry_flags, uvmtier_service_group_t *__restrict__ service)
{
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= prm.n_batches || !batch_unmapped[b])
        return;
    const uint32_t base = b * prm.B, ng = batches[b].num_block_groups;
    uint32_t ip_sum = 0, dup_sum = 0;
    for (uint32_t g = 0; g < ng; g++) {
        const uvmtier_fault_group_t grp = groups[base + g];
        if (grp.flags & 1u)
            continue;
        const uint32_t va_space = tbl[grp.ctx].va_space;
        if (svc_slot(prm, va_space, grp.block_start) != 0xffffffffu)
            continue;
        for (uint32_t i = 0; i < grp.block_faults; i++) {
            const uint32_t s = grp.first_fault_index + i;
            const uvmtier_ordered_fault_t e = ordered[base + s];
            const uint32_t acc = e.fault_access_type & 7u;
            uint32_t fl = 0;
            bool dup = false;
            if (s > 0 && !prm.non_replayable) {                           // (non-replayable faults know no duplicates)
                const uvmtier_ordered_fault_t pe = ordered[base + s - 1];
                dup = tbl[pe.ctx].va_space == va_space && pe.fault_address == e.fault_address;
            }
            if (dup && (entry_flags[base + s - 1] & UVMTIER_SF_INVALID_PREFETCH)) {
                fl |= UVMTIER_SF_INVALID_PREFETCH;
                ip_sum += e.num_instances;
            }
            if (dup)
                fl |= UVMTIER_SF_DUPLICATE;
            dup_sum += dup ? e.num_instances : (uint32_t)e.num_instances - 1u;   // update_batch_and_notify_fault (:1692, :1200-1203)
            if (acc == ACC_PREFETCH) {
                fl |= UVMTIER_SF_INVALID_PREFETCH;
                ip_sum += e.num_instances;
            }
            else {
                fl |= UVMTIER_SF_FATAL | (REASON_INVALID_ADDRESS << UVMTIER_SF_REASON_SHIFT) | (CANCEL_ALL << UVMTIER_SF_CANCEL_SHIFT);
            }
            entry_flags[base + s] = (uint16_t)fl;
        }
        uvmtier_service_group_t rec;
        rec.slot = 0xffffffffu; rec.region_first = 0; rec.region_outer = 0; rec.page_fault_count = 0; rec.prefetch_count = 0;
        rec.migrated[0] = 0; rec.migrated[1] = 0; rec.first_touch = 0; rec.hint_residency = UVMTIER_ID_INVALID;
        rec.run_offset = 0; rec.run_count = 0; rec.batch = (uint16_t)(b + prm.batch_base); rec.group = g;
        for (int k = 0; k < 8; k++) rec.new_residency.bitmap[k] = 0;
        service[base + g] = rec;
    }
    if (ip_sum)
        atomicAdd(&batches[b].num_invalid_prefetch_faults, ip_sum);
    if (dup_sum)
        atomicAdd(&batches[b].num_duplicate_faults, dup_sum);
}

} // namespace uvmtier_k
