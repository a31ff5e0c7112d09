// resident_kernels.cuh -- uvm_va_block_make_resident on the flat state table, one request per warp.
// Device code only (host side: resident.cu); also compiled by tests/simt for the CPU emulator.
//
// Reference (UVM = kernel-open/nvidia-uvm, uvm_va_block.c): uvm_va_block_make_resident (:5086-5114) =
// _copy (:4877-4976: unmap, populate, block_copy_resident_pages :4660-4847) + _finish (:5042-5084:
// block_make_resident_update_state :4990-5040, block_copy_set_first_touch_residency :4581-4619,
// break_read_duplication_in_region :4538-4579); with UVMTIER_MR_READ_DUPLICATE the COPY variant
// uvm_va_block_make_resident_read_duplicate (:5164-5332, block_prep_read_duplicate_mapping :5120-5162).  Callers: uvm_va_block_migrate_locked (uvm_migrate.c:225-289),
// uvm_va_block_evict_chunks (:12999-13130, cause EVICTION, page_mask = chunk pages & gpu resident), the CPU fault
// handler, and the GPU fault path (which this library runs inside the servicing pass instead).
//
// Processors of the table: this GPU, the CPU with two NUMA nodes (0 = host DRAM, 1 = CXL), and up to four other GPUs
// of the box that hold pages of the same va_blocks (residency only; uvmtier_block_state_add_peer).  CPU pages are
// allocated on the node the request names (block_populate_pages_cpu's first choice, :1887-2018).  A page lives on
// one CPU node at most; it can be on the CPU and on the GPU only while read-duplicated, and is then copied from
// the closest holder (block_copy_resident_pages_mask :4486-4497).
// Source order with other GPUs (block_copy_resident_pages :4707-4829, closest first = uvm_va_space.c:1879-1924):
// destination this GPU: the GPUs it can copy from directly in ascending id, CPU node 0, CPU node 1, then the pages
// only GPUs without direct access hold, staged through CPU pages on node 0 (one run GPU -> DRAM per such GPU, one run
// DRAM -> HBM for their union); destination CPU: the other CPU node, then all GPUs in ascending id, this one among them.
// Masks: one 32-bit word per lane, lanes 16-31 mirror lanes 0-15 (mask512.cuh helpers are half-warp wide).
#pragma once

#include "service_kernels.cuh"

namespace uvmtier_k {

constexpr int kResThreads = 128;

struct resident_kparams {
    uint32_t n_requests, n_slots, run_cap;
    uint32_t gpu_id;                           // uvm_gpu_id_t of this GPU (places it among the other GPUs)
    const uint32_t *n_requests_dev;            // when set: the request count lives on the device (at most n_requests)
};

// the other GPUs of the table (uvmtier_block_state_t.peer_*), ascending gpu id
struct peer_state_ptrs {
    uint32_t *resident[UVMTIER_MAX_PEER_GPUS];
    uint32_t gpu_id[UVMTIER_MAX_PEER_GPUS];
    uint32_t n_peers, direct;
};

constexpr int kResSegs = 8;                    // most copy-list segments of one request (4 GPUs, 2 CPU legs, staging, zero fill)

__global__ void __launch_bounds__(kResThreads)
k_make_resident(resident_kparams prm, const uvmtier_make_resident_req_t *__restrict__ reqs,
                const uint32_t *__restrict__ page_masks, const uint32_t *__restrict__ prefetch_masks, block_state_ptrs st,
                peer_state_ptrs peers, uvmtier_make_resident_result_t *__restrict__ results, uvmtier_copy_run_t *__restrict__ runs,
                uint32_t *__restrict__ counters)
{
    const int lane = threadIdx.x & 31, l = lane & 15;
    const uint32_t n_warps = gridDim.x * (kResThreads / 32);
    const uint32_t n_requests = prm.n_requests_dev ? min(prm.n_requests, *prm.n_requests_dev) : prm.n_requests;
    for (uint32_t q = blockIdx.x * (kResThreads / 32) + (threadIdx.x >> 5); q < n_requests; q += n_warps) {
        const uvmtier_make_resident_req_t rq = reqs[q];
        uint32_t *res = (uint32_t *)&results[q];
        const uint32_t slot = rq.slot;
        const bool dst_gpu = rq.dest_tier == UVMTIER_TIER_HBM;
        const bool rdup = (rq.flags & UVMTIER_MR_READ_DUPLICATE) != 0;
        const bool valid = slot < prm.n_slots && rq.region_first <= rq.region_outer && rq.region_outer <= 512u &&
                           (dst_gpu || rq.dest_tier == UVMTIER_TIER_DRAM || rq.dest_tier == UVMTIER_TIER_CXL) &&
                           !(rdup && (!st.read_duplicated || rq.cause == UVMTIER_CAUSE_EVICTION || peers.n_peers ||
                                      (rq.flags & UVMTIER_MR_EVICT_GPU_PAGES)));   // :5304 asserts cause != EVICTION
        if (!valid) {                                                     // uniform across the warp
            if (lane < 16) {
                if (l < 8) res[l] = l == 0 ? 0xffffffffu : 0u;            // slot 0xFFFFFFFF marks a rejected request
                res[8 + l] = 0;
                res[24 + l] = 0;
            }
            continue;
        }
        const size_t w = (size_t)slot * 16 + l;
        uint32_t gres = st.gpu_resident[w], c0 = st.cpu_resident[0][w], c1 = st.cpu_resident[1][w];
        uint32_t pte0 = st.gpu_pte[0][w], pte1 = st.gpu_pte[1][w], pte2 = st.gpu_pte[2][w];
        uint32_t evicted = st.gpu_evicted ? st.gpu_evicted[w] : 0u;
        uint32_t rd = st.read_duplicated ? st.read_duplicated[w] : 0u;

        // page_mask restricted to the region; NULL page_mask = the whole region
        uint32_t pm = hw_range((int)rq.region_first, (int)rq.region_outer, l);
        if (rq.flags & UVMTIER_MR_HAS_PAGE_MASK)
            pm &= page_masks[(size_t)rq.mask_row * 16 + l];
        if (rq.flags & UVMTIER_MR_EVICT_GPU_PAGES)
            pm &= gres;                                                   // pages_to_evict (:13078)
        const uint32_t pf = (rq.flags & UVMTIER_MR_HAS_PREFETCH_MASK) ? prefetch_masks[(size_t)rq.mask_row * 16 + l] : 0u;

        // block_copy_resident_pages: what is missing at the destination, and where each page comes from
        const int nid = rq.dest_tier == UVMTIER_TIER_CXL ? 1 : 0;
        const uint32_t dst_res = dst_gpu ? gres : (nid ? c1 : c0);
        const uint32_t copy = pm & ~dst_res;                              // :4687-4692
        uint32_t from[3] = {0u, 0u, 0u};                                  // CPU node 0, CPU node 1, this GPU
        uint32_t pres[UVMTIER_MAX_PEER_GPUS], from_peer[UVMTIER_MAX_PEER_GPUS], staged = 0;
#pragma unroll
        for (uint32_t j = 0; j < UVMTIER_MAX_PEER_GPUS; j++) {
            pres[j] = j < peers.n_peers ? peers.resident[j][w] : 0u;
            from_peer[j] = 0;
        }
        const uint32_t dst_code = dst_gpu ? UVMTIER_RUN_DST_HBM : (nid ? UVMTIER_RUN_DST_CXL : UVMTIER_RUN_DST_DRAM);
        // copy list segments in the order the reference visits the sources, the zero fills last
        uint32_t seg_m[kResSegs], seg_src[kResSegs], seg_dst[kResSegs];
        int ns = 0;
#define RES_SEG(mask_, src_, dst_) do { seg_m[ns] = (mask_); seg_src[ns] = (src_); seg_dst[ns] = (dst_); ns++; } while (0)
        uint32_t rest = copy;
        if (dst_gpu) {
#pragma unroll
            for (uint32_t j = 0; j < UVMTIER_MAX_PEER_GPUS; j++)           // GPUs with direct copy access (:4707-4745)
                if (j < peers.n_peers && ((peers.direct >> j) & 1u)) {
                    from_peer[j] = rest & pres[j];
                    rest &= ~from_peer[j];
                    RES_SEG(from_peer[j], UVMTIER_RUN_SRC_PEER0 + j, dst_code);
                }
            from[0] = rest & c0;                                          // CPU nodes in ascending order (:4490)
            rest &= ~from[0];
            RES_SEG(from[0], UVMTIER_RUN_SRC_DRAM, dst_code);
            from[1] = rest & c1;
            rest &= ~from[1];
            RES_SEG(from[1], UVMTIER_RUN_SRC_CXL, dst_code);
            if (peers.n_peers && (peers.direct ^ ((1u << peers.n_peers) - 1u))) {
                // what only GPUs without direct access hold goes through staging pages on CPU node 0
                // (block_populate_pages pages_staged :3140-3170; copies :4754-4785 then :4803-4829)
#pragma unroll
                for (uint32_t j = 0; j < UVMTIER_MAX_PEER_GPUS; j++)
                    if (j < peers.n_peers && !((peers.direct >> j) & 1u)) {
                        const uint32_t hop = rest & pres[j];
                        rest &= ~hop;
                        from_peer[j] = hop;
                        staged |= hop;
                        RES_SEG(hop, UVMTIER_RUN_SRC_PEER0 + j, UVMTIER_RUN_DST_DRAM);
                    }
                RES_SEG(staged, UVMTIER_RUN_SRC_DRAM, dst_code);
            }
        }
        else {
            // the CPU itself is the closest source: the other node first (:4486-4497), then the GPUs in ascending id
            from[nid ^ 1] = rest & (nid ? c0 : c1);
            rest &= ~from[nid ^ 1];
            RES_SEG(from[nid ^ 1], (uint32_t)(nid ^ 1), dst_code);
            bool self_done = false;
#pragma unroll
            for (uint32_t j = 0; j <= UVMTIER_MAX_PEER_GPUS; j++) {
                if (j > peers.n_peers)
                    break;
                if (!self_done && (j == peers.n_peers || peers.gpu_id[j < UVMTIER_MAX_PEER_GPUS ? j : 0] > prm.gpu_id)) {
                    from[2] = rest & gres;
                    rest &= ~from[2];
                    RES_SEG(from[2], UVMTIER_RUN_SRC_HBM, dst_code);
                    self_done = true;
                }
                if (j < peers.n_peers && j < UVMTIER_MAX_PEER_GPUS) {
                    from_peer[j] = rest & pres[j];
                    rest &= ~from_peer[j];
                    RES_SEG(from_peer[j], UVMTIER_RUN_SRC_PEER0 + j, dst_code);
                }
            }
        }
        // pages resident nowhere are populated (and zeroed) at the destination: first touch (:4581-4619)
        const uint32_t touch = rest;
        const uint32_t migrated = copy & ~touch;
        RES_SEG(touch, UVMTIER_RUN_SRC_ZERO, dst_code);
#undef RES_SEG

        uint32_t seg_at[kResSegs], run_cnt = 0;
#pragma unroll
        for (int i = 0; i < kResSegs; i++) {
            seg_at[i] = run_cnt;
            if (i < ns)
                run_cnt += svc_emit_runs(seg_m[i], pf, slot, seg_src[i], NULL, 0u, 0u, lane);
        }
        uint32_t run_off = 0;
        if (lane == 0 && run_cnt)
            run_off = atomicAdd(&counters[0], run_cnt);
        run_off = __shfl_sync(HW_FULL, run_off, 0);
#pragma unroll
        for (int i = 0; i < kResSegs; i++)
            if (i < ns)
                svc_emit_runs(seg_m[i], pf, slot, seg_src[i], runs, run_off + seg_at[i], prm.run_cap, lane, seg_dst[i]);

        const uint32_t now = migrated | touch;
        uint32_t unmap, changed;
        if (rdup) {
            // pages entering read duplication (:5222-5243, block_prep_read_duplicate_mapping): everybody but the holder
            // is unmapped and the holder loses write access.  The only mapper in the table is this GPU: it loses its
            // mappings of such CPU pages and WRITE / ATOMIC on such pages of its own.
            const uint32_t fresh = pm & ~rd, on_cpu = fresh & (c0 | c1), on_gpu = fresh & gres;
            unmap = on_cpu & pte0;
            pte0 &= ~on_cpu; pte1 &= ~(on_cpu | on_gpu); pte2 &= ~(on_cpu | on_gpu);
            // :5259-5307: the destination gains first touch + migrated, the sources keep theirs -- except the other
            // CPU node (one CPU page table: those pages switch node) -- and what migrated is read-duplicated
            if (dst_gpu) {
                gres |= now;
                evicted &= ~migrated;
            }
            else if (nid) { c1 |= now; c0 &= ~migrated; }
            else          { c0 |= now; c1 &= ~migrated; }
            rd |= migrated;
            changed = touch;                                              // only first touch reports (:4614-4617)
        }
        else {
            // make_resident_copy unmaps every processor from the pages that are not resident on the destination
            // (:4919-4926) and every processor but the destination from the read-duplicated pages (:4931-4939); the
            // only mapper in the table is this GPU
            const uint32_t gone = copy | (dst_gpu ? 0u : (pm & rd));
            unmap = gone & pte0;
            pte0 &= ~gone; pte1 &= ~gone; pte2 &= ~gone;
            // _finish: the destination becomes resident (update_state + first touch), everybody else loses the pages
            // of page_mask (break_read_duplication_in_region clears all of region & page_mask, not only what moved)
            rd &= ~pm;
#pragma unroll
            for (uint32_t j = 0; j < UVMTIER_MAX_PEER_GPUS; j++)
                pres[j] &= ~pm;
            if (dst_gpu) {
                gres |= now;
                c0 &= ~pm;
                c1 &= ~pm;
                evicted &= ~migrated;                                     // block_make_resident_clear_evicted :4978-4988
            }
            else {
                if (nid) { c1 |= now; c0 &= ~pm; }
                else     { c0 |= now; c1 &= ~pm; }
                gres &= ~pm;
                // :5023-5037: every GPU that was a source gets ALL the migrated pages marked (copy_mask)
                if (rq.cause == UVMTIER_CAUSE_EVICTION && hw_any(from[2], lane))
                    evicted |= migrated;
            }
            changed = now;                                                // pages_changed_residency (:5005-5008, :4614-4617)
        }

        const uint32_t n0 = hw_weight(from[0]) + hw_weight(staged), n1 = hw_weight(from[1]), n2 = hw_weight(from[2]);
        uint32_t np[UVMTIER_MAX_PEER_GPUS];
#pragma unroll
        for (uint32_t j = 0; j < UVMTIER_MAX_PEER_GPUS; j++)
            np[j] = j < peers.n_peers ? hw_weight(from_peer[j]) : 0u;
        const uint32_t nt = hw_weight(touch), nu = hw_weight(unmap);
        if (lane < 16) {
            st.gpu_resident[w] = gres; st.cpu_resident[0][w] = c0; st.cpu_resident[1][w] = c1;
            st.gpu_pte[0][w] = pte0; st.gpu_pte[1][w] = pte1; st.gpu_pte[2][w] = pte2;
            if (st.gpu_evicted)
                st.gpu_evicted[w] = evicted;
            if (st.read_duplicated)
                st.read_duplicated[w] = rd;
#pragma unroll
            for (uint32_t j = 0; j < UVMTIER_MAX_PEER_GPUS; j++)
                if (j < peers.n_peers)
                    peers.resident[j][w] = pres[j];
            // result: 8 header words, then pages_migrated and pages_changed_residency
            if (l < 8) {
                uint32_t v = 0;
                if (l == 0) v = slot;
                if (l == 1) v = n0 | (n1 << 16);
                if (l == 2) v = n2 | (nt << 16);
                if (l == 3) v = run_off;
                if (l == 4) v = run_cnt | (nu << 16);
                if (l == 5) v = np[0] | (np[1] << 16);                    // migrated_peer[0..3]
                if (l == 6) v = np[2] | (np[3] << 16);
                res[l] = v;
            }
            res[8 + l] = migrated;                                        // make_resident.pages_migrated
            res[24 + l] = changed;
        }
    }
}

} // namespace uvmtier_k
