// prefetch_kernels.cuh -- K2 kernels (device code only; host entry points are in prefetch.cu).
// Kept in a header so that tests/simt can compile the very same source for the CPU SIMT emulator.
#pragma once

#include "uvmtier.h"
#include "devio.cuh"
#include "prefetch.cuh"

namespace uvmtier_k {


constexpr int kPfThreads = 256;

struct pf_tunables {
    uint32_t enable, threshold, min_faults;
};


__global__ void __launch_bounds__(kPfThreads)
k_prefetch_compute_ats(size_t n, const uint32_t *__restrict__ faulted, const uvmtier_va_block_region_t *__restrict__ fregion,
                       const uvmtier_va_block_region_t *__restrict__ mregion, const uint32_t *__restrict__ residency,
                       uint32_t *__restrict__ out, pf_tunables tun)
{
    const int lane = threadIdx.x & 31, l = lane & 15, grp = lane >> 4;
    const size_t warps = (size_t)gridDim.x * (kPfThreads / 32);
    for (size_t pair = (size_t)blockIdx.x * (kPfThreads / 32) + (threadIdx.x >> 5); pair * 2 < n; pair += warps) {
        const size_t i = pair * 2 + grp;
        const bool live = i < n;
        uint32_t f = 0, r = 0;
        int ff = 0, fo = 0, mf = 0, mo = 0;
        if (live) {
            f = ld_stream(faulted + i * 16 + l);
            if (residency)
                r = ld_stream(residency + i * 16 + l);
            const uvmtier_va_block_region_t fr = fregion[i], mr = mregion[i];
            ff = fr.first; fo = fr.outer; mf = mr.first; mo = mr.outer;
        }
        uint32_t o = 0;
        if (tun.enable) {   // uniform
            // init_bitmap_tree_from_region (:222-238): pages = (residency | faulted) >> max.first
            const uint32_t pages_t = hw_shr(r | f, mf, l);
            const uint32_t fwork = f & hw_range(ff, fo, l) & hw_range(mf, mo, l);
            o = pf_prefetch_mask(pages_t, fwork, mf, mo, 0, tun.threshold, l);
        }
        if (live)
            out[i * 16 + l] = o;
    }
}

__global__ void __launch_bounds__(kPfThreads)
k_prefetch_get_hint(size_t n, const uvmtier_prefetch_block_t *__restrict__ blocks, const uint32_t *__restrict__ resident,
                    const uint32_t *__restrict__ faulted, const uint32_t *__restrict__ thrashing,
                    const uint32_t *__restrict__ cpu_read, uvmtier_prefetch_info_t *__restrict__ info,
                    uint32_t *__restrict__ out, uint32_t *__restrict__ out_residency, pf_tunables tun)
{
    const int lane = threadIdx.x & 31, l = lane & 15, grp = lane >> 4;
    const size_t warps = (size_t)gridDim.x * (kPfThreads / 32);
    for (size_t pair = (size_t)blockIdx.x * (kPfThreads / 32) + (threadIdx.x >> 5); pair * 2 < n; pair += warps) {
        const size_t i = pair * 2 + grp;
        const bool live = i < n;

        uint32_t f = 0, res = 0, thr = 0, mapped = 0;
        uint64_t start = 0;
        int num_pages = 0, ff = 0, fo = 0;
        uint32_t dst = 0, pref = UVMTIER_ID_INVALID, flags = 0, big_pages = 16;
        uvmtier_prefetch_info_t pi = {UVMTIER_ID_INVALID, 0};
        if (live) {
            const uvmtier_prefetch_block_t b = blocks[i];
            start = b.start; num_pages = b.num_pages; ff = b.faulted_first; fo = b.faulted_outer;
            dst = b.new_residency; pref = b.preferred_location; flags = b.flags;
            // big page size of the destination: the GPU's if it has a registered VA space, else 64K (:258-270)
            if (dst != UVMTIER_ID_CPU && (flags & UVMTIER_PF_GPU_VA_SPACE_REGISTERED))
                big_pages = b.big_page_size / UVMTIER_PAGE_SIZE;
            f = ld_stream(faulted + i * 16 + l);
            if (dst == UVMTIER_ID_CPU || (flags & UVMTIER_PF_HAS_RESIDENT_MASK))
                res = ld_stream(resident + i * 16 + l);
            if (thrashing && (flags & UVMTIER_PF_HAS_THRASHING))
                thr = ld_stream(thrashing + i * 16 + l);
            if (dst == UVMTIER_ID_CPU && cpu_read)
                mapped = ld_stream(cpu_read + i * 16 + l);
            pi = info[i];
        }
        const bool have_thr = thrashing && (flags & UVMTIER_PF_HAS_THRASHING);

        uint32_t o = 0;
        uint32_t hint = UVMTIER_ID_INVALID;
        if (tun.enable)   // uniform
            o = pf_get_hint_block(f, res, thr, mapped, have_thr, start, num_pages, ff, fo, dst, pref, flags, big_pages,
                                  tun.threshold, tun.min_faults, &pi, &hint, lane);
        if (live) {
            out[i * 16 + l] = o;
            if (l == 0) {
                out_residency[i] = hint;
                info[i] = pi;
            }
        }
    }
}

} // namespace uvmtier_k
