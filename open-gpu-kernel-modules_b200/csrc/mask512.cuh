// mask512.cuh -- a 512-bit uvm_page_mask_t spread over a half-warp: lane l (0..15) of the group
// holds 32-bit word l (bits 32l .. 32l+31).  Two va_blocks per warp.  All helpers are
// warp-synchronous and must be called by all 32 lanes; shuffles use width 16 so the two groups
// never mix.
#pragma once

#include <stdint.h>

#define HW_FULL 0xffffffffu

__device__ __forceinline__ uint32_t hw_ones(unsigned n) { return n >= 32 ? 0xffffffffu : ((1u << n) - 1u); }

// This lane's word of the mask with bits [first, outer) set.
__device__ __forceinline__ uint32_t hw_range(int first, int outer, int l)
{
    int lo = first - 32 * l, hi = outer - 32 * l;
    lo = lo < 0 ? 0 : (lo > 32 ? 32 : lo);
    hi = hi < 0 ? 0 : (hi > 32 ? 32 : hi);
    return hi > lo ? (hw_ones(hi) & ~hw_ones(lo)) : 0u;
}

// 512-bit logical shift right by k (0..511): result bit i = input bit i+k.
__device__ __forceinline__ uint32_t hw_shr(uint32_t w, unsigned k, int l)
{
    unsigned q = k >> 5, r = k & 31;
    uint32_t lo = __shfl_down_sync(HW_FULL, w, q, 16);
    uint32_t hi = __shfl_down_sync(HW_FULL, w, q + 1, 16);
    if (l + (int)q > 15) lo = 0;
    if (l + (int)q + 1 > 15) hi = 0;
    return __funnelshift_r(lo, hi, r);
}

// 512-bit logical shift left by k (0..511): result bit i = input bit i-k.
__device__ __forceinline__ uint32_t hw_shl(uint32_t w, unsigned k, int l)
{
    unsigned q = k >> 5, r = k & 31;
    uint32_t hi = __shfl_up_sync(HW_FULL, w, q, 16);
    uint32_t lo = __shfl_up_sync(HW_FULL, w, q + 1, 16);
    if (l - (int)q < 0) hi = 0;
    if (l - (int)q - 1 < 0) lo = 0;
    return __funnelshift_l(lo, hi, r);
}

// Sum / OR / max of a per-lane value over the 16 lanes of the group (result in every lane).
__device__ __forceinline__ uint32_t hw_sum(uint32_t v)
{
    v += __shfl_xor_sync(HW_FULL, v, 1);
    v += __shfl_xor_sync(HW_FULL, v, 2);
    v += __shfl_xor_sync(HW_FULL, v, 4);
    v += __shfl_xor_sync(HW_FULL, v, 8);
    return v;
}
__device__ __forceinline__ uint32_t hw_weight(uint32_t w) { return hw_sum(__popc(w)); }

// true in every lane of the group iff any lane of the group has a non-zero word
__device__ __forceinline__ bool hw_any(uint32_t w, int lane)
{
    uint32_t b = __ballot_sync(HW_FULL, w != 0);
    return ((b >> (lane & 16)) & 0xffffu) != 0;
}

// Index of the first / last set bit of the 512-bit mask; 512 if empty (first) / -1 if empty (last).
__device__ __forceinline__ int hw_first_bit(uint32_t w, int lane)
{
    uint32_t b = (__ballot_sync(HW_FULL, w != 0) >> (lane & 16)) & 0xffffu;
    int src = b ? __ffs(b) - 1 : 0;
    uint32_t ww = __shfl_sync(HW_FULL, w, src, 16);   // both groups always shuffle: no divergence
    return b ? 32 * src + __ffs(ww) - 1 : 512;
}
__device__ __forceinline__ int hw_last_bit(uint32_t w, int lane)
{
    uint32_t b = (__ballot_sync(HW_FULL, w != 0) >> (lane & 16)) & 0xffffu;
    int src = b ? 31 - __clz(b) : 0;
    uint32_t ww = __shfl_sync(HW_FULL, w, src, 16);
    return b ? 32 * src + 31 - __clz(ww) : -1;
}
