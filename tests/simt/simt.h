// simt.h -- a tiny CPU SIMT emulator used ONLY by the "not gpu" tests.
//
// It compiles the product's device source (open-gpu-kernel-modules_b200/csrc/*.cuh) with g++ and
// runs each CUDA thread of a CTA as a pthread, so that warp shuffles/ballots, __syncthreads,
// shared memory and atomics behave as on the GPU.  CTAs of a grid run one after the other.
// It exists because this container has no GPU: it lets the CPU suite check the kernels' LOGIC
// against the oracle before a B200 run.  It is test infrastructure: it is never linked into
// libuvmtier.so and is no fallback -- the product fails loudly without a GPU.
//
// Limits: kernels must not wait on other CTAs (they are serialised); every lane named in the mask of a
// *_sync intrinsic must reach it, as the product code guarantees.
#pragma once
#define UVMTIER_SIMT_EMU 1

#include <pthread.h>
#include <sched.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <functional>
#include <vector>

struct dim3 {
    unsigned x, y, z;
    dim3(unsigned x_ = 1, unsigned y_ = 1, unsigned z_ = 1) : x(x_), y(y_), z(z_) {}
};
struct uint2 { uint32_t x, y; };
struct alignas(16) uint4 { uint32_t x, y, z, w; };
struct alignas(16) ulonglong2 { unsigned long long x, y; };
static inline uint4 make_uint4(uint32_t a, uint32_t b, uint32_t c, uint32_t d) { return uint4{a, b, c, d}; }
static inline uint2 make_uint2(uint32_t a, uint32_t b) { return uint2{a, b}; }

#define __device__
#define __global__ static    /* kernels live in headers that several harness files include */
#define __host__
#define __forceinline__ inline
#define __launch_bounds__(...)
#define __shared__ static
#define __constant__ static
#ifndef __restrict__
#define __restrict__
#endif

namespace simt {
struct Warp {
    uint64_t slot[32];
    pthread_barrier_t bar;
    uint32_t m_arrive[32], m_gen[32];    // partial-mask barriers, indexed by the lowest lane of the mask
};
struct Cta {
    pthread_barrier_t bar;
    uint64_t red;          // scratch for __syncthreads_or / _count
    pthread_mutex_t mu;
};
extern thread_local Warp *tl_warp;
extern thread_local Cta *tl_cta;
extern thread_local int tl_lane;
extern char g_dynamic_smem[256 * 1024];
// Runs kernel body `fn` for every thread of every CTA of the grid (CTAs sequentially).
void launch(dim3 grid, dim3 block, const std::function<void()> &fn);
} // namespace simt

extern thread_local dim3 threadIdx, blockIdx;
extern dim3 blockDim, gridDim;
static const int warpSize = 32;

// ---------------------------------------------------------------- warp intrinsics
// Every *_sync intrinsic takes its member mask seriously: the lanes named in the mask (and only they) meet at the
// call.  Full masks use the warp's pthread barrier; partial masks (sub-warp groups that have diverged from the rest
// of the warp) use a small spinning barrier keyed by the lowest lane of the mask, so a warp must always be split
// into the same groups while it uses partial masks.
namespace simt {
template <class T> static inline uint64_t pack(T v) { uint64_t u = 0; memcpy(&u, &v, sizeof(T)); return u; }
template <class T> static inline T unpack(uint64_t u) { T v; memcpy(&v, &u, sizeof(T)); return v; }
static inline void wbar() { pthread_barrier_wait(&tl_warp->bar); }
static inline void mbar(unsigned mask)
{
    if (mask == 0xffffffffu) {
        wbar();
        return;
    }
    const int leader = __builtin_ctz(mask), n = __builtin_popcount(mask);
    if (n <= 1)
        return;
    Warp *w = tl_warp;
    const uint32_t g = __atomic_load_n(&w->m_gen[leader], __ATOMIC_ACQUIRE);
    if (__atomic_add_fetch(&w->m_arrive[leader], 1u, __ATOMIC_ACQ_REL) == (uint32_t)n) {
        __atomic_store_n(&w->m_arrive[leader], 0u, __ATOMIC_RELAXED);
        __atomic_add_fetch(&w->m_gen[leader], 1u, __ATOMIC_RELEASE);
    }
    else {
        while (__atomic_load_n(&w->m_gen[leader], __ATOMIC_ACQUIRE) == g)
            sched_yield();
    }
}
template <class T> static inline T exchange(unsigned mask, T v, int src_lane)
{
    static_assert(sizeof(T) <= 8, "shuffle payload");
    tl_warp->slot[tl_lane] = pack(v);
    mbar(mask);
    T r = unpack<T>(tl_warp->slot[src_lane]);
    mbar(mask);
    return r;
}
template <class F> static inline unsigned reduce(unsigned mask, unsigned v, F f)
{
    tl_warp->slot[tl_lane] = v;
    mbar(mask);
    unsigned r = 0;
    bool first = true;
    for (int i = 0; i < 32; i++)
        if ((mask >> i) & 1u) {
            r = first ? (unsigned)tl_warp->slot[i] : f(r, (unsigned)tl_warp->slot[i]);
            first = false;
        }
    mbar(mask);
    return r;
}
} // namespace simt

template <class T> static inline T __shfl_sync(unsigned mask, T v, int src, int width = 32)
{
    int base = simt::tl_lane & ~(width - 1);
    return simt::exchange(mask, v, base | (src & (width - 1)));
}
template <class T> static inline T __shfl_up_sync(unsigned mask, T v, unsigned delta, int width = 32)
{
    int base = simt::tl_lane & ~(width - 1);
    int src = simt::tl_lane - (int)delta;
    return simt::exchange(mask, v, src < base ? simt::tl_lane : src);
}
template <class T> static inline T __shfl_down_sync(unsigned mask, T v, unsigned delta, int width = 32)
{
    int base = simt::tl_lane & ~(width - 1);
    int src = simt::tl_lane + (int)delta;
    return simt::exchange(mask, v, src >= base + width ? simt::tl_lane : src);
}
template <class T> static inline T __shfl_xor_sync(unsigned mask, T v, int m, int width = 32)
{
    int base = simt::tl_lane & ~(width - 1);
    int src = simt::tl_lane ^ m;
    return simt::exchange(mask, v, src >= base + width ? simt::tl_lane : src);
}
static inline unsigned __ballot_sync(unsigned mask, int pred)
{
    simt::tl_warp->slot[simt::tl_lane] = pred ? 1 : 0;
    simt::mbar(mask);
    unsigned b = 0;
    for (int i = 0; i < 32; i++)
        if ((mask >> i) & 1u)
            b |= (unsigned)(simt::tl_warp->slot[i] & 1) << i;
    simt::mbar(mask);
    return b;
}
static inline int __any_sync(unsigned m, int pred) { return __ballot_sync(m, pred) != 0; }
static inline int __all_sync(unsigned m, int pred) { return __ballot_sync(m, pred) == m; }
template <class T> static inline unsigned __match_any_sync(unsigned mask, T v)
{
    simt::tl_warp->slot[simt::tl_lane] = simt::pack(v);
    simt::mbar(mask);
    unsigned b = 0;
    for (int i = 0; i < 32; i++)
        if (((mask >> i) & 1u) && simt::tl_warp->slot[i] == simt::pack(v))
            b |= 1u << i;
    simt::mbar(mask);
    return b;
}
static inline unsigned __reduce_add_sync(unsigned m, unsigned v) { return simt::reduce(m, v, [](unsigned a, unsigned b) { return a + b; }); }
static inline unsigned __reduce_or_sync(unsigned m, unsigned v) { return simt::reduce(m, v, [](unsigned a, unsigned b) { return a | b; }); }
static inline unsigned __reduce_and_sync(unsigned m, unsigned v) { return simt::reduce(m, v, [](unsigned a, unsigned b) { return a & b; }); }
static inline unsigned __reduce_min_sync(unsigned m, unsigned v) { return simt::reduce(m, v, [](unsigned a, unsigned b) { return a < b ? a : b; }); }
static inline unsigned __reduce_max_sync(unsigned m, unsigned v) { return simt::reduce(m, v, [](unsigned a, unsigned b) { return a > b ? a : b; }); }
static inline void __syncwarp(unsigned m = 0xffffffffu) { simt::mbar(m); }
static inline unsigned __activemask() { return 0xffffffffu; }

// ---------------------------------------------------------------- CTA barriers
static inline void __syncthreads() { pthread_barrier_wait(&simt::tl_cta->bar); }
static inline int __syncthreads_or(int pred)
{
    __syncthreads();
    if (threadIdx.x == 0) simt::tl_cta->red = 0;
    __syncthreads();
    if (pred) __atomic_fetch_or(&simt::tl_cta->red, 1, __ATOMIC_SEQ_CST);
    __syncthreads();
    return simt::tl_cta->red != 0;
}
static inline int __syncthreads_count(int pred)
{
    __syncthreads();
    if (threadIdx.x == 0) simt::tl_cta->red = 0;
    __syncthreads();
    if (pred) __atomic_fetch_add(&simt::tl_cta->red, 1, __ATOMIC_SEQ_CST);
    __syncthreads();
    return (int)simt::tl_cta->red;
}
static inline void __threadfence() { __sync_synchronize(); }
static inline void __threadfence_block() { __sync_synchronize(); }
static inline void __threadfence_system() { __sync_synchronize(); }

// ---------------------------------------------------------------- integer intrinsics
static inline int __popc(unsigned v) { return __builtin_popcount(v); }
static inline int __popcll(unsigned long long v) { return __builtin_popcountll(v); }
static inline int __clz(int v) { return v ? __builtin_clz((unsigned)v) : 32; }
static inline int __clzll(long long v) { return v ? __builtin_clzll((unsigned long long)v) : 64; }
static inline int __ffs(int v) { return __builtin_ffs(v); }
static inline int __ffsll(long long v) { return __builtin_ffsll(v); }
static inline unsigned __brev(unsigned v)
{
    unsigned r = 0;
    for (int i = 0; i < 32; i++) r |= ((v >> i) & 1u) << (31 - i);
    return r;
}
static inline unsigned __funnelshift_r(unsigned lo, unsigned hi, unsigned s)
{
    uint64_t v = ((uint64_t)hi << 32) | lo;
    return (unsigned)(v >> (s & 31));
}
static inline unsigned __funnelshift_l(unsigned lo, unsigned hi, unsigned s)
{
    uint64_t v = ((uint64_t)hi << 32) | lo;
    return (unsigned)((v << (s & 31)) >> 32);
}
using std::max;
using std::min;
static inline int max(int a, unsigned b) { return a > (int)b ? a : (int)b; }
template <class T> static inline T __ldg(const T *p) { return *p; }

// ---------------------------------------------------------------- atomics
template <class T> static inline T atomicAdd(T *p, T v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
template <class T> static inline T atomicSub(T *p, T v) { return __atomic_fetch_sub(p, v, __ATOMIC_SEQ_CST); }
template <class T> static inline T atomicOr(T *p, T v) { return __atomic_fetch_or(p, v, __ATOMIC_SEQ_CST); }
template <class T> static inline T atomicAnd(T *p, T v) { return __atomic_fetch_and(p, v, __ATOMIC_SEQ_CST); }
template <class T> static inline T atomicExch(T *p, T v) { return __atomic_exchange_n(p, v, __ATOMIC_SEQ_CST); }
template <class T> static inline T atomicCAS(T *p, T cmp, T v)
{
    __atomic_compare_exchange_n(p, &cmp, v, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST);
    return cmp;
}
template <class T> static inline T atomicMax(T *p, T v)
{
    T old = __atomic_load_n(p, __ATOMIC_SEQ_CST);
    while (old < v && !__atomic_compare_exchange_n(p, &old, v, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST)) {}
    return old;
}
template <class T> static inline T atomicMin(T *p, T v)
{
    T old = __atomic_load_n(p, __ATOMIC_SEQ_CST);
    while (old > v && !__atomic_compare_exchange_n(p, &old, v, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST)) {}
    return old;
}
