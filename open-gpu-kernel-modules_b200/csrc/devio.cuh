// devio.cuh -- global-memory access helpers with explicit cache policy (inline PTX).
// When compiled for the CPU SIMT emulator (tests/simt, UVMTIER_SIMT_EMU) they degrade to plain
// loads/stores so that the same kernel source runs under the emulator.
#pragma once

#include <stdint.h>

// A kernel defined in a header that more than one .cu file includes: internal linkage.
#ifdef UVMTIER_SIMT_EMU
#define UVMTIER_KERNEL static
#else
#define UVMTIER_KERNEL static __global__
#endif

// Dynamic shared memory of the running CTA (the emulator gives every CTA the same static buffer).
#ifdef UVMTIER_SIMT_EMU
#define UVMTIER_DYN_SMEM(name) unsigned char *name = (unsigned char *)simt::g_dynamic_smem
#else
#define UVMTIER_DYN_SMEM(name) extern __shared__ __align__(16) unsigned char name[]
#endif

#ifdef UVMTIER_SIMT_EMU

static inline uint32_t ld_stream(const uint32_t *p) { return *p; }
static inline uint4 ld_stream_v4(const uint4 *p) { return *p; }
static inline void st_stream_v4(uint4 *p, uint4 v) { *p = v; }
static inline void st_stream(uint32_t *p, uint32_t v) { *p = v; }

#else

// streaming (read-once) loads: read-only path, do not allocate in L1
__device__ __forceinline__ uint32_t ld_stream(const uint32_t *p)
{
    uint32_t v;
    asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ uint4 ld_stream_v4(const uint4 *p)
{
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ void st_stream_v4(uint4 *p, uint4 v)
{
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void st_stream(uint32_t *p, uint32_t v)
{
    asm volatile("st.global.L1::no_allocate.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}

#endif
