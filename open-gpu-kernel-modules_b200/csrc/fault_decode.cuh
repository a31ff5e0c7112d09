// fault_decode.cuh -- F1: decode of one 32-byte NVC369_BUF_ENTRY fault packet (UVM/clc369.h:33-66)
// into the fields of uvm_fault_buffer_entry_t, for a Blackwell parent GPU.
// Reference: parse_fault_entry_common (UVM/uvm_volta_fault_buffer.c:312-391) with the HAL overrides
// a Blackwell GPU installs (UVM/uvm_hal.c:520-534): Blackwell get_fault_type / get_mmu_engine_type
// (UVM/uvm_blackwell_fault_buffer.c:34-122), Hopper get_ve_id (UVM/uvm_hopper_fault_buffer.c:31-44),
// Blackwell mmu_client_id_to_utlb_id (UVM/uvm_blackwell_mmu.c:128-201), and
// uvm_parent_gpu_canonical_address (UVM/uvm_gpu.c:428-459).
#pragma once

#include <stdint.h>

#include "uvmtier.h"

namespace uvmtier_k {

// uvm_fault_access_type_t, UVM/uvm_hal_types.h:204-212 (ordered by intrusiveness)
enum : uint32_t { ACC_PREFETCH = 0, ACC_READ = 1, ACC_WRITE = 2, ACC_ATOMIC_WEAK = 3, ACC_ATOMIC_STRONG = 4, ACC_COUNT = 5 };
// uvm_aperture_t, UVM/uvm_hal_types.h:33-51
enum : uint32_t { APERTURE_SYS = 9, APERTURE_VID = 10, APERTURE_MAX = 12 };
// uvm_fault_type_t, UVM/uvm_hal_types.h:275-305
enum : uint32_t { FAULT_TYPE_FATAL = 5, FAULT_TYPE_COUNT = 17 };
// uvm_mmu_engine_type_t, UVM/uvm_hal_types.h:319-325
enum : uint32_t { ENGINE_GRAPHICS = 0, ENGINE_HOST = 1, ENGINE_CE = 2 };

struct fault_fields {
    uint64_t fault_address;     // canonical, 4K aligned
    uint64_t timestamp;
    uint64_t inst_address;      // instance_ptr.address
    uint32_t inst_aperture;
    uint32_t fault_type, access_type, client_type, client_id, gpc_id, utlb_id, engine_id, engine_type, ve_id;
    uint32_t flags;             // UVMTIER_FE_*
    uint32_t valid;
};

// hw NV_PFAULT_FAULT_TYPE_* (hwref/blackwell/gb100/dev_fault.h:269-285) -> uvm_fault_type_t, one
// nibble-pair per hw value packed in two 64-bit constants (values up to 16 need 5 bits: use bytes).
__device__ __forceinline__ uint32_t fd_fault_type(uint32_t hw)
{
    // index:            0  1  2  3  4  5  6  7   8   9  10  11  12  13  14  15
    // uvm_fault_type:   0  5  1  6  7  8  3  4   9  10  11  16  13  14  15   2
    const uint64_t lo = 0x0403080706010500ull, hi = 0x020f0e0d100b0a09ull;
    if (hw >= 16)
        return FAULT_TYPE_COUNT;
    return (uint32_t)(((hw < 8 ? lo : hi) >> ((hw & 7) * 8)) & 0xff);
}

// get_fault_access_type + is_fault_address_virtual (UVM/uvm_volta_fault_buffer.c:105-152);
// hw values hwref/volta/gv100/dev_fault.h:246-255.
__device__ __forceinline__ uint32_t fd_access_type(uint32_t hw, bool *is_virtual)
{
    *is_virtual = hw < 8 || hw > 11;   // invalid encodings: the reference falls out of the switch with "true"
    switch (hw) {
        case 0: case 8: return ACC_READ;
        case 1: case 9: return ACC_WRITE;
        case 2: case 10: return ACC_ATOMIC_STRONG;
        case 3: case 11: return ACC_PREFETCH;
        case 4: return ACC_ATOMIC_WEAK;
    }
    return ACC_COUNT;
}

// UVM/uvm_blackwell_mmu.c:128-201, client ids hwref/blackwell/gb100/dev_fault.h:287-333
__device__ __forceinline__ uint32_t fd_client_to_utlb(uint32_t c)
{
    if (c <= 0x07) return 1 + c;                            // T1_0..7    -> LTP0..7
    if (c <= 0x0f) return 2 + 2 * (c - 0x08);               // PE_0..7    -> LTP1,3,..,15
    if (c <= 0x12) return 0;                                // RAST, GCC, GPCCS -> RGG
    if (c >= 0x21 && c <= 0x28) return 9 + (c - 0x21);      // T1_8..15   -> LTP8..15
    if (c >= 0x29 && c <= 0x30) return 2 + 2 * (c - 0x29);  // TPCCS_0..7 -> LTP1,3,..,15
    if (c == 0x31 || c == 0x32) return 18 + 2 * (c - 0x31); // PE_8,9     -> LTP17,19
    if (c == 0x33 || c == 0x34) return 18 + 2 * (c - 0x33); // TPCCS_8,9  -> LTP17,19
    if (c >= 0x35 && c <= 0x38) return 17 + (c - 0x35);     // T1_16..19  -> LTP16..19
    return 0;                                               // the reference asserts, then returns 0
}

// UVM/uvm_blackwell_fault_buffer.c:78-122
__device__ __forceinline__ uint32_t fd_engine_type(uint32_t hub, uint32_t client)
{
    if (hub) {
        if ((client >= 0x20 && client <= 0x27) || client == 0x01 || client == 0x02 || client == 0x18 || client == 0x0c)
            return ENGINE_CE;
        if (client == 0x06 || (client >= 0x63 && client <= 0x6e))
            return ENGINE_HOST;
    }
    return ENGINE_GRAPHICS;
}

// w0..w7 = the packet's eight 32-bit words.
__device__ __forceinline__ void fd_decode(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3, uint32_t w4, uint32_t w5,
                                          uint32_t w6, uint32_t w7, uint32_t utlb_per_gpc, uint32_t va_bits, fault_fields *f)
{
    const uint32_t ap = (w0 >> 8) & 3;
    f->inst_address = (((uint64_t)w1 << 20) | (w0 >> 12)) << 12;
    f->inst_aperture = ap == 0 ? APERTURE_VID : (ap >= 2 ? APERTURE_SYS : APERTURE_MAX);
    uint64_t addr = (((uint64_t)w3 << 20) | (w2 >> 12)) << 12;
    addr = (uint64_t)((int64_t)(addr << (64 - va_bits)) >> (64 - va_bits));
    f->fault_address = addr;                                 // already 4K aligned (:919 is a no-op at PAGE_SIZE 4K)
    f->timestamp = ((uint64_t)w5 << 32) | w4;
    f->fault_type = fd_fault_type(w7 & 0x1f);
    bool is_virtual;
    f->access_type = fd_access_type((w7 >> 16) & 0xf, &is_virtual);
    const uint32_t hub = (w7 >> 20) & 1;
    f->client_type = hub;
    f->client_id = (w7 >> 8) & 0x7f;
    f->gpc_id = (w7 >> 24) & 0x1f;
    f->utlb_id = hub ? 0u : f->gpc_id * utlb_per_gpc + fd_client_to_utlb(f->client_id);   // :356-368
    f->engine_id = w6 & 0x1ff;
    f->engine_type = fd_engine_type(hub, f->client_id);
    f->ve_id = f->engine_type == ENGINE_GRAPHICS ? ((f->engine_id - 384u) & 0xffu) : 0u;
    f->flags = ((w7 >> 7) & 1 ? UVMTIER_FE_REPLAYABLE : 0u) | (is_virtual ? UVMTIER_FE_VIRTUAL : 0u) |
               ((w7 >> 29) & 1 ? UVMTIER_FE_PROTECTED : 0u) | (f->fault_type >= FAULT_TYPE_FATAL ? UVMTIER_FE_FATAL : 0u);
    f->valid = w7 >> 31;
}

} // namespace uvmtier_k
