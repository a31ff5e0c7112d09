/*
 * ref_faults_blackwell.c -- TEST INFRASTRUCTURE ONLY (oracle/_ref build, TU 2 of 3).
 * Compiles kernel-open/nvidia-uvm/uvm_blackwell_fault_buffer.c UNMODIFIED (Blackwell get_fault_type /
 * get_mmu_engine_type, the overrides installed at uvm_hal.c:528-534) and the client-id -> uTLB map
 * of uvm_blackwell_mmu.c:128-201 (verbatim extract) with the Blackwell dev_fault.h values.
 */
#include "shim/kshim.h"
#include "nvmisc.h"
#include "uvm_types.h"
#include "x_push_macros.h"

typedef struct uvm_va_space_struct uvm_va_space_t;
typedef struct uvm_gpu_struct uvm_gpu_t;
typedef struct uvm_parent_gpu_struct uvm_parent_gpu_t;

#define _UVM_LINUX_H
#define __UVM_COMMON_H__
#define __UVM_GLOBAL_H__
#define __UVM_GPU_H__
#define __UVM_HAL_H__
#define __UVM_HAL_TYPES_H__
#include "x_hal_types.h"

#include "uvm_blackwell_fault_buffer.c"

#include "x_blackwell_utlb_enum.h"
#include "x_blackwell_client_to_utlb.h"
