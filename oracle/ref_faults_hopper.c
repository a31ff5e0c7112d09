/*
 * ref_faults_hopper.c -- TEST INFRASTRUCTURE ONLY (oracle/_ref build, TU 3 of 3).
 * Compiles kernel-open/nvidia-uvm/uvm_hopper_fault_buffer.c UNMODIFIED: its get_ve_id is what
 * Blackwell inherits (uvm_hal.c:520-526).
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

#include "uvm_hopper_fault_buffer.c"
