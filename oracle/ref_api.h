/*
 * ref_api.h -- flat C harness types shared by the two CPU checkers:
 *   oracle/_ref/libref_*.so  (the reference's own C compiled against shims; symbols ref_*)
 *   oracle/liboracle.so      (the CPU restatement; symbols oracle_*)
 * TEST INFRASTRUCTURE ONLY.  Layouts are byte-identical to the product structs in
 * include/uvmtier.h so one set of numpy arrays drives all three implementations.
 */
#ifndef ORACLE_REF_API_H
#define ORACLE_REF_API_H

#include <stddef.h>
#include <stdint.h>

/* flags of ref_prefetch_block_t / uvmtier_prefetch_block_t */
#define REF_PF_HAS_RESIDENT_MASK 0x01      /* va_block->gpus[dst] != NULL (uvm_perf_prefetch.c:357) */
#define REF_PF_BLOCK_RESIDENT_EMPTY 0x02   /* uvm_processor_mask_empty(&va_block->resident) (:363) */
#define REF_PF_HAS_THRASHING 0x04          /* uvm_perf_thrashing_get_thrashing_pages() != NULL */
#define REF_PF_GPU_VA_SPACE_REGISTERED 0x08 /* dst GPU has a registered gpu_va_space (:261) */

typedef struct {
    uint64_t start;              /* va_block->start (page aligned)                         */
    uint32_t big_page_size;      /* uvm_va_block_gpu_big_page_size(): 65536 or 131072      */
    uint16_t num_pages;          /* uvm_va_block_num_cpu_pages()                           */
    uint16_t faulted_first;      /* faulted_region.first                                   */
    uint16_t faulted_outer;      /* faulted_region.outer                                   */
    uint16_t new_residency;      /* uvm_processor_id_t value: 0 CPU, 1.. GPUs              */
    uint16_t preferred_location; /* policy->preferred_location value, 257 = UVM_ID_INVALID */
    uint16_t flags;              /* REF_PF_*                                               */
} ref_prefetch_block_t;          /* 24 bytes */

typedef struct {
    uint16_t last_migration_proc_id;        /* va_block->prefetch_info (uvm_va_block.h:478-483) */
    uint16_t fault_migrations_to_last_proc; /* NvU16 in the reference: wraps at 65536           */
} ref_prefetch_info_t;

#endif
