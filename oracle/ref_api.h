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


/* ---------------- fault leg ---------------- */

/* parsed fault entry + fetch-time coalescing state (view of uvm_fault_buffer_entry_t,
 * uvm_hal_types.h:376-453); same layout as uvmtier_fault_buffer_entry_t */
#define REF_FE_FATAL      0x01
#define REF_FE_REPLAYABLE 0x02
#define REF_FE_VIRTUAL    0x04
#define REF_FE_PROTECTED  0x08
#define REF_FE_FILTERED   0x10
typedef struct {
    uint64_t fault_address;         /* canonical, page aligned                               */
    uint64_t timestamp;
    uint64_t instance_ptr_address;
    uint16_t client_id;
    uint16_t mmu_engine_id;
    uint16_t utlb_id;
    uint16_t num_instances;         /* valid for non-filtered entries                        */
    uint8_t instance_ptr_aperture;  /* uvm_aperture_t                                        */
    uint8_t fault_type;             /* uvm_fault_type_t                                      */
    uint8_t fault_access_type;      /* uvm_fault_access_type_t                               */
    uint8_t client_type;
    uint8_t mmu_engine_type;
    uint8_t gpc_id;
    uint8_t ve_id;
    uint8_t flags;                  /* REF_FE_*                                              */
    uint8_t access_type_mask;       /* valid for non-filtered entries                        */
    uint8_t pad[7];
} ref_fault_entry_t;                /* 48 bytes */

#define REF_VE_ID_ANY 0xFF
/* one row of the instance-pointer translation table (stand-in for
 * uvm_parent_gpu_fault_entry_to_va_space, uvm_gpu.c:3491-3552) */
typedef struct {
    uint64_t instance_ptr;          /* instance_ptr.address (4K aligned)                     */
    uint8_t aperture;               /* uvm_aperture_t value                                  */
    uint8_t ve_id;                  /* subcontext, REF_VE_ID_ANY = any                       */
    uint16_t va_space;              /* index of the uvm_va_space_t (pointer order = index order) */
    uint16_t gpu_id;                /* uvm_gpu_id_t value                                    */
    uint16_t sort_rank;             /* dense rank of (va_space, gpu_id) over the table       */
} ref_instance_ptr_entry_t;         /* 16 bytes */

typedef struct {
    uint32_t max_batch_size;        /* uvm_perf_fault_batch_count                            */
    uint32_t utlb_per_gpc_count;    /* parent_gpu->utlb_per_gpc_count (Blackwell: 2*TPC/GPC+1) */
    uint32_t gpc_count;             /* rm_info.maxGpcCount                                   */
    uint32_t va_bits;               /* canonical address width: 57 (Hopper+), 49 before      */
    uint32_t coalesce;              /* uvm_perf_fault_coalesce                               */
    uint32_t reserved[3];
} ref_fault_gpu_params_t;

typedef struct {
    uint32_t num_cached_faults;
    uint32_t num_coalesced_faults;
    uint32_t is_single_instance_ptr;
    uint32_t max_utlb_id;
    uint32_t status;                /* NV_OK or NV_WARN_MORE_PROCESSING_REQUIRED (0x10005)   */
    uint32_t num_duplicate_faults;
    uint32_t num_invalid_prefetch_faults;
    uint32_t num_block_groups;
} ref_fault_batch_t;                /* 32 bytes */

/* ---------------- service pass (stage C/D/E) ---------------- */

/* one element of the sorted view and one va_block run, same layout as the product's
 * uvmtier_ordered_fault_t / uvmtier_fault_group_t */
typedef struct {
    uint64_t fault_address;
    uint16_t fault_index;
    uint16_t ctx;
    uint16_t num_instances;
    uint8_t access_type_mask;
    uint8_t fault_access_type;      /* low 3 bits; 0x80 = fatal */
} ref_ordered_fault_t;
typedef struct {
    uint64_t block_start;
    uint16_t ctx;
    uint16_t first_fault_index;
    uint16_t block_faults;
    uint16_t flags;
} ref_fault_group_t;

/* VA layout of the harness: va_space s owns [va_base, va_base + blocks_per_space * 2MB); its va_block
 * b has state slot s * blocks_per_space + b (stand-in for the va_range tree) */
typedef struct {
    uint64_t va_base;
    uint32_t blocks_per_space;
    uint32_t n_spaces;
    uint32_t gpu_id;                /* uvm_gpu_id_t value of the faulting GPU (this context) */
    uint32_t reserved;
} ref_va_layout_t;

/* per-slot va_block state, SoA (host pointers here; device pointers in the product) */
typedef struct {
    uint64_t *gpu_resident;         /* gpu_state->resident                 [n_slots][8] */
    uint64_t *cpu_resident[2];      /* cpu node_state->resident, node 0 = DRAM, 1 = CXL */
    uint64_t *gpu_pte[3];           /* gpu_state->pte_bits[READ, WRITE, ATOMIC] */
    ref_prefetch_info_t *prefetch_info;
    uint8_t *logical_prot;          /* compute_logical_prot(): 0 NONE, 1 READ_ONLY, 3 READ_WRITE_ATOMIC */
    uint16_t *preferred_location;   /* policy->preferred_location, 257 = none */
    uint64_t *gpu_evicted;          /* gpu_state->evicted (may be NULL) */
    uint64_t *read_duplicated;      /* va_block->read_duplicated_pages (NULL = read duplication off) */
    /* other GPUs of the box that hold pages of these va_blocks (uvmtier_block_state_t, same layout): their
     * gpu_state->resident masks, ascending uvm_gpu_id_t; bit k of peer_direct = this GPU can copy from peer k directly
     * (block_get_can_copy_from_mask), clear = through the CPU (uvm_va_block.c:3140-3170, :4769-4829) */
    uint64_t *peer_resident[4];
    uint16_t peer_gpu_id[4];
    uint32_t n_peers;
    uint32_t peer_direct;
} ref_block_state_t;

/* uvm_va_block_make_resident request / result, same layout as uvmtier_make_resident_req_t / _result_t */
typedef struct {
    uint32_t slot;
    uint16_t region_first, region_outer;
    uint8_t  dest_tier;             /* 0 HBM (this GPU), 1 CPU node 0 (DRAM), 2 CPU node 1 (CXL) */
    uint8_t  cause;                 /* uvm_make_resident_cause_t; 4 = EVICTION */
    uint16_t flags;                 /* 1 page mask given, 2 prefetch mask given, 4 page_mask = gpu resident pages,
                                       8 uvm_va_block_make_resident_read_duplicate (copy, sources keep their pages) */
    uint32_t mask_row;
} ref_make_resident_req_t;          /* 16 bytes */
typedef struct {
    uint32_t slot;
    uint16_t migrated[3];
    uint16_t first_touch;
    uint32_t run_offset;
    uint16_t run_count;
    uint16_t unmapped;
    uint16_t migrated_peer[4];      /* pages copied from peer GPU k */
    uint32_t reserved;
    uint64_t pages_migrated[8];
    uint64_t pages_changed_residency[8];
} ref_make_resident_result_t;       /* 160 bytes */

#define REF_SF_SERVICED         0x0001
#define REF_SF_DUPLICATE        0x0002
#define REF_SF_FATAL            0x0004
#define REF_SF_INVALID_PREFETCH 0x0008
#define REF_SF_AUTHORIZED       0x0010
#define REF_SF_THROTTLED        0x2000   /* thrashing hint THROTTLE: mark_fault_throttled, not serviced */
#define REF_SF_PINNED           0x4000   /* thrashing hint PIN on the faulting GPU: in thrashing_pin_mask, serviced */
#define REF_SF_PIN_REMOTE       0x8000   /* thrashing hint PIN elsewhere: left to the caller */
#define REF_SF_ACCESS_SHIFT     5       /* 3 bits: service access type */
#define REF_SF_REASON_SHIFT     8       /* 3 bits: UvmEventFatalReason */
#define REF_SF_CANCEL_SHIFT     12      /* 1 bit: uvm_fault_cancel_va_mode_t */

/* what servicing one va_block run did (service_fault_batch_block_locked + uvm_va_block_service_locked) */
typedef struct {
    uint32_t slot;                  /* block state slot, 0xFFFFFFFF = address outside every va_range */
    uint16_t region_first, region_outer;   /* service region after the prefetch hint (:11875) */
    uint16_t page_fault_count;      /* pages set by faults */
    uint16_t prefetch_count;        /* pages added by the hint */
    uint16_t migrated[2];           /* pages copied from CPU node 0 / 1 */
    uint16_t first_touch;           /* pages populated (zeroed) at the destination */
    uint16_t hint_residency;
    uint32_t run_offset;            /* first copy run of this group in the run array */
    uint16_t run_count;
    uint16_t batch;
    uint32_t group;                 /* index of the run inside its batch */
    uint64_t new_residency[8];      /* per_processor_masks[gpu].new_residency after the hint */
} ref_service_group_t;              /* 96 bytes */

/* one contiguous page run to move: block_copy_resident_pages_between coalesces consecutive pages
 * when both sides are block-contiguous (uvm_va_block.c:4278-4288) */
typedef struct {
    uint32_t slot;
    uint16_t first_page;
    uint16_t num_pages : 10;
    uint16_t src : 3;               /* 0 CPU node 0, 1 CPU node 1, 2 this GPU, 7 = zero-fill (first touch) */
    uint16_t prefetch : 1;          /* run cause is PREFETCH (runs also split where the cause changes) */
    uint16_t dst : 2;               /* 0 this GPU, 1 CPU node 0, 2 CPU node 1 */
} ref_copy_run_t;                   /* 8 bytes */

#endif
