/*
 * uvmtier.h -- C ABI of libuvmtier: a B200-native (sm_100a) implementation of the data-parallel
 * core of the nvidia-uvm tiering path of CXLMemUring/open-gpu-kernel-modules (575.57.08):
 *
 *   (1) replayable-fault batch preprocessing  (uvm_gpu_replayable_faults.c)
 *   (2) the uvm_perf_prefetch bitmap-tree prefetch hint per 2 MB va_block
 *   (3) va_block make_resident page-list migration between host DRAM, a CXL NUMA node, HBM and
 *       peer GPUs
 *
 * Every entry point cites the reference interface it replaces (paths relative to the reference
 * root; UVM = kernel-open/nvidia-uvm).  Plain C: pointers and sizes only.  Error convention is
 * the reference's NV_STATUS (kernel-open/common/inc/nvstatuscodes.h).  There is no CPU fallback:
 * every compute entry point launches CUDA kernels on the context's device and fails with
 * NV_ERR_GPU_* / NV_ERR_GENERIC when that is impossible.
 *
 * Threading (mirrors UVM/uvm_gpu_isr.c:598-625): one servicing thread per GPU.  Calls on one
 * uvmtier_ctx_t are not re-entrant; different contexts (GPUs) may be driven concurrently.
 *
 * Pointer naming: `d_` = device pointer (HBM of the context's GPU), `h_` = host pointer.
 * Functions ending in _device take device pointers and only enqueue work on the context stream
 * (call uvmtier_synchronize to wait); functions without the suffix take host buffers and include
 * the host<->device copies.
 */
#ifndef UVMTIER_H
#define UVMTIER_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------------------------------ */
/* NV_STATUS subset (kernel-open/common/inc/nvstatuscodes.h:27,59,106,110,172,173)             */
/* ------------------------------------------------------------------------------------------ */
typedef uint32_t NV_STATUS;
#ifndef NV_OK
#define NV_OK                              0x00000000u
#define NV_ERR_GENERIC                     0x0000FFFFu
#define NV_ERR_INVALID_ADDRESS             0x0000001Eu
#define NV_ERR_INVALID_ARGUMENT            0x0000001Fu
#define NV_ERR_INVALID_STATE               0x00000040u
#define NV_ERR_MORE_PROCESSING_REQUIRED    0x0000004Du
#define NV_ERR_NO_MEMORY                   0x00000051u
#define NV_ERR_NOT_SUPPORTED               0x00000056u
#define NV_ERR_INVALID_CHANNEL             0x00000024u
#define NV_ERR_GPU_NOT_FULL_POWER          0x00000011u /* used for "no usable CUDA device" */
#define NV_ERR_GPU_IS_LOST                 0x0000000Fu
#define NV_ERR_INSUFFICIENT_RESOURCES      0x0000001Au
#define NV_WARN_MISMATCHED_TARGET          0x00010004u
#define NV_WARN_MORE_PROCESSING_REQUIRED   0x00010005u
#endif

/* ------------------------------------------------------------------------------------------ */
/* Basic types                                                                                 */
/* ------------------------------------------------------------------------------------------ */
#define UVMTIER_PAGE_SIZE            4096u
#define UVMTIER_VA_BLOCK_SIZE        (2u << 20)   /* UVM_VA_BLOCK_SIZE, UVM/uvm_va_block_types.h:42-45 */
#define UVMTIER_PAGES_PER_VA_BLOCK   512u         /* PAGES_PER_UVM_VA_BLOCK, :50 */
#define UVMTIER_ID_CPU               0u           /* UVM_ID_CPU_VALUE,  UVM/uvm_processors.h:280 */
#define UVMTIER_ID_GPU0              1u           /* UVM_ID_GPU0_VALUE, :281 */
#define UVMTIER_ID_INVALID           257u         /* UVM_ID_INVALID = UVM_ID_MAX_PROCESSORS, :294,300 */

/* uvm_page_mask_t (UVM/uvm_va_block_types.h:77-80): 512 bits, bit p = page p of the va_block.
 * Arrays of masks are dense: mask i occupies 64 bytes at offset 64*i. */
typedef struct { uint64_t bitmap[8]; } uvmtier_page_mask_t;

/* uvm_va_block_region_t (UVM/uvm_va_block_types.h:70-75): [first, outer) */
typedef struct { uint16_t first, outer; } uvmtier_va_block_region_t;

/* va_block->prefetch_info (UVM/uvm_va_block.h:478-483); in/out of the prefetch hint. */
typedef struct {
    uint16_t last_migration_proc_id;
    uint16_t fault_migrations_to_last_proc;   /* NvU16 in the reference: wraps at 65536 */
} uvmtier_prefetch_info_t;

typedef struct uvmtier_ctx uvmtier_ctx_t;

/* ------------------------------------------------------------------------------------------ */
/* Context                                                                                     */
/* ------------------------------------------------------------------------------------------ */

/* Tunables; names, defaults and validation are the reference's module parameters
 * (UVM/uvm_perf_prefetch.c:39-61,504-535; UVM/uvm_gpu_replayable_faults.c:72-115,152-213). */
typedef struct {
    uint32_t uvm_perf_prefetch_enable;      /* default 1 */
    uint32_t uvm_perf_prefetch_threshold;   /* default 51; >100 -> 51 */
    uint32_t uvm_perf_prefetch_min_faults;  /* default 1; outside [1,20] -> 1 */
    uint32_t uvm_perf_fault_batch_count;    /* default 256 */
    uint32_t uvm_perf_fault_coalesce;       /* default 1 */
    uint32_t reserved[11];
} uvmtier_config_t;

/* Fills *cfg with the reference defaults. */
void uvmtier_config_default(uvmtier_config_t *cfg);

/* Creates a context bound to CUDA device `device` (one per GPU, like the reference's per-parent-GPU
 * batch_service_context, UVM/uvm_gpu.h:411).  Applies the reference's tunable validation
 * (uvm_perf_prefetch_init, UVM/uvm_perf_prefetch.c:504-535).  cfg may be NULL (defaults). */
NV_STATUS uvmtier_init(int device, const uvmtier_config_t *cfg, uvmtier_ctx_t **out_ctx);
void      uvmtier_deinit(uvmtier_ctx_t *ctx);

/* Effective (validated) tunables. */
void      uvmtier_get_config(const uvmtier_ctx_t *ctx, uvmtier_config_t *out);

/* The context's CUDA stream (cudaStream_t as void*), and wait for all enqueued work. */
void     *uvmtier_stream(uvmtier_ctx_t *ctx);
NV_STATUS uvmtier_synchronize(uvmtier_ctx_t *ctx);

/* Library build info: returns a static string "uvmtier <version> sm_100a". */
const char *uvmtier_version(void);
/* Number of CUDA kernels this library has launched in this process (all contexts). */
uint64_t  uvmtier_kernel_launch_count(void);
/* Last CUDA error text seen by the library on this thread ("" if none). */
const char *uvmtier_last_error(void);

/* Timing on the context's stream with CUDA events (the stream the kernels are launched on).
 * uvmtier_event_record enqueues event `slot` (0..UVMTIER_MAX_EVENTS-1); uvmtier_event_elapsed_ms returns the
 * device time between two recorded slots after synchronising on the later one. */
#define UVMTIER_MAX_EVENTS 1024
NV_STATUS uvmtier_event_record(uvmtier_ctx_t *ctx, uint32_t slot);
NV_STATUS uvmtier_event_elapsed_ms(uvmtier_ctx_t *ctx, uint32_t slot_start, uint32_t slot_stop, float *out_ms);
/* cudaMemsetAsync on the context's stream (used by the harness to flush L2 between timed iterations). */
NV_STATUS uvmtier_memset_device(uvmtier_ctx_t *ctx, void *d_ptr, int value, size_t bytes);
/* Asynchronous copies on the context's stream (host memory should be pinned for true asynchrony). */
NV_STATUS uvmtier_memcpy_h2d_async(uvmtier_ctx_t *ctx, void *d_dst, const void *h_src, size_t bytes);
NV_STATUS uvmtier_memcpy_d2h_async(uvmtier_ctx_t *ctx, void *h_dst, const void *d_src, size_t bytes);
NV_STATUS uvmtier_memcpy_d2d_async(uvmtier_ctx_t *ctx, void *d_dst, const void *d_src, size_t bytes);

/* uvm_test_rng_32 stream (UVM/uvm_test_rng.c:37-62, Marsaglia KISS with the reference's constants):
 * fills out[0..n) with the first n outputs for `seed`.  Host-only; the synthetic-trace generators of
 * the harness shape all their inputs from this stream. */
void      uvmtier_test_rng_fill(uint32_t seed, size_t n, uint32_t *out);

/* Device memory helpers for C callers without their own CUDA code (thin cudaMalloc/cudaMemcpy). */
NV_STATUS uvmtier_mem_alloc_device(uvmtier_ctx_t *ctx, size_t bytes, void **out_d_ptr);
NV_STATUS uvmtier_mem_free_device(uvmtier_ctx_t *ctx, void *d_ptr);
NV_STATUS uvmtier_mem_alloc_host_pinned(uvmtier_ctx_t *ctx, size_t bytes, void **out_h_ptr);
NV_STATUS uvmtier_mem_free_host_pinned(uvmtier_ctx_t *ctx, void *h_ptr);
NV_STATUS uvmtier_memcpy_h2d(uvmtier_ctx_t *ctx, void *d_dst, const void *h_src, size_t bytes);
NV_STATUS uvmtier_memcpy_d2h(uvmtier_ctx_t *ctx, void *h_dst, const void *d_src, size_t bytes);

/* ------------------------------------------------------------------------------------------ */
/* (1) Replayable-fault batch preprocessing -- UVM/uvm_gpu_replayable_faults.c                 */
/* ------------------------------------------------------------------------------------------ */

/* One HW fault packet: NVC369_BUF_ENTRY, 32 bytes (UVM/clc369.h:33-66). */
typedef struct { uint32_t w[8]; } uvmtier_fault_packet_t;

/* Flat view of uvm_fault_buffer_entry_t (UVM/uvm_hal_types.h:376-453) after
 * parse_fault_entry_common (UVM/uvm_volta_fault_buffer.c:312-391) and the fetch-time coalescing of
 * fetch_fault_buffer_entries (UVM/uvm_gpu_replayable_faults.c:844-986).  Enum values are the
 * reference's (uvm_aperture_t, uvm_fault_type_t, uvm_fault_access_type_t, ...). */
#define UVMTIER_FE_FATAL      0x01u  /* is_fatal: fault_type >= UVM_FAULT_TYPE_FATAL (:922) */
#define UVMTIER_FE_REPLAYABLE 0x02u  /* is_replayable */
#define UVMTIER_FE_VIRTUAL    0x04u  /* is_virtual */
#define UVMTIER_FE_PROTECTED  0x08u  /* in_protected_mode */
#define UVMTIER_FE_FILTERED   0x10u  /* filtered: merged into another entry at fetch time (:753-780) */
typedef struct {
    uint64_t fault_address;          /* canonical, page aligned */
    uint64_t timestamp;
    uint64_t instance_ptr_address;
    uint16_t client_id;
    uint16_t mmu_engine_id;
    uint16_t utlb_id;
    uint16_t num_instances;          /* defined for non-filtered entries only (as in the reference) */
    uint8_t  instance_ptr_aperture;
    uint8_t  fault_type;
    uint8_t  fault_access_type;
    uint8_t  client_type;
    uint8_t  mmu_engine_type;
    uint8_t  gpc_id;
    uint8_t  ve_id;
    uint8_t  flags;                  /* UVMTIER_FE_* */
    uint8_t  access_type_mask;       /* defined for non-filtered entries only */
    uint8_t  pad[7];
} uvmtier_fault_buffer_entry_t;      /* 48 bytes */

/* One row of the instance-pointer translation table.  Replaces the lookup
 * uvm_parent_gpu_fault_entry_to_va_space (UVM/uvm_gpu.c:3491-3552): (instance_ptr, ve_id) ->
 * (va_space, gpu).  First matching row wins.  sort_rank is the dense rank of (va_space, gpu_id) over
 * the table -- the order cmp_va_space/cmp_gpu (UVM/uvm_gpu_replayable_faults.c:707-719) induce when
 * va_spaces are allocated from one array (pointer order == index order). */
#define UVMTIER_VE_ID_ANY 0xFFu
typedef struct {
    uint64_t instance_ptr;           /* instance_ptr.address */
    uint8_t  aperture;               /* uvm_aperture_t value (UVM_APERTURE_VID = 10, _SYS = 9) */
    uint8_t  ve_id;                  /* subcontext id or UVMTIER_VE_ID_ANY */
    uint16_t va_space;               /* VA space index */
    uint16_t gpu_id;                 /* uvm_gpu_id_t value */
    uint16_t sort_rank;
} uvmtier_instance_ptr_entry_t;      /* 16 bytes */

/* The uvm_parent_gpu_t fields the fault path reads. */
typedef struct {
    uint32_t max_batch_size;         /* fault_buffer.max_batch_size = uvm_perf_fault_batch_count, 1..16384 */
    uint32_t utlb_per_gpc_count;     /* Blackwell: 2 * maxTpcPerGpcCount + 1 (UVM/uvm_blackwell_fault_buffer.h:85-90) */
    uint32_t gpc_count;              /* rm_info.maxGpcCount; utlb_count = gpc_count * utlb_per_gpc_count */
    uint32_t va_bits;                /* canonical-address width (UVM/uvm_gpu.c:428-459): 57 on Hopper+ */
    uint32_t coalesce;               /* uvm_perf_fault_coalesce */
    uint32_t reserved[3];
} uvmtier_fault_gpu_params_t;        /* 32 bytes */

/* Per-batch results: the counters of uvm_fault_service_batch_context_t (UVM/uvm_gpu.h:286-339). */
typedef struct {
    uint32_t num_cached_faults;
    uint32_t num_coalesced_faults;
    uint32_t is_single_instance_ptr;
    uint32_t max_utlb_id;
    uint32_t status;                 /* NV_OK, or NV_WARN_MORE_PROCESSING_REQUIRED when an instance pointer
                                        did not translate (stale channel: UVM/uvm_gpu_replayable_faults.c:1097-1107) */
    uint32_t num_duplicate_faults;   /* filled by the service pass (:1200-1203) */
    uint32_t num_invalid_prefetch_faults;
    uint32_t num_block_groups;       /* va_block groups found in this batch */
} uvmtier_fault_batch_t;             /* 32 bytes */

/* One element of ordered_fault_cache[] (the sorted view, UVM/uvm_gpu_replayable_faults.c:1134-1178):
 * instead of a pointer into fault_cache, the index plus the fields servicing needs. */
#define UVMTIER_OF_FATAL     0x80u   /* in fault_access_type: entry is fatal (skipped by servicing, :2294) */
typedef struct {
    uint64_t fault_address;
    uint16_t fault_index;            /* index into this batch's fault_cache (ring order) */
    uint16_t ctx;                    /* row of the instance-pointer table -> va_space, gpu */
    uint16_t num_instances;
    uint8_t  access_type_mask;
    uint8_t  fault_access_type;      /* low 3 bits: uvm_fault_access_type_t; UVMTIER_OF_FATAL */
} uvmtier_ordered_fault_t;           /* 16 bytes */

/* Per input entry: what fetch-time coalescing did with it. */
#define UVMTIER_DISP_INDEX_MASK 0x0000FFFFu /* fault_cache index of the entry it was merged into (itself if it opened a new one) */
#define UVMTIER_DISP_FILTERED   0x00010000u
#define UVMTIER_DISP_FATAL      0x00020000u
#define UVMTIER_DISP_VALID      0x00040000u /* entry was fetched (valid bit set, inside the batch) */
#define UVMTIER_DISP_RULE_SHIFT 24          /* 0 new representative, 1 merged by the per-uTLB rule, 2 by the batch-wide rule */

/* A run of sorted entries that service_fault_batch_block_locked (:1375-1597) handles in one call:
 * same va_space, same gpu, same 2 MB va_block, no fatal entry. */
typedef struct {
    uint64_t block_start;            /* va_block->start (2 MB aligned VA) */
    uint16_t ctx;                    /* instance-pointer table row of the first entry */
    uint16_t first_fault_index;      /* position in this batch's ordered array */
    uint16_t block_faults;           /* entries in the run */
    uint16_t flags;                  /* bit0: run is a single fatal entry (not serviced) */
} uvmtier_fault_group_t;             /* 16 bytes */

/* Configures the context's replayable-fault state: the parent-GPU parameters and the instance-pointer
 * table (copied to the device).  Mirrors fault_buffer_init_replayable_faults
 * (UVM/uvm_gpu_replayable_faults.c:139-213): allocate-once, reused by every batch.
 * NV_ERR_INVALID_ARGUMENT for max_batch_size outside 1..16384, va_bits outside 40..57, n_rows > 256 or
 * sort ranks that do not fit the packed sort key (64 bits = rank | 45 page bits | 3 access-type bits | index bits with
 * 2^index_bits >= max_batch_size: rank < 256 for batches <= 256, < 64 up to 1024, < 16 up to 4096, < 4 up to 16384).
 * The reference clamps uvm_perf_fault_batch_count to the size of the hardware fault buffer (:153-165), not to a CTA:
 * up to 1024 entries a batch is one CTA with one thread per entry; above that, one CTA of 1024 threads walks the batch
 * in strides with its per-entry arrays in a working buffer the library keeps (59 bytes per entry), and the fetch-time
 * coalescing is the reference's sequential scan done by one thread per batch -- slower per entry (there are few such
 * batches to run side by side), same results.  The access-counter path takes max_batch_size <= 1024 only. */
NV_STATUS uvmtier_fault_buffer_init(uvmtier_ctx_t *ctx, const uvmtier_fault_gpu_params_t *gpu_params,
                                    const uvmtier_instance_ptr_entry_t *h_table, size_t n_rows);

/* Optional output of every fetch entry point below: batch_context->utlbs[u].num_pending_faults (uvm_fault_utlb_info_t,
 * UVM/uvm_gpu.h:62-84; cleared at UVM/uvm_gpu_replayable_faults.c:879-883, counted at :966 for every entry that
 * opens a new representative, read by the cancel path at :2302).  Batch b of a call writes utlb_count =
 * gpc_count * utlb_per_gpc_count counters at d_utlb_pending[b * utlb_count]; the caller sizes the device array for
 * the batches of its largest call.  NULL (the state after uvmtier_fault_buffer_init) switches the output off. */
NV_STATUS uvmtier_fault_buffer_set_utlb_pending_output(uvmtier_ctx_t *ctx, uint16_t *d_utlb_pending);

/* F1 only: parse n packets (uvm_hal_volta_fault_buffer_parse_replayable_entry without clearing the
 * valid bit).  num_instances / access_type_mask are 0. */
NV_STATUS uvmtier_fault_buffer_parse_entries_device(uvmtier_ctx_t *ctx, const uvmtier_fault_packet_t *d_packets,
                                                    size_t n, uvmtier_fault_buffer_entry_t *d_out);
NV_STATUS uvmtier_fault_buffer_parse_entries(uvmtier_ctx_t *ctx, const uvmtier_fault_packet_t *h_packets, size_t n,
                                             uvmtier_fault_buffer_entry_t *h_out);

/* fetch_fault_buffer_entries (:844-986, mode BATCH_READY) + preprocess_fault_batch (:1134-1178) + the
 * va_block grouping of service_fault_batch (:2232-2377) for ALL batches between GET and PUT at once:
 * batch b covers ring entries get + b*max_batch_size .. (modulo ring_entries), one CTA per batch.
 * Inside a batch, entries after the first one whose VALID bit is clear are not fetched.
 *
 *   d_ring            the fault ring (ring_entries packets); VALID bits of fetched entries are cleared
 *                     when clear_valid != 0 (entry_clear_valid, UVM/uvm_pascal_fault_buffer.c:304-318)
 *   d_disposition     [n] one word per ring entry in fetch order (UVMTIER_DISP_*)
 *   d_ordered         [n] batch b's sorted view starts at b*max_batch_size
 *   d_groups          [n] batch b's va_block groups start at b*max_batch_size
 *   d_batches         [n_batches]
 *   d_entries         optional [n] full parsed entries (parity/debug view), may be NULL
 * n = (put - get) mod ring_entries.  Returns the new GET in *out_get (the reference's write_get). */
NV_STATUS uvmtier_fetch_preprocess_fault_batches_device(uvmtier_ctx_t *ctx, uvmtier_fault_packet_t *d_ring,
                                                        uint32_t ring_entries, uint32_t get, uint32_t put,
                                                        int clear_valid, uint32_t *d_disposition,
                                                        uvmtier_ordered_fault_t *d_ordered,
                                                        uvmtier_fault_group_t *d_groups,
                                                        uvmtier_fault_batch_t *d_batches,
                                                        uvmtier_fault_buffer_entry_t *d_entries, uint32_t *out_get);
/* The same for a ring that lives in PINNED HOST memory (where the hardware writes it in the reference: the fault
 * buffer is sysmem, UVM/uvm_gpu_replayable_faults.c:149-184), results stay on the device.  The packets of
 * [get, put) are copied to the same positions of d_ring in chunks of chunk_batches batches on a second stream,
 * and every chunk is preprocessed on the context's stream as soon as it has arrived: the copy of chunk c+1
 * overlaps the kernels of chunk c.  chunk_batches = 0 picks one (at most UVMTIER_MAX_COPY_CHUNKS chunks per
 * call).  The host ring is not modified: VALID bits are cleared in d_ring only (clear_valid), the host owner
 * clears its own for [get, *out_get).  Enqueues only; uvmtier_synchronize waits. */
#define UVMTIER_MAX_COPY_CHUNKS 64
NV_STATUS uvmtier_fetch_preprocess_fault_batches_streamed(uvmtier_ctx_t *ctx, const uvmtier_fault_packet_t *h_ring_pinned,
                                                          uvmtier_fault_packet_t *d_ring, uint32_t ring_entries,
                                                          uint32_t get, uint32_t put, int clear_valid,
                                                          uint32_t chunk_batches, uint32_t *d_disposition,
                                                          uvmtier_ordered_fault_t *d_ordered,
                                                          uvmtier_fault_group_t *d_groups,
                                                          uvmtier_fault_batch_t *d_batches,
                                                          uvmtier_fault_buffer_entry_t *d_entries, uint32_t *out_get);
/* Host-buffer form of the above: h_ring is copied in, results are copied out (h_entries may be NULL).
 * The caller's ring is updated (cleared VALID bits) when clear_valid != 0. */
NV_STATUS uvmtier_fetch_preprocess_fault_batches(uvmtier_ctx_t *ctx, uvmtier_fault_packet_t *h_ring,
                                                 uint32_t ring_entries, uint32_t get, uint32_t put, int clear_valid,
                                                 uint32_t *h_disposition, uvmtier_ordered_fault_t *h_ordered,
                                                 uvmtier_fault_group_t *h_groups, uvmtier_fault_batch_t *h_batches,
                                                 uvmtier_fault_buffer_entry_t *h_entries, uint32_t *out_get);

/* ------------------------------------------------------------------------------------------ */
/* (1b) Servicing the sorted batches per va_block -- UVM/uvm_gpu_replayable_faults.c:1375-2377  */
/* ------------------------------------------------------------------------------------------ */

/* VA layout of the harness (stand-in for the va_range tree): va_space s owns
 * [va_base, va_base + blocks_per_space * 2 MB); its va_block b has state slot s * blocks_per_space + b. */
typedef struct {
    uint64_t va_base;
    uint32_t blocks_per_space;
    uint32_t n_spaces;
    uint32_t gpu_id;                 /* uvm_gpu_id_t value of the faulting GPU this context serves */
    uint32_t reserved;
} uvmtier_va_layout_t;

/* Per-slot va_block state (the uvm_va_block_t fields the hot path reads and writes), SoA, device memory:
 *   gpu_resident      gpu_state->resident                                    (UVM/uvm_va_block.h:101-230)
 *   cpu_resident[n]   cpu node state resident, node 0 = host DRAM, 1 = CXL   (:232-266)
 *   gpu_pte[k]        gpu_state->pte_bits[UVM_PTE_BITS_GPU_READ/WRITE/ATOMIC]
 *   prefetch_info     va_block->prefetch_info                                 (:478-483)
 *   logical_prot      compute_logical_prot() of the block's VMA: 0 NONE, 1 READ_ONLY, 3 READ_WRITE_ATOMIC
 *   preferred_location policy->preferred_location (UVMTIER_ID_INVALID = none)
 *   gpu_evicted       gpu_state->evicted: pages an eviction moved off this GPU and that have not come back
 *                     (UVM/uvm_va_block.h:161-164, set :5023-5037, cleared :4978-4988).  May be NULL when the
 *                     caller never evicts (the servicing pass then skips it).
 *   read_duplicated   va_block->read_duplicated_pages (UVM/uvm_va_block.h:436-441): pages with a copy on more than
 *                     one processor, made by UVMTIER_MR_READ_DUPLICATE requests (:5164-5332) and broken again by
 *                     every other make_resident, the servicing pass included (:4538-4579).  NULL = read
 *                     duplication is off for the table (the reference's default policy): uvmtier_block_state_alloc
 *                     leaves it NULL, uvmtier_block_state_enable_read_duplication adds it.  Invariant the library
 *                     keeps and expects of tables it is handed: a read_duplicated page has a CPU copy (true of
 *                     every page the reference marks in a block with one GPU, :5279-5310). */
typedef struct {
    uvmtier_page_mask_t *gpu_resident;
    uvmtier_page_mask_t *cpu_resident[2];
    uvmtier_page_mask_t *gpu_pte[3];
    uvmtier_prefetch_info_t *prefetch_info;
    uint8_t *logical_prot;
    uint16_t *preferred_location;
    uvmtier_page_mask_t *gpu_evicted;
    uvmtier_page_mask_t *read_duplicated;
    /* Other GPUs of the box that hold pages of these va_blocks: their gpu_state->resident masks (the reference keeps
     * one uvm_va_block_gpu_state_t per GPU in the va_block, UVM/uvm_va_block.h:101-230, 304-310), in ascending
     * uvm_gpu_id_t, which is also "closest first" among GPUs (uvm_processor_mask_find_closest_id, UVM/uvm_va_space.c:
     * 1879-1924).  Bit k of peer_direct: this GPU copies from peer k directly (block_get_can_copy_from_mask); clear:
     * the pages only such a GPU holds are staged through CPU pages (UVM/uvm_va_block.c:3140-3170, :4754-4829).  Peer
     * k's HBM arena is tier UVMTIER_TIER_PEER0 + k of the context that executes the copy runs.  n_peers = 0 (what
     * uvmtier_block_state_alloc leaves): a one-GPU table. */
    uvmtier_page_mask_t *peer_resident[4];
    uint16_t peer_gpu_id[4];
    uint32_t n_peers;
    uint32_t peer_direct;
} uvmtier_block_state_t;
#define UVMTIER_MAX_PEER_GPUS 4u

#define UVMTIER_SF_SERVICED         0x0001u /* page added to new_residency (:1560-1579) */
#define UVMTIER_SF_DUPLICATE        0x0002u /* check_fault_entry_duplicate (:1180-1191) */
#define UVMTIER_SF_FATAL            0x0004u /* mark_fault_fatal (:1235-1255) */
#define UVMTIER_SF_INVALID_PREFETCH 0x0008u /* mark_fault_invalid_prefetch (:1214-1228) */
#define UVMTIER_SF_AUTHORIZED       0x0010u /* uvm_va_block_page_is_gpu_authorized: nothing to do (:1518) */
#define UVMTIER_SF_ACCESS_SHIFT     5       /* 3 bits: service access type */
#define UVMTIER_SF_REASON_SHIFT     8       /* 3 bits: UvmEventFatalReason */
#define UVMTIER_SF_CANCEL_SHIFT     12      /* 1 bit: uvm_fault_cancel_va_mode_t */
/* with thrashing hints attached (uvmtier_service_fault_batches_thrashing_device) */
#define UVMTIER_SF_THROTTLED        0x2000u /* hint THROTTLE: mark_fault_throttled (:1537), not serviced; replay later */
#define UVMTIER_SF_PINNED           0x4000u /* hint PIN on the faulting GPU: the page is in thrashing_pin_mask (:1541-1545), serviced */
#define UVMTIER_SF_PIN_REMOTE       0x8000u /* hint PIN on another processor: not serviced by this pass (it moves pages to the
                                             * faulting GPU only); the caller moves / maps the page at the pin's residency */

/* What servicing one va_block run did: the uvm_service_block_context_t fields that survive the call
 * (UVM/uvm_va_block_types.h), plus the copy list. */
typedef struct {
    uint32_t slot;                   /* block state slot; 0xFFFFFFFF = address outside every va_range */
    uint16_t region_first, region_outer;
    uint16_t page_fault_count;
    uint16_t prefetch_count;
    uint16_t migrated[2];            /* pages copied from CPU node 0 / 1 */
    uint16_t first_touch;            /* pages populated (zeroed) at the destination */
    uint16_t hint_residency;
    uint32_t run_offset;             /* first copy run of this record in the run array */
    uint16_t run_count;
    uint16_t batch;
    uint32_t group;
    uvmtier_page_mask_t new_residency; /* per_processor_masks[gpu].new_residency after the prefetch hint */
} uvmtier_service_group_t;           /* 96 bytes */

/* One contiguous page run to move between tiers (block_copy_resident_pages_between coalesces consecutive pages
 * with the same cause when both sides are block-contiguous, UVM/uvm_va_block.c:4242-4335).  The servicing pass
 * only produces runs into HBM (dst 0); uvmtier_make_resident_device also produces runs into the CPU nodes. */
#define UVMTIER_RUN_SRC_DRAM 0u
#define UVMTIER_RUN_SRC_CXL  1u
#define UVMTIER_RUN_SRC_HBM  2u
#define UVMTIER_RUN_SRC_PEER0 3u        /* 3 + k: HBM of peer GPU k of the state table (tier UVMTIER_TIER_PEER0 + k) */
#define UVMTIER_RUN_SRC_ZERO 7u
#define UVMTIER_RUN_DST_HBM  0u
#define UVMTIER_RUN_DST_DRAM 1u
#define UVMTIER_RUN_DST_CXL  2u
typedef struct {
    uint32_t slot;
    uint16_t first_page;
    uint16_t num_pages : 10;
    uint16_t src : 3;                /* UVMTIER_RUN_SRC_*: CPU node 0 / 1, this GPU, or zero fill (first touch) */
    uint16_t prefetch : 1;           /* cause is PREFETCH rather than the fault */
    uint16_t dst : 2;                /* UVMTIER_RUN_DST_* */
} uvmtier_copy_run_t;                /* 8 bytes */

/* Allocates (zeroed, logical_prot = READ_WRITE_ATOMIC, no preferred location, prefetch_info =
 * {UVMTIER_ID_INVALID, 0}) / frees the state table of n_spaces * blocks_per_space slots. */
NV_STATUS uvmtier_block_state_alloc(uvmtier_ctx_t *ctx, const uvmtier_va_layout_t *layout, uvmtier_block_state_t *out);
NV_STATUS uvmtier_block_state_free(uvmtier_ctx_t *ctx, uvmtier_block_state_t *state);
/* Adds another GPU to a table made by uvmtier_block_state_alloc: a zeroed resident mask array (freed with the table)
 * for the GPU with id peer_gpu_id, which must be above the ids of the peers added before and differ from
 * layout->gpu_id; direct != 0: this GPU can copy from it directly.  NV_ERR_INSUFFICIENT_RESOURCES beyond
 * UVMTIER_MAX_PEER_GPUS.  The servicing passes (sections 1b, 1e, 1g) take one-GPU tables only and answer
 * NV_ERR_NOT_SUPPORTED otherwise: in the sharded set-up every GPU services the faults on the va_blocks it
 * owns; pages that sit on another GPU are moved with uvmtier_make_resident_device. */
NV_STATUS uvmtier_block_state_add_peer(uvmtier_ctx_t *ctx, const uvmtier_va_layout_t *layout, uvmtier_block_state_t *state,
                                       uint32_t peer_gpu_id, int direct);
/* Adds a zeroed read_duplicated mask array to a table made by uvmtier_block_state_alloc (freed with it). */
NV_STATUS uvmtier_block_state_enable_read_duplication(uvmtier_ctx_t *ctx, const uvmtier_va_layout_t *layout,
                                                      uvmtier_block_state_t *state);

/* service_fault_batch (:2232-2377) for ALL batches of a preprocess call: every va_block run is serviced
 * like service_fault_batch_block_locked (:1375-1597) -> uvm_va_block_service_locked
 * (UVM/uvm_va_block.c:12307): duplicate / permission / already-authorised filtering, prefetch hint,
 * make_resident residency algebra and copy list, mapping state.  Runs that touch the same va_block are
 * applied in batch order (the order the reference services them); different va_blocks run in parallel.
 *
 *   d_ordered, d_groups, d_batches   outputs of uvmtier_fetch_preprocess_fault_batches_device
 *                                    (d_batches: num_duplicate_faults / num_invalid_prefetch_faults accumulate)
 *   d_entry_flags   [n_batches*max_batch_size] UVMTIER_SF_* per sorted entry
 *   d_service       [n_batches*max_batch_size] record of run g of batch b at b*max_batch_size + g
 *                   (page_fault_count == 0 and slot set: examined, nothing to service)
 *   d_runs/run_cap  copy runs; d_counters[0] = runs needed (if > run_cap the list is truncated and the call
 *                   must be repeated on a fresh state with a larger buffer), d_counters[1] = 0 (reserved; the
 *                   number of records of batch b is d_batches[b].num_block_groups)
 * Scratch (per-slot counters and the per-block run lists) lives in the context and is reused. */
NV_STATUS uvmtier_service_fault_batches_device(uvmtier_ctx_t *ctx, uint32_t n_batches,
                                               const uvmtier_ordered_fault_t *d_ordered,
                                               const uvmtier_fault_group_t *d_groups, uvmtier_fault_batch_t *d_batches,
                                               const uvmtier_va_layout_t *layout, const uvmtier_block_state_t *state,
                                               uint16_t *d_entry_flags, uvmtier_service_group_t *d_service,
                                               uvmtier_copy_run_t *d_runs, uint32_t run_cap, uint32_t *d_counters);

/* uvm_parent_gpu_service_replayable_faults (UVM/uvm_gpu_replayable_faults.c:2906-2960) for a ring in PINNED HOST
 * memory: the reference's loop is fetch -> preprocess -> service, one batch after the other; here the packets of
 * [get, put) cross the link in chunks of chunk_batches batches (0 picks one) on a second stream and every chunk is
 * preprocessed AND serviced as soon as it has arrived, while the next chunk is still in flight.  Runs that touch
 * the same va_block are applied in batch order, exactly as when the whole trace is serviced in one pass (chunk
 * order is batch order); records carry trace-wide batch numbers; d_counters accumulate over the chunks.
 *
 * What the caller of service_fault_batch consumes afterwards -- the batch headers, the per-fault flags and the
 * two counters -- can be copied back chunk by chunk on a third stream: pass pinned buffers in `readback`
 * (NULL, or NULL members, skip it).  Service records and copy runs stay on the device, where
 * uvmtier_copy_runs_device consumes them.  Enqueues only; uvmtier_synchronize waits for everything, copies back
 * included.  Other arguments as in uvmtier_fetch_preprocess_fault_batches_streamed and
 * uvmtier_service_fault_batches_device. */
typedef struct {
    uvmtier_fault_batch_t *h_batches;   /* [n_batches] pinned */
    uint16_t *h_entry_flags;            /* [n] pinned, n = (put - get) mod ring_entries */
    uint32_t *h_counters;               /* [2] pinned */
} uvmtier_service_readback_t;
NV_STATUS uvmtier_service_replayable_faults_streamed(uvmtier_ctx_t *ctx, const uvmtier_fault_packet_t *h_ring_pinned,
                                                     uvmtier_fault_packet_t *d_ring, uint32_t ring_entries,
                                                     uint32_t get, uint32_t put, int clear_valid, uint32_t chunk_batches,
                                                     uint32_t *d_disposition, uvmtier_ordered_fault_t *d_ordered,
                                                     uvmtier_fault_group_t *d_groups, uvmtier_fault_batch_t *d_batches,
                                                     const uvmtier_va_layout_t *layout, const uvmtier_block_state_t *state,
                                                     uint16_t *d_entry_flags, uvmtier_service_group_t *d_service,
                                                     uvmtier_copy_run_t *d_runs, uint32_t run_cap, uint32_t *d_counters,
                                                     const uvmtier_service_readback_t *readback, uint32_t *out_get);

/* The same servicing pass for callers that are not allowed to allocate.  (This entry point has no counterpart
 * in the reference; it exists for the user-space setting.)
 *
 * A fault-servicing thread is the code that runs when GPU memory is exhausted -- servicing is what evicts and
 * frees it -- so it cannot depend on cudaMalloc succeeding, or returning promptly, for its own working space;
 * integrations that fix their memory budget at start-up and forbid allocator calls afterwards are in the same
 * position.  uvmtier_service_fault_batches_device grows its working space with cudaMalloc; this variant uses a
 * fixed array that is part of the library image instead and calls no cudaMalloc / cudaFree or any other
 * allocating routine (it enqueues kernels, two memsets and event records / waits on existing streams).  Arguments, results and semantics are those of uvmtier_service_fault_batches_device.
 *
 * Limits (NV_ERR_INVALID_ARGUMENT beyond them, uvmtier_last_error() says which):
 *   n_spaces * blocks_per_space <= 65536 va_block slots (128 GB of VA),
 *   n_batches * max_batch_size  <= 262144 entries,  n_batches <= 4096.
 * The array is about 4.5 MB and every GPU that loads the library holds it until the process exits.
 * Nothing is created by the call itself: uvmtier_init has already looked up the array's device address (with the
 * context's GPU current; that lookup is also what makes the CUDA runtime load the array) and created the pass's two
 * side streams and their events -- so create the context while memory is still to be had, and before any stream
 * capture is begun.  All calls on a GPU share the one array, which keeps the previous call's
 * contents: the function clears the parts that have to start at zero itself, but two calls whose GPU work
 * overlaps (different streams or contexts on the same GPU) race on it -- serialise them. */
NV_STATUS uvmtier_service_fault_batches_noalloc_device(uvmtier_ctx_t *ctx, uint32_t n_batches,
                                                       const uvmtier_ordered_fault_t *d_ordered,
                                                       const uvmtier_fault_group_t *d_groups,
                                                       uvmtier_fault_batch_t *d_batches,
                                                       const uvmtier_va_layout_t *layout,
                                                       const uvmtier_block_state_t *state, uint16_t *d_entry_flags,
                                                       uvmtier_service_group_t *d_service, uvmtier_copy_run_t *d_runs,
                                                       uint32_t run_cap, uint32_t *d_counters);

/* ------------------------------------------------------------------------------------------ */
/* (1c) make_resident on the state table -- UVM/uvm_va_block.c:4877-5114                       */
/* ------------------------------------------------------------------------------------------ */

/* uvm_make_resident_cause_t (UVM/uvm_va_block_types.h:116-129) */
#define UVMTIER_CAUSE_REPLAYABLE_FAULT     0u
#define UVMTIER_CAUSE_NON_REPLAYABLE_FAULT 1u
#define UVMTIER_CAUSE_ACCESS_COUNTER       2u
#define UVMTIER_CAUSE_PREFETCH             3u
#define UVMTIER_CAUSE_EVICTION             4u
#define UVMTIER_CAUSE_API_TOOLS            5u
#define UVMTIER_CAUSE_API_MIGRATE          6u
#define UVMTIER_CAUSE_API_SET_RANGE_GROUP  7u
#define UVMTIER_CAUSE_API_HINT             8u

#define UVMTIER_MR_HAS_PAGE_MASK     0x1u  /* page_mask = d_page_masks[mask_row], else NULL (= the whole region) */
#define UVMTIER_MR_HAS_PREFETCH_MASK 0x2u  /* prefetch_page_mask = d_prefetch_masks[mask_row] (fault causes only) */
#define UVMTIER_MR_EVICT_GPU_PAGES   0x4u  /* page_mask = the block's gpu_resident pages in the region: what
                                              uvm_va_block_evict_chunks builds for a 2 MB root chunk (:13055-13080) */
#define UVMTIER_MR_READ_DUPLICATE    0x8u  /* uvm_va_block_make_resident_read_duplicate (:5164-5332): COPY instead
                                              of MOVE -- the sources keep their pages, write access to them is
                                              revoked, read_duplicated |= pages_migrated; a CPU destination on
                                              another NUMA node switches node instead (one CPU page table).  Needs
                                              state->read_duplicated; not with cause EVICTION or _EVICT_GPU_PAGES
                                              (the reference asserts), such requests are rejected. */

/* One uvm_va_block_make_resident call (UVM/uvm_va_block.h:770-777): move the pages of `region` (and page_mask)
 * to dest_tier.  dest_tier = UVMTIER_TIER_HBM is dest_id = this GPU; UVMTIER_TIER_DRAM / _CXL is dest_id = CPU
 * with make_resident.dest_nid = node 0 / 1. */
typedef struct {
    uint32_t slot;
    uint16_t region_first, region_outer;
    uint8_t  dest_tier;
    uint8_t  cause;                  /* UVMTIER_CAUSE_* */
    uint16_t flags;                  /* UVMTIER_MR_* */
    uint32_t mask_row;
} uvmtier_make_resident_req_t;       /* 16 bytes */

/* What the call leaves in va_block_context->make_resident (UVM/uvm_va_block_types.h:197-242). */
typedef struct {
    uint32_t slot;
    uint16_t migrated[3];            /* pages copied from CPU node 0, CPU node 1, this GPU */
    uint16_t first_touch;            /* pages resident nowhere: populated and zeroed at the destination */
    uint32_t run_offset;             /* first copy run of this request in the run array */
    uint16_t run_count;
    uint16_t unmapped;               /* GPU mappings removed (pages the GPU had mapped that left it) */
    uint16_t migrated_peer[4];       /* pages copied from peer GPU k of the table */
    uint32_t reserved;
    uvmtier_page_mask_t pages_migrated;
    uvmtier_page_mask_t pages_changed_residency;
} uvmtier_make_resident_result_t;    /* 160 bytes */

/* Batched uvm_va_block_make_resident (= _copy + _finish, :5086-5114) on the flat state table: residency algebra
 * (block_copy_resident_pages :4660-4847: copy = page_mask & ~resident(dest); GPU destination takes from the other
 * GPUs of the table it can copy from directly, in ascending id, then from CPU node 0, then node 1, and what only a GPU
 * without direct access holds goes through CPU pages on node 0 in two runs; CPU destination takes from the other CPU
 * node first, then from the GPUs in ascending id, this one among them), first touch
 * (:4581-4619), state update (:4990-5040: resident masks, gpu_evicted set on EVICTION / cleared on return),
 * clearing of every other residency, the other GPUs' included (break_read_duplication_in_region :4538-4579), removal of the GPU's mappings
 * of pages that leave it (:4919-4926), and the copy list: runs d_runs[run_offset .. run_offset + run_count) per
 * request (uvmtier_copy_runs_device executes them).  Requests of one call must name distinct slots (a va_block
 * is locked for the duration of a make_resident; two requests on one block in the same call would race).
 * d_page_masks / d_prefetch_masks may be NULL when no request refers to them.  d_counters[0] += number of runs
 * (the caller zeroes it, or keeps accumulating behind a servicing pass). */
NV_STATUS uvmtier_make_resident_device(uvmtier_ctx_t *ctx, uint32_t n_requests,
                                       const uvmtier_make_resident_req_t *d_requests,
                                       const uvmtier_page_mask_t *d_page_masks,
                                       const uvmtier_page_mask_t *d_prefetch_masks,
                                       const uvmtier_va_layout_t *layout, const uvmtier_block_state_t *state,
                                       uvmtier_make_resident_result_t *d_results, uvmtier_copy_run_t *d_runs,
                                       uint32_t run_cap, uint32_t *d_counters);

/* ------------------------------------------------------------------------------------------ */
/* (1d) Which HBM root chunk to evict -- UVM/uvm_pmm_gpu.c:1395-1500                           */
/* ------------------------------------------------------------------------------------------ */
/* The reference keeps every allocated 2 MB root chunk on one of two lists, root_chunks.va_block_unused (its
 * va_block has no page resident on the GPU) and root_chunks.va_block_used; every list move goes to the tail
 * (root_chunk_update_eviction_list :1430-1448), and pick_root_chunk_to_evict (:1460-1500) takes the head of
 * the unused list, else the head of the used list: least recently serviced first.  A va_block of the state table
 * is one root chunk (2 MB blocks, UVM/uvm_va_block.c:3547-3566), so the two lists are one 64-bit stamp per slot:
 *   0                          the va_block holds no root chunk
 *   bit 63 | sequence number   on va_block_used;  bit 63 clear: on va_block_unused
 * where the sequence number is that of the last move.  The moves come from block_mark_memory_used (after every
 * make_resident that leaves the GPU with resident pages, :5076-5077 and the fault path), from
 * block_clear_resident_processor (the GPU's last page left, :3579-3603), from chunk allocation
 * (chunk_update_lists_locked :639-643) and from the free after an eviction; they are applied by the two
 * "note" calls below, in the order the reference would have made them ((batch, run) order of a servicing pass,
 * request order of a make_resident call). */
#define UVMTIER_PMM_USED (1ull << 63)
typedef struct {
    uint64_t *d_stamp;               /* [n_slots], device */
    uint32_t *d_scratch;             /* working space of the calls below, device */
    uint64_t  clock;                 /* sequence numbers handed out so far */
    uint32_t  n_slots;
    uint32_t  capacity_chunks;       /* root chunks HBM can hold: what the caller keeps chunks_in_use below */
    /* uvmtier_pmm_enable_chunk_table: the HBM arena holds capacity_chunks chunks of 2 MB instead of one per va_block;
     * d_chunk_of_slot[slot] = the chunk behind the va_block (0xFFFFFFFF = none), d_free_bits = one bit per chunk.  NULL:
     * the arena is VA-sized and slot s lives at s * 2 MB. */
    uint32_t *d_chunk_of_slot;
    uint32_t *d_free_bits;
    uint64_t *d_dram_stamp;          /* [n_slots]: round of the tiered loop in which the va_block's pages last reached DRAM
                                        from HBM (0 = never / none left there): the order of DRAM -> CXL demotion */
    uint64_t  round;                 /* rounds of the tiered loop so far */
} uvmtier_pmm_t;
NV_STATUS uvmtier_pmm_create(uvmtier_ctx_t *ctx, const uvmtier_va_layout_t *layout, uint32_t capacity_chunks,
                             uvmtier_pmm_t *out);
NV_STATUS uvmtier_pmm_destroy(uvmtier_ctx_t *ctx, uvmtier_pmm_t *pmm);
/* Puts root chunks behind the va_blocks (block_populate_gpu_chunk, UVM/uvm_va_block.c:2974-3049; the free after an
 * eviction, UVM/uvm_pmm_gpu.c:1290-1380): from now on a va_block takes one of the capacity_chunks 2 MB chunks of the HBM
 * arena when its first page arrives on the GPU (the "note" calls below hand it out) and gives it back when an eviction
 * empties it, and every copy of this context that names the HBM tier through copy runs (uvmtier_copy_runs_device, the
 * tiered loop) goes to chunk_of_slot[slot] * 2 MB + page * 4 KB.  The HBM tier then needs capacity_chunks * 2 MB, not the
 * size of the VA.  Call it on a table without GPU-resident pages (all stamps 0), once per context; which chunk a
 * va_block gets is arbitrary.  The table belongs to pmm; uvmtier_pmm_destroy detaches it. */
NV_STATUS uvmtier_pmm_enable_chunk_table(uvmtier_ctx_t *ctx, uvmtier_pmm_t *pmm);
/* The list moves of one servicing pass (d_service as written by uvmtier_service_fault_batches_device for
 * n_batches batches): every serviced run moves its va_block to the tail of va_block_used. */
NV_STATUS uvmtier_pmm_note_service_device(uvmtier_ctx_t *ctx, uvmtier_pmm_t *pmm, uint32_t n_batches,
                                          const uvmtier_fault_batch_t *d_batches, const uvmtier_service_group_t *d_service);
/* The list moves of one uvmtier_make_resident_device call: to the GPU -> tail of va_block_used; the GPU's last
 * pages left -> tail of va_block_unused; an UVMTIER_MR_EVICT_GPU_PAGES request that emptied the GPU -> the chunk
 * is freed (stamp 0), as evict_root_chunk does after uvm_va_block_evict_chunks (:1290-1380). */
NV_STATUS uvmtier_pmm_note_make_resident_device(uvmtier_ctx_t *ctx, uvmtier_pmm_t *pmm, uint32_t n_requests,
                                                const uvmtier_make_resident_req_t *d_requests,
                                                const uvmtier_make_resident_result_t *d_results,
                                                const uvmtier_block_state_t *state);
/* How many va_blocks touched by these preprocessed batches hold no root chunk (each needs one when it is
 * serviced: block_populate_gpu_chunk, UVM/uvm_va_block.c:2974-3049), and how many chunks are in use now:
 * d_out[0] = needed, d_out[1] = in use. */
NV_STATUS uvmtier_pmm_count_needed_device(uvmtier_ctx_t *ctx, uvmtier_pmm_t *pmm, uint32_t n_batches,
                                          const uvmtier_fault_group_t *d_groups, const uvmtier_fault_batch_t *d_batches,
                                          const uvmtier_va_layout_t *layout, uint32_t *d_out);
/* pick_root_chunk_to_evict, k times: writes one eviction request per victim (whole block, cause EVICTION, flag
 * UVMTIER_MR_EVICT_GPU_PAGES, destination dest_tier = UVMTIER_TIER_DRAM or _CXL) to d_requests[0 .. *d_count),
 * in no particular order; fewer than k when fewer chunks are allocated.  The victims keep their stamps until
 * uvmtier_pmm_note_make_resident_device sees their evictions. */
NV_STATUS uvmtier_pmm_pick_chunks_to_evict_device(uvmtier_ctx_t *ctx, uvmtier_pmm_t *pmm, uint32_t k, uint32_t dest_tier,
                                                  uvmtier_make_resident_req_t *d_requests, uint32_t *d_count);
/* Pages resident on the GPU, on CPU node 0 (DRAM) and on CPU node 1 (CXL): d_out[0..2]. */
NV_STATUS uvmtier_block_state_count_pages_device(uvmtier_ctx_t *ctx, const uvmtier_va_layout_t *layout,
                                                 const uvmtier_block_state_t *state, uint32_t *d_out);

/* Servicing under HBM oversubscription (BASELINE config D): the preprocessed batches are serviced in chunks of
 * chunk_batches batches ("rounds"); before a round is serviced, room is made for the root chunks it needs.  With
 *     overflow = chunks in use + va_blocks the round faults on that hold no chunk - pmm->capacity_chunks
 * victims are taken in the reference's order (pick_root_chunk_to_evict: ascending stamp) until `overflow` of them are
 * va_blocks the round does not fault on -- a victim the round faults on needs a chunk again at once, so it frees
 * nothing -- then evicted with make_resident (cause EVICTION) and their pages copied out; then the round is serviced
 * and its pages copied in.  The reference evicts on demand, one root chunk at a time, when block_populate_gpu_chunk
 * finds no free chunk and retries the block (NV_ERR_MORE_PROCESSING_REQUIRED, UVM/uvm_va_block.h:736-745,
 * UVM/uvm_pmm_gpu.c:1502-1560); evicting a round's worth ahead takes the same victims from the same lists as long as
 * no victim is faulted on again within the round (then the reference would have evicted it a little later) --
 * chunk_batches = 1 is the closest.  A round that alone faults on more va_blocks than HBM has chunks is serviced over
 * capacity and counted (not allowed with a chunk table: chunk_batches * max_batch_size <= capacity_chunks then).
 * Where evicted pages go is this library's policy, the reference has none: host DRAM while it then holds
 * at most dram_capacity_pages pages of this table, else the CXL node; with UVMTIER_TIERED_DEMOTE room is first made in
 * DRAM by moving the DRAM pages of the va_blocks that left HBM longest ago (never-promoted ones first; not those the
 * round faults on) to the CXL node, so that DRAM keeps the warmer pages.
 * Every decision of a round is taken on the device (counts, weighted radix selection over the stamps, request lists,
 * copy lists): the call enqueues the whole trace and synchronises once at the end.
 * d_entry_flags / d_service / d_batches are filled as by uvmtier_service_fault_batches_device; the copy runs are
 * consumed round by round (UVMTIER_TIERED_EXECUTE_COPIES runs them: the tiers must exist), record.run_offset counts
 * within the record's round. */
#define UVMTIER_TIERED_EXECUTE_COPIES 1
#define UVMTIER_TIERED_DEMOTE         2
typedef struct {
    uint32_t passes;                 /* chunks of batches serviced */
    uint32_t eviction_waves;
    uint32_t evicted_to_dram, evicted_to_cxl;   /* root chunks */
    uint32_t peak_chunks_in_use, chunks_in_use; /* after the last pass */
    uint64_t pages_evicted, pages_promoted;     /* pages copied out of / into HBM (first touch included) */
    uint32_t overcommitted_passes;              /* chunks of batches that alone needed more root chunks than HBM holds:
                                                   everything else was evicted and they were serviced over capacity */
    uint32_t blocks_demoted;                    /* va_blocks whose DRAM pages were moved to the CXL node */
    uint32_t reserved[2];
    uint64_t pages_demoted;                     /* pages copied DRAM -> CXL */
} uvmtier_tiering_stats_t;
NV_STATUS uvmtier_service_fault_batches_tiered_device(uvmtier_ctx_t *ctx, uvmtier_pmm_t *pmm, uint32_t n_batches,
                                                      const uvmtier_ordered_fault_t *d_ordered,
                                                      const uvmtier_fault_group_t *d_groups, uvmtier_fault_batch_t *d_batches,
                                                      const uvmtier_va_layout_t *layout, const uvmtier_block_state_t *state,
                                                      uint16_t *d_entry_flags, uvmtier_service_group_t *d_service,
                                                      uint32_t chunk_batches, uint64_t dram_capacity_pages, int flags,
                                                      uvmtier_tiering_stats_t *out_stats);

/* ------------------------------------------------------------------------------------------ */
/* (1e) Access-counter notifications -- UVM/uvm_gpu_access_counters.c                          */
/* ------------------------------------------------------------------------------------------ */
/* The sibling front end on the same back end: the GPU reports regions it keeps reaching remotely and the driver
 * migrates them to it.  32-byte NVC365 packets (UVM/clc365.h:33-80) are fetched (:862-940), parsed
 * (UVM/uvm_turing_access_counter_buffer.c:154-198), sorted by (va_space, gpu, address) (preprocess_notifications
 * :974-992), expanded per va_block into a page mask (expand_notification_block :1287-1354) and passed through the
 * fault path's back half (service_va_block_locked :1079-1249 -> uvm_va_block_service_locked).  Uses the
 * instance-pointer table and max_batch_size of uvmtier_fault_buffer_init (uvm_perf_access_counter_batch_count has
 * the same default, 256). */
typedef struct { uint32_t w[8]; } uvmtier_access_counter_packet_t;       /* NVC365_NOTIFY_BUF_SIZE = 32 */

/* uvm_access_counter_buffer_entry_t (UVM/uvm_hal_types.h:481-520) */
typedef struct {
    uint64_t address;                /* canonical VA of the tracked region */
    uint64_t instance_ptr;           /* instance_ptr.address */
    uint32_t instance_aperture;      /* uvm_aperture_t value (UVM_APERTURE_VID = 10, _SYS = 9) */
    uint32_t mmu_engine_id;
    uint32_t ve_id;
    uint32_t counter_value;
    uint32_t sub_granularity;        /* one bit per 1/32 of the tracked region */
    uint32_t bank, tag;
    uint32_t valid;                  /* entry was fetched */
} uvmtier_access_counter_entry_t;    /* 48 bytes */

/* One element of batch_context->notifications[] after preprocess_notifications (the sorted pointer array). */
#define UVMTIER_AC_CTX_NONE 0xFFFFu  /* instance pointer did not translate: va_space NULL, entry ignored (:957-960) */
typedef struct {
    uint64_t address;
    uint16_t index;                  /* index into this batch's notification_cache (ring order) */
    uint16_t ctx;                    /* row of the instance-pointer table, or UVMTIER_AC_CTX_NONE */
    uint32_t sub_granularity;
} uvmtier_ordered_notification_t;    /* 16 bytes */

/* current_config of the notification buffer (:328-353) and the one harness choice the expansion depends on: the
 * size of the CPU chunks backing a va_block (uvm_va_block_get_physical_size, UVM/uvm_va_block.c:2353-2373).  With
 * 2 MB tracking and 2 MB CPU chunks a notification expands to the 64 KB sub-regions its sub_granularity names,
 * otherwise to the one page at its address (:1325-1352); other tracking granularities are ignored (:1303-1306). */
typedef struct {
    uint64_t granularity_bytes;      /* default 2 MB */
    uint64_t cpu_chunk_bytes;        /* default 2 MB (tiers back a va_block contiguously); 4096 = page-sized chunks */
} uvmtier_access_counter_config_t;
NV_STATUS uvmtier_access_counters_init(uvmtier_ctx_t *ctx, const uvmtier_access_counter_config_t *cfg /* NULL: defaults */);

/* fetch + parse + preprocess for all batches between GET and PUT, one CTA per batch; outputs sliced like the fault
 * path's (batch b at b * max_batch_size).  d_groups: runs of notifications of one (va_space, gpu, va_block), what
 * service_notifications_in_block (:1356-1427) handles in one call; untranslated notifications sort first and belong
 * to no group.  d_batches: num_cached_faults = num_coalesced_faults = notifications fetched, is_single_instance_ptr
 * (:925-930), num_block_groups.  d_entries: optional parsed entries in ring order.  Ordering compares page numbers
 * (tracked regions are at least 64 KB aligned); ties keep cache order. */
NV_STATUS uvmtier_access_counters_fetch_preprocess_device(uvmtier_ctx_t *ctx, uvmtier_access_counter_packet_t *d_ring,
                                                          uint32_t ring_entries, uint32_t get, uint32_t put, int clear_valid,
                                                          uvmtier_ordered_notification_t *d_ordered,
                                                          uvmtier_fault_group_t *d_groups, uvmtier_fault_batch_t *d_batches,
                                                          uvmtier_access_counter_entry_t *d_entries, uint32_t *out_get);
/* service_notifications_batch (:1513-1615) for all batches: per va_block run, in batch order, the notifications are
 * expanded into accessed pages -- only while the GPU still maps something in the block (:1103-1104), pages already
 * on the GPU are stale (:1122-1123) -- then prefetch hint, make_resident to the GPU, mapping.  Records, copy runs and
 * counters as uvmtier_service_fault_batches_device (page_fault_count = pages the notifications named); the records
 * are also what uvmtier_pmm_note_service_device takes, so a promotion moves its va_block to the tail of the
 * eviction order like a serviced fault does. */
NV_STATUS uvmtier_access_counters_service_device(uvmtier_ctx_t *ctx, uint32_t n_batches,
                                                 const uvmtier_ordered_notification_t *d_ordered,
                                                 const uvmtier_fault_group_t *d_groups, uvmtier_fault_batch_t *d_batches,
                                                 const uvmtier_va_layout_t *layout, const uvmtier_block_state_t *state,
                                                 uvmtier_service_group_t *d_service, uvmtier_copy_run_t *d_runs,
                                                 uint32_t run_cap, uint32_t *d_counters);

/* ------------------------------------------------------------------------------------------ */
/* (1f) Thrashing detection -- UVM/uvm_perf_thrashing.c:1230-1380                              */
/* ------------------------------------------------------------------------------------------ */

/* The reference learns about thrashing from two perf events, raised under the va_block lock: a MIGRATION
 * (uvm_va_block.c:4291,4303) and a permission REVOCATION.  thrashing_event_cb (:1230-1380) keeps, per page, the time
 * of the last event, a 3-bit count of events that followed the previous one within `lapse` and the processors
 * involved; a page whose count reaches `threshold` enters the va_block's thrashing_pages mask (thrashing_detected
 * :1023-1040) -- the mask uvm_perf_thrashing_get_thrashing_pages (:1803) hands to the prefetcher, which leaves those
 * pages alone (uvm_perf_prefetch.c:378-381,407) -- and leaves it when the window passes without an event
 * (thrashing_reset_page :1045-1073) or a user-driven migration covers it (:1272-1285).  Per-page state only exists
 * once two different processors touched the va_block inside the window (:1316-1337).
 * Events of evictions and of the first half of staged copies are not sent (:1254-1260).
 * The second half is the hint (below): what a fault on a thrashing page should do -- wait (throttle), or have the page
 * pinned somewhere all the thrashing processors can reach and map it from there. */
#define UVMTIER_PERF_EVENT_MIGRATION  0u
#define UVMTIER_PERF_EVENT_REVOCATION 1u
#define UVMTIER_PERF_EVENT_RESETS     0x1u /* flags: a migration whose cause is not a fault, an access counter or a
                                            * prefetch, or that copies instead of moving, or under read duplication */
#define UVMTIER_PERF_EVENT_PREFETCH   0x2u /* flags: a migration whose cause is PREFETCH (matters only with the hint half
                                            * on: a fault / access-counter migration onto pinned pages is the pin being
                                            * carried out and is not an event, is_migration_pinned_pages_update :1198-1226) */
typedef struct {
    uint64_t time_stamp;     /* ns (NV_GETTIME() at the callback); events of a va_block arrive in time order */
    uint32_t slot;           /* va_block: its row of the state table                                          */
    uint16_t first_page;     /* region: (address - block start) / 4K                                          */
    uint16_t num_pages;      /* bytes / 4K; first_page + num_pages <= 512                                      */
    uint16_t processor;      /* migration.dst / revocation.proc_id, < 32                                       */
    uint8_t kind;            /* UVMTIER_PERF_EVENT_MIGRATION / _REVOCATION                                     */
    uint8_t flags;           /* UVMTIER_PERF_EVENT_RESETS                                                      */
    uint32_t reserved;
} uvmtier_perf_event_t;      /* 24 bytes */

typedef struct {
    uint64_t last_time_stamp;           /* block_thrashing_info_t (:88-119), the detector's fields */
    uint64_t last_thrashing_time_stamp;
    uint16_t last_processor;
    uint16_t num_thrashing_pages;
    uint32_t tracked;                   /* pages != NULL */
    uint16_t pinned_count;              /* pinned_pages.count (:107): pins ever made, the reference never counts down */
    uint8_t  thrashing_reset_count;     /* epoch resets so far (:94) */
    uint8_t  reserved;
    uint32_t throttling_count;          /* statistic (:103) */
} uvmtier_thrashing_block_t;            /* 32 bytes */

/* The hint half's parameters (va_space_thrashing_info_t.params :172-199; nap / epoch / pin are the module parameters
 * times the lapse, :246-310) and the processor topology the decision reads (uvm_va_space_t.accessible_from, can_access,
 * has_fast_link, has_native_atomics: one 32-bit processor mask per processor id; id 0 = the CPU, GPUs from 1, ids that
 * share (id - 1) / 8 are SMC partitions of one GPU).  has_memory: uvm_processor_has_memory. */
typedef struct {
    uint64_t lapse_ns, nap_ns, epoch_ns, pin_ns;
    uint32_t threshold, pin_threshold, max_resets, has_memory;
    uint32_t accessible_from[32], can_access[32], has_fast_link[32], has_native_atomics[32];
} uvmtier_thrashing_hint_config_t;      /* 560 bytes */

/* One call of uvm_perf_thrashing_get_hint(va_block, address, requester) (:1615): a fault of `requester` on page `page`
 * of va_block `slot` at time `time_stamp`.  What the reference reads from the va_block comes with the query:
 * resident_processors = where the page lives now (uvm_va_block_page_resident_processors), preferred_location = the
 * policy's (0xFFFF = none). */
typedef struct {
    uint64_t time_stamp;
    uint32_t slot;
    uint16_t page;
    uint8_t  requester;                 /* < 32 */
    uint8_t  reserved;
    uint32_t resident_processors;
    uint16_t preferred_location;
    uint16_t reserved2;
} uvmtier_thrashing_query_t;            /* 24 bytes */

/* uvm_perf_thrashing_hint_t (UVM/uvm_perf_thrashing.h:33-72) */
#define UVMTIER_THRASHING_HINT_NONE     0u
#define UVMTIER_THRASHING_HINT_PIN      1u
#define UVMTIER_THRASHING_HINT_THROTTLE 2u
typedef struct {
    uint64_t throttle_end_time_stamp;   /* THROTTLE: when the requester may service faults on the page again */
    uint32_t pin_processors;            /* PIN: the processors to map to the new residency */
    uint8_t  type;
    uint8_t  pin_residency;             /* PIN: where the page is to live; 0xFF otherwise */
    uint16_t reserved;
} uvmtier_thrashing_hint_t;             /* 16 bytes */

/* Device tables, one row per va_block slot of the layout.  d_page_word: the reference's first bit-field word of
 * page_thrashing_info_t (:46-63) -- bits 0-57 last_time_stamp in 64 ns units, 58 has_migration_events,
 * 59 has_revocation_events, 60-62 num_thrashing_events, 63 pinned (always 0 here).  d_thrashing_pages: 16 words per
 * slot, the layout of the residency masks, so it can be passed wherever a thrashing mask is taken (section 2). */
typedef struct {
    uvmtier_thrashing_block_t *d_blocks;
    uint64_t *d_page_word;        /* n_slots * 512 */
    uint32_t *d_page_processors;  /* n_slots * 512: page_thrashing_info_t.processors */
    uint32_t *d_thrashing_pages;  /* n_slots * 16 */
    uint32_t *d_scratch;
    uint64_t lapse_ns;            /* uvm_perf_thrashing_lapse_usec * 1000 (default 500 us, :264) */
    uint32_t threshold;           /* uvm_perf_thrashing_threshold (default 3, :246); 1..7 */
    uint32_t n_slots;
    /* the hint half (uvmtier_thrashing_enable_hints; NULL until then).  Per page: d_page_throttle = throttling end time
     * stamp in 64 ns units (:66); d_page_hint = bits 0-7 throttling_count (:71), 8-15 pinned_residency_id + 1 (:81),
     * 16-23 do_not_throttle_processor_id + 1 (:84); d_page_throttled = throttled_processors (:78); d_page_deadline = when
     * the pin expires (:132).  d_pinned_pages: pinned_pages.mask, 16 words per slot.  Bit 63 of d_page_word is `pinned`. */
    uint64_t *d_page_throttle;
    uint32_t *d_page_hint;
    uint32_t *d_page_throttled;
    uint64_t *d_page_deadline;
    uint32_t *d_pinned_pages;
    uvmtier_thrashing_hint_config_t *d_hint_config;
    uvmtier_thrashing_hint_config_t hint_config;
} uvmtier_thrashing_t;

/* lapse_ns = 0 and threshold = 0 take the reference's defaults.  Tables start out all zero (no history). */
NV_STATUS uvmtier_thrashing_create(uvmtier_ctx_t *ctx, const uvmtier_va_layout_t *layout, uint64_t lapse_ns, uint32_t threshold,
                                   uvmtier_thrashing_t *out);
NV_STATUS uvmtier_thrashing_destroy(uvmtier_ctx_t *ctx, uvmtier_thrashing_t *t);

/* Applies n_events events, asynchronously on the context's stream.  d_events must be grouped by slot (all events of a
 * va_block next to each other, in time order); the groups may come in any order.  d_counters (4 words, zeroed
 * here): [0] pages detected as thrashing (the reference's num_thrashing statistic), [1] pages reset, [2] events
 * dropped -- malformed, or belonging to a second group of a va_block that already had one in this call. */
NV_STATUS uvmtier_thrashing_note_events_device(uvmtier_ctx_t *ctx, uvmtier_thrashing_t *t, uint32_t n_events,
                                               const uvmtier_perf_event_t *d_events, uint32_t *d_counters);

/* Turns the hint half on: allocates its (zeroed) tables.  cfg->lapse_ns / threshold are replaced by the detector's.
 * From now on uvmtier_thrashing_note_events_device also applies what the reference's callback does with pinned and
 * throttled pages: a reset ends the page's throttling and its pin (:1057-1061), and a fault / access-counter migration
 * whose whole region is pinned is not an event (:1198-1226, :1293-1295).
 * uvmtier_thrashing_default_hint_config fills cfg with the module-parameter defaults (pin_threshold 10, nap 1, epoch 2000,
 * pin 300 lapses, max_resets 4; :246-310) for a box with one CPU (id 0) and the GPUs of gpu_mask: with fast_links = 0 a PCIe
 * box (the CPU reaches no GPU memory, GPUs reach the CPU's and each other's); with fast_links = 1 NVLink between the GPUs
 * (fast links and native atomics among them); with fast_links = 2 a coherent box where the CPU is part of that too. */
void uvmtier_thrashing_default_hint_config(uvmtier_thrashing_hint_config_t *cfg, uint32_t gpu_mask, int fast_links);
NV_STATUS uvmtier_thrashing_enable_hints(uvmtier_ctx_t *ctx, uvmtier_thrashing_t *t, const uvmtier_thrashing_hint_config_t *cfg);

/* uvm_perf_thrashing_get_hint (:1615-1780) for n_queries faults, asynchronously on the context's stream: d_hints[i]
 * answers d_queries[i] and the pages' throttling / pinning state advances as in the reference (throttle periods and the
 * processor exempt from them :772-832, pinning after pin_threshold periods and where :1455-1595, the quiet-epoch reset of
 * a va_block :1644-1674).  d_queries must be grouped by slot, in time order inside a group (as the events).
 * d_counters (8 words, zeroed here): [2] queries dropped (malformed, or a second group of a va_block), [3] throttles,
 * [4] pins on the requester, [5] pins elsewhere (the reference's num_throttle / num_pin_local / num_pin_remote).
 * What the caller does with a hint is the fault path's business (UVM/uvm_gpu_replayable_faults.c:1524-1545): THROTTLE =
 * leave the fault unserviced until the time given, PIN = service it with the pin's residency as destination and map the
 * processors named. */
NV_STATUS uvmtier_thrashing_get_hints_device(uvmtier_ctx_t *ctx, uvmtier_thrashing_t *t, uint32_t n_queries,
                                             const uvmtier_thrashing_query_t *d_queries, uvmtier_thrashing_hint_t *d_hints,
                                             uint32_t *d_counters);

/* The reference's delayed worker (thrashing_unpin_pages :1822-1906) at time now_ns: pins whose deadline has passed end
 * and their pages' tracking state is cleared (they are no longer in thrashing_pages either).  Removing the remote
 * mappings of such pages (:1892) is the caller's.  d_counters (8 words, zeroed here): [1] pages reset. */
NV_STATUS uvmtier_thrashing_unpin_expired_device(uvmtier_ctx_t *ctx, uvmtier_thrashing_t *t, uint64_t now_ns, uint32_t *d_counters);

/* uvmtier_service_fault_batches_device with the hint inside the fault loop, where the reference asks for it
 * (service_fault_batch_block_locked, UVM/uvm_gpu_replayable_faults.c:1524-1545): a fault that passed the duplicate,
 * permission and already-authorised checks calls uvm_perf_thrashing_get_hint(va_block, address, this GPU) -- requester =
 * layout->gpu_id, residency read from the state table, preferred location from the table, time = d_batch_time_ns[batch]
 * (NV_GETTIME() when the batch is serviced) -- and
 *   THROTTLE   the fault is marked (mark_fault_throttled :1537 -> UVMTIER_SF_THROTTLED) and not serviced;
 *   PIN        the page joins thrashing_pin_mask (UVMTIER_SF_PINNED) and uvm_va_block_select_residency takes the pin's
 *              residency (UVM/uvm_va_block.c:11624-11628): on this GPU the fault is serviced as ever; elsewhere (the CPU on a
 *              PCIe box) it is marked UVMTIER_SF_PIN_REMOTE and left to the caller -- this pass moves pages to the faulting
 *              GPU only, it does not service a GPU fault by making the page resident on another processor and mapping it
 *              remotely.  The hint's state change (the pin) is recorded either way;
 *   and the prefetcher leaves the va_block's thrashing pages alone (UVM/uvm_perf_prefetch.c:378-381, 407).
 * Only va_blocks the detector tracks page by page (uvmtier_thrashing_block_t.tracked) are affected: they are serviced by
 * the warp-per-block kernel one entry at a time; everything else takes the regular path.  The detector is fed by the
 * caller (uvmtier_thrashing_note_events_device with the migrations the records / copy runs describe), as the reference's
 * is by its perf events.  d_thrashing_counters (8 words, zeroed here): [3] throttles, [4] pins on this GPU, [5] pins
 * elsewhere.  Needs uvmtier_thrashing_enable_hints; layout->gpu_id < 32. */
NV_STATUS uvmtier_service_fault_batches_thrashing_device(uvmtier_ctx_t *ctx, uint32_t n_batches,
                                                         const uvmtier_ordered_fault_t *d_ordered,
                                                         const uvmtier_fault_group_t *d_groups, uvmtier_fault_batch_t *d_batches,
                                                         const uvmtier_va_layout_t *layout, const uvmtier_block_state_t *state,
                                                         uvmtier_thrashing_t *thrashing, const uint64_t *d_batch_time_ns,
                                                         uint16_t *d_entry_flags, uvmtier_service_group_t *d_service,
                                                         uvmtier_copy_run_t *d_runs, uint32_t run_cap, uint32_t *d_counters,
                                                         uint32_t *d_thrashing_counters);

/* ------------------------------------------------------------------------------------------ */
/* (1g) Non-replayable faults -- UVM/uvm_gpu_non_replayable_faults.c                           */
/* ------------------------------------------------------------------------------------------ */
/* The other fault front end on the same back end: faults of engines that cannot replay (copy engines, host) arrive
 * in a buffer RM owns; UVM receives a copy of the pending packets (the "shadow buffer",
 * fetch_non_replayable_fault_buffer_entries :179-236), parses them with the same packet parser
 * (uvm_hal_volta_fault_buffer_parse_non_replayable_entry, UVM/uvm_volta_fault_buffer.c:436-445) and services them
 * one by one in buffer order (uvm_parent_gpu_service_non_replayable_fault_buffer :777-802 -> service_fault :757 ->
 * service_fault_once :606-712 -> service_managed_fault_in_block_locked :345-445): no coalescing, no sort, no duplicate
 * or already-authorised shortcut -- logical permission check, new_residency = the one page, uvm_va_block_service_locked.
 *
 *   d_shadow_buffer   n_faults packets (NVC369 layout; the VALID bit is not looked at, nothing is written back)
 *   d_entries         optional [n_faults] parsed entries (num_instances 1, access_type_mask = the one bit)
 *   d_entry_flags     [ceil(n_faults / B) * B], B = max_batch_size of uvmtier_fault_buffer_init: per fault
 *                     UVMTIER_SF_SERVICED | access type, or UVMTIER_SF_FATAL | reason (UvmEventFatalReason: 1 address
 *                     in no va_range, :570-574 with no ATS; 2 logical permissions, :383-387; 3 fatal fault type,
 *                     :221-224), or 0 for a fault whose channel is gone (:623-632, dropped)
 *   d_service         [ceil(n_faults / B) * B] record i belongs to fault i (all zero for dropped and fatal-type faults)
 *   d_runs / d_counters   as uvmtier_service_fault_batches_device
 * Faults on one va_block are applied in buffer order; different va_blocks run in parallel.  What the reference
 * does after servicing -- clearing the channel's faulted bit (:328-343) or killing the channel (:500-533) -- is
 * hardware control and stays with the caller, which reads d_entry_flags. */
NV_STATUS uvmtier_service_non_replayable_faults_device(uvmtier_ctx_t *ctx, const uvmtier_fault_packet_t *d_shadow_buffer,
                                                       uint32_t n_faults, const uvmtier_va_layout_t *layout,
                                                       const uvmtier_block_state_t *state,
                                                       uvmtier_fault_buffer_entry_t *d_entries, uint16_t *d_entry_flags,
                                                       uvmtier_service_group_t *d_service, uvmtier_copy_run_t *d_runs,
                                                       uint32_t run_cap, uint32_t *d_counters);

/* ------------------------------------------------------------------------------------------ */
/* (2) Prefetch hint -- UVM/uvm_perf_prefetch.c                                                */
/* ------------------------------------------------------------------------------------------ */

/* Per-block inputs of uvm_perf_prefetch_get_hint_va_block that the reference reads out of
 * uvm_va_block_t / uvm_va_policy_t / uvm_va_space_t (UVM/uvm_perf_prefetch.c:327-413,447-502). */
#define UVMTIER_PF_HAS_RESIDENT_MASK       0x01u /* va_block->gpus[dst] != NULL (:357); always true for a CPU dst */
#define UVMTIER_PF_BLOCK_RESIDENT_EMPTY    0x02u /* uvm_processor_mask_empty(&va_block->resident) (:363) */
#define UVMTIER_PF_HAS_THRASHING           0x04u /* uvm_perf_thrashing_get_thrashing_pages(va_block) != NULL */
#define UVMTIER_PF_GPU_VA_SPACE_REGISTERED 0x08u /* dst GPU is in va_space->registered_gpu_va_spaces (:261) */

typedef struct {
    uint64_t start;              /* va_block->start (page aligned; need not be 64K aligned)          */
    uint32_t big_page_size;      /* uvm_va_block_gpu_big_page_size(): 65536 (or 131072)              */
    uint16_t num_pages;          /* uvm_va_block_num_cpu_pages(va_block), 1..512                     */
    uint16_t faulted_first;      /* faulted_region.first                                             */
    uint16_t faulted_outer;      /* faulted_region.outer                                             */
    uint16_t new_residency;      /* uvm_processor_id_t value                                         */
    uint16_t preferred_location; /* policy->preferred_location value, UVMTIER_ID_INVALID if unset    */
    uint16_t flags;              /* UVMTIER_PF_*                                                     */
} uvmtier_prefetch_block_t;      /* 24 bytes */


/* Batched uvm_perf_prefetch_compute_ats (UVM/uvm_perf_prefetch.h:76-82, .c:425-445): the pure
 * masks-in / mask-out form.  For block i: faulted pages d_faulted_pages[i], faulted region
 * d_faulted_region[i], window d_max_prefetch_region[i], residency d_residency_mask[i]
 * (d_residency_mask may be NULL = no residency information, like the reference's NULL argument).
 * Writes d_out_prefetch_mask[i].  The bitmap tree is scratch in the reference and lives in
 * registers here. */
NV_STATUS uvmtier_perf_prefetch_compute_ats_device(uvmtier_ctx_t *ctx, size_t n_blocks,
                                                   const uvmtier_page_mask_t *d_faulted_pages,
                                                   const uvmtier_va_block_region_t *d_faulted_region,
                                                   const uvmtier_va_block_region_t *d_max_prefetch_region,
                                                   const uvmtier_page_mask_t *d_residency_mask,
                                                   uvmtier_page_mask_t *d_out_prefetch_mask);
NV_STATUS uvmtier_perf_prefetch_compute_ats(uvmtier_ctx_t *ctx, size_t n_blocks,
                                            const uvmtier_page_mask_t *h_faulted_pages,
                                            const uvmtier_va_block_region_t *h_faulted_region,
                                            const uvmtier_va_block_region_t *h_max_prefetch_region,
                                            const uvmtier_page_mask_t *h_residency_mask,
                                            uvmtier_page_mask_t *h_out_prefetch_mask);

/* Batched uvm_perf_prefetch_get_hint_va_block (UVM/uvm_perf_prefetch.h:92-98, .c:447-502) for
 * n independent managed va_blocks.  d_resident[i] = resident mask of new_residency on block i
 * (uvm_va_block_resident_mask_get, ignored when UVMTIER_PF_HAS_RESIDENT_MASK is clear);
 * d_thrashing may be NULL (no block has thrashing info); d_cpu_read_mapped =
 * va_block->cpu.pte_bits[UVM_PTE_BITS_CPU_READ], only read for a CPU destination, may be NULL
 * (treated as empty).  d_info is updated in place.  Outputs: uvm_perf_prefetch_hint_t
 * {prefetch_pages_mask -> d_out_prefetch_mask[i], residency -> d_out_residency[i]}.
 * Range-group clipping (.c:484-494) is not modelled: the harness has no range groups. */
NV_STATUS uvmtier_perf_prefetch_get_hint_va_block_device(uvmtier_ctx_t *ctx, size_t n_blocks,
                                                         const uvmtier_prefetch_block_t *d_blocks,
                                                         const uvmtier_page_mask_t *d_resident,
                                                         const uvmtier_page_mask_t *d_faulted_pages,
                                                         const uvmtier_page_mask_t *d_thrashing,
                                                         const uvmtier_page_mask_t *d_cpu_read_mapped,
                                                         uvmtier_prefetch_info_t *d_info,
                                                         uvmtier_page_mask_t *d_out_prefetch_mask,
                                                         uint32_t *d_out_residency);
NV_STATUS uvmtier_perf_prefetch_get_hint_va_block(uvmtier_ctx_t *ctx, size_t n_blocks,
                                                  const uvmtier_prefetch_block_t *h_blocks,
                                                  const uvmtier_page_mask_t *h_resident,
                                                  const uvmtier_page_mask_t *h_faulted_pages,
                                                  const uvmtier_page_mask_t *h_thrashing,
                                                  const uvmtier_page_mask_t *h_cpu_read_mapped,
                                                  uvmtier_prefetch_info_t *h_info,
                                                  uvmtier_page_mask_t *h_out_prefetch_mask,
                                                  uint32_t *h_out_residency);

/* ------------------------------------------------------------------------------------------ */
/* (3) Page-list migration between tiers -- UVM/uvm_va_block.c block_copy_* / make_resident    */
/* ------------------------------------------------------------------------------------------ */

/* Memory tiers of one context.  A tier is a contiguous arena; a va_block slot s is backed in every tier at
 * byte offset s * 2 MB (block-contiguous backing: is_block_phys_contig() true on both sides, so consecutive
 * pages coalesce into one copy, UVM/uvm_va_block.c:4278-4288). */
#define UVMTIER_TIER_HBM    0u   /* this context's GPU memory                                              */
#define UVMTIER_TIER_DRAM   1u   /* host DRAM (CPU NUMA node with CPUs), pinned + mapped                    */
#define UVMTIER_TIER_CXL    2u   /* CXL memory = CPU-less NUMA node; emulated on a DRAM node when absent    */
#define UVMTIER_TIER_PEER0  8u   /* UVMTIER_TIER_PEER0 + i: HBM of another GPU, mapped through NVLink/NVSwitch */
#define UVMTIER_MAX_TIERS   24u

typedef struct {
    uint64_t bytes;
    void    *device_ptr;         /* address the kernels use                                                */
    void    *host_ptr;           /* CPU address (host tiers), NULL for HBM / peer                          */
    int32_t  numa_node;          /* node the pages were bound to, -1 = not bound                           */
    uint32_t flags;              /* UVMTIER_TIERF_*                                                        */
} uvmtier_tier_info_t;
#define UVMTIER_TIERF_CXL_EMULATED 0x1u  /* no CPU-less NUMA node on this machine: a DRAM node stands in   */
#define UVMTIER_TIERF_HUGEPAGES    0x2u
#define UVMTIER_TIERF_EXTERNAL     0x4u  /* memory owned by the caller (uvmtier_tier_register)             */
#define UVMTIER_TIERF_IPC          0x8u  /* peer HBM of another process, opened from a CUDA IPC handle      */

/* NUMA probe of /sys/devices/system/node: out_dram_node = a node with CPUs and memory, out_cxl_node = a
 * node with memory and no CPUs (-1 if none).  Returns the number of online nodes. */
int uvmtier_numa_probe(int *out_dram_node, int *out_cxl_node);

/* Creates tier `tier` with `bytes` of backing store.  HBM: cudaMalloc.  DRAM/CXL: anonymous mmap bound to
 * `numa_node` (-1: probe; the reference allocates with alloc_pages_node(nid, __GFP_THISNODE),
 * UVM/uvm_pmm_sysmem.c:829-853) and registered with CUDA as mapped pinned memory so that SMs can address it. */
NV_STATUS uvmtier_tier_create(uvmtier_ctx_t *ctx, uint32_t tier, uint64_t bytes, int numa_node, uvmtier_tier_info_t *out);
/* Registers caller-owned memory as a tier: device_ptr must be addressable by this context's GPU (peer HBM
 * after cudaDeviceEnablePeerAccess, or host memory already registered as mapped). */
NV_STATUS uvmtier_tier_register(uvmtier_ctx_t *ctx, uint32_t tier, void *device_ptr, void *host_ptr, uint64_t bytes);
/* Maps the HBM tier of `peer_ctx` (another GPU of this process) as tier UVMTIER_TIER_PEER0 + peer_index of
 * `ctx`, enabling peer access both ways (the reference: UVM_GPU_LINK_NVLINK_5 peer apertures, UVM/uvm_gpu.h:611-640). */
NV_STATUS uvmtier_tier_map_peer(uvmtier_ctx_t *ctx, uint32_t peer_index, uvmtier_ctx_t *peer_ctx);
/* Same across processes (one process per GPU): export this context's HBM tier as a 64-byte CUDA IPC handle,
 * import a peer's handle as tier UVMTIER_TIER_PEER0 + i (bytes = size of the peer's arena). */
NV_STATUS uvmtier_tier_ipc_export(uvmtier_ctx_t *ctx, uint32_t tier, uint8_t handle[64]);
NV_STATUS uvmtier_tier_ipc_import(uvmtier_ctx_t *ctx, uint32_t tier, const uint8_t handle[64], uint64_t bytes);
NV_STATUS uvmtier_tier_info(uvmtier_ctx_t *ctx, uint32_t tier, uvmtier_tier_info_t *out);
NV_STATUS uvmtier_tier_destroy(uvmtier_ctx_t *ctx, uint32_t tier);

/* The copy primitive: replaces block_copy_pages -> block_copy_push -> ce_hal->memcopy
 * (UVM/uvm_va_block.c:3863-3991, UVM/uvm_volta_ce.c:119-183) and, for CPU<->CPU, the kmap+memcpy page loop
 * (:3972-3984).  Copies n pages of page_size bytes (4096, 65536 or 2097152): page i from
 * src_tier + d_src_off[i] to dst_tier + d_dst_off[i].  Offsets are bytes, page_size aligned.  SM-driven:
 * persistent CTAs stage 4 KB pieces through shared memory with 1-D TMA bulk copies.
 * method: 0 = TMA bulk pipeline (default), 1 = plain 16-byte ld/st (for comparison). */
NV_STATUS uvmtier_copy_page_list_device(uvmtier_ctx_t *ctx, uint32_t dst_tier, uint32_t src_tier, size_t n,
                                        const uint64_t *d_dst_off, const uint64_t *d_src_off, uint32_t page_size,
                                        int method);
/* Host-list form (lists are copied to the device first). */
NV_STATUS uvmtier_copy_page_list(uvmtier_ctx_t *ctx, uint32_t dst_tier, uint32_t src_tier, size_t n,
                                 const uint64_t *h_dst_off, const uint64_t *h_src_off, uint32_t page_size, int method);

/* Executes the copy runs produced by the servicing / make_resident passes: run r moves pages
 * [first_page, first_page + num_pages) of slot r.slot from the tier r.src names (DRAM, CXL, HBM) to the tier
 * r.dst names, at the same block offset (slot * 2 MB + page * 4 KB in every arena); UVMTIER_RUN_SRC_ZERO runs
 * zero-fill the destination (block_populate_gpu_chunk zeroes new chunks, UVM/uvm_va_block.c:3011; CPU pages
 * :1977-1983).  DRAM <-> CXL runs replace the reference's kmap + memcpy page loop (:3972-3984) with the same SM
 * copy kernel reading and writing host-mapped memory.  n_runs is read from the device counter d_n_runs (clamped
 * to run_cap).  Runs are grouped by (src, dst) pair; the pairs execute one after the other on the stream. */
NV_STATUS uvmtier_copy_runs_device(uvmtier_ctx_t *ctx, const uvmtier_copy_run_t *d_runs, const uint32_t *d_n_runs,
                                   uint32_t run_cap, int method);

/* 64-bit checksum of n pages of a tier (order-sensitive FNV-style mix per page, one value per page):
 * lets the harness verify migrated contents without copying them back. */
NV_STATUS uvmtier_checksum_pages_device(uvmtier_ctx_t *ctx, uint32_t tier, size_t n, const uint64_t *d_off,
                                        uint32_t page_size, uint64_t *d_out);

/* Link-peak probe: best-of-`reps` time of one cudaMemcpyAsync (the copy engine, i.e. what the reference's
 * ce_hal->memcopy does) of `bytes` from the start of src_tier to the start of dst_tier.  Defines the
 * "measured link peak" the copy kernels are compared with. */
NV_STATUS uvmtier_tier_memcpy_peak(uvmtier_ctx_t *ctx, uint32_t dst_tier, uint32_t src_tier, uint64_t bytes, int reps,
                                   float *out_best_ms);

/* ------------------------------------------------------------------------------------------ */
/* (3b) The fork's CXL P2P DMA surface -- src/nvidia/src/kernel/mem_mgr/p2p_cxl.c              */
/* ------------------------------------------------------------------------------------------ */
/* What CXLMemUring added to the driver as a second front door to the copy path: register a CXL-backed buffer,
 * ask for DMA between it and GPU memory, unregister (RM controls NV2080_CTRL_CMD_BUS_REGISTER_CXL_BUFFER /
 * _CXL_P2P_DMA_REQUEST / _UNREGISTER_CXL_BUFFER, src/common/sdk/nvidia/inc/ctrl/ctrl2080/ctrl2080bus.h:1400-1560;
 * user-space flow tests/cxl_p2p_test.c:532-632).  Same validation rules, limits and status codes; the copy is
 * this library's TMA page-copy pipeline instead of memmgrMemCopy on a copy engine.
 *
 * Differences a user-space library cannot avoid: the buffer is pinned with cudaHostRegister (the reference pins
 * with pin_user_pages, kernel-open/nvidia/nv-p2p.c:1140-1260), the "physical" page array holds the pages' host
 * virtual addresses, and gpuOffset is an offset into this context's HBM tier (the reference: an FB offset). */
#define UVMTIER_CXL_P2P_PAGE_SIZE_4K          4096u          /* p2p_cxl.c:38-39 */
#define UVMTIER_CXL_P2P_PAGE_SIZE_2M          (2u << 20)
#define UVMTIER_CXL_P2P_MAX_BUFFER_SIZE       (1ull << 40)   /* :218 */
#define UVMTIER_CXL_P2P_MAX_REGISTERED_BUFFERS 256u          /* :221 */
#define UVMTIER_CXL_P2P_DMA_FLAG_CXL_TO_GPU   0x1u           /* :35; clear = GPU to CXL */
#define UVMTIER_CXL_P2P_DMA_FLAG_ASYNC        0x2u           /* ctrl2080bus.h:1495: enqueue only */
#define UVMTIER_CXL_P2P_DMA_MAX_TRANSFER      0xFFFFF000ull  /* a request above 4 GB - 1 is clamped to this (:617-622) */
typedef struct uvmtier_cxl_buffer uvmtier_cxl_buffer_t;

/* RmP2PGetCxlSystemInfo (:99-137): CXL memory devices = CPU-less NUMA nodes with memory; version 2 when none is
 * found (the reference's default when enumeration fails).  Any out pointer may be NULL. */
NV_STATUS uvmtier_p2p_get_cxl_system_info(uint32_t *num_devices, uint32_t *num_memory_devices, uint32_t *link_up,
                                          uint32_t *cxl_version);
/* RmP2PRegisterCxlBuffer (:225-352).  NV_ERR_INVALID_ARGUMENT: size 0, NULL out, version outside 1..3, size above
 * 1 TB, base + size overflow; NV_ERR_INSUFFICIENT_RESOURCES: 256 buffers registered already.  2 MB pages when base
 * and size are 2 MB multiples (_cxlP2PCanUseHugePages :149-160), else 4 KB pages covering [base, base + size). */
NV_STATUS uvmtier_p2p_register_cxl_buffer(uvmtier_ctx_t *ctx, void *base, uint64_t size, uint32_t cxl_version,
                                          uvmtier_cxl_buffer_t **out_handle);
/* RmP2PUnregisterCxlBuffer (:361-421): NV_ERR_INVALID_ARGUMENT for NULL, NV_ERR_INVALID_STATE if not registered. */
NV_STATUS uvmtier_p2p_unregister_cxl_buffer(uvmtier_ctx_t *ctx, uvmtier_cxl_buffer_t *handle);
/* RmP2PGetCxlPages (:423-481) / RmP2PPutCxlPages (:483-514): the page array of [offset, offset + size). */
NV_STATUS uvmtier_p2p_get_cxl_pages(uvmtier_cxl_buffer_t *handle, uint64_t offset, uint64_t size, uint64_t *page_addrs,
                                    uint32_t *page_count, uint32_t *page_size);
NV_STATUS uvmtier_p2p_put_cxl_pages(uvmtier_cxl_buffer_t *handle, uint64_t offset, uint64_t size);
/* RmP2PCxlDmaRequest (:516-677): copies min(size, UVMTIER_CXL_P2P_DMA_MAX_TRANSFER) bytes between
 * HBM tier + gpu_offset and buffer + cxl_offset.  NV_ERR_INVALID_ARGUMENT: NULL ctx / handle, size 0, range
 * outside the buffer (or outside the HBM tier); NV_ERR_INVALID_STATE: buffer not registered, no HBM tier.
 * Waits for the copy unless UVMTIER_CXL_P2P_DMA_FLAG_ASYNC is set.  *out_bytes (may be NULL) = bytes moved. */
NV_STATUS uvmtier_p2p_cxl_dma_request(uvmtier_ctx_t *ctx, uvmtier_cxl_buffer_t *handle, uint64_t gpu_offset,
                                      uint64_t cxl_offset, uint64_t size, uint32_t flags, uint64_t *out_bytes);
/* Registered buffers / bytes in this process (g_cxlRegisteredBufferCount, g_cxlTotalRegisteredSize :143-144). */
void uvmtier_p2p_cxl_registered(uint32_t *count, uint64_t *bytes);

/* ------------------------------------------------------------------------------------------ */
/* Ops table -- the reference dispatches through per-arch function-pointer tables chosen at GPU */
/* init (uvm_fault_buffer_hal_struct UVM/uvm_hal.h:848-864, uvm_ce_hal_struct :813-838); a      */
/* harness built from the reference's C swaps implementations the same way.   */
/* ------------------------------------------------------------------------------------------ */
typedef struct {
    /* fetch_fault_buffer_entries + preprocess_fault_batch (UVM/uvm_gpu_replayable_faults.c:844, :1134), host buffers */
    NV_STATUS (*fetch_preprocess)(uvmtier_ctx_t *, uvmtier_fault_packet_t *h_ring, uint32_t ring_entries, uint32_t get,
                                  uint32_t put, int clear_valid, uint32_t *h_disposition, uvmtier_ordered_fault_t *h_ordered,
                                  uvmtier_fault_group_t *h_groups, uvmtier_fault_batch_t *h_batches,
                                  uvmtier_fault_buffer_entry_t *h_entries, uint32_t *out_get);
    /* the same on device buffers */
    NV_STATUS (*fetch_preprocess_device)(uvmtier_ctx_t *, uvmtier_fault_packet_t *d_ring, uint32_t ring_entries, uint32_t get,
                                         uint32_t put, int clear_valid, uint32_t *d_disposition,
                                         uvmtier_ordered_fault_t *d_ordered, uvmtier_fault_group_t *d_groups,
                                         uvmtier_fault_batch_t *d_batches, uvmtier_fault_buffer_entry_t *d_entries,
                                         uint32_t *out_get);
    /* service_fault_batch (:2232) down to uvm_va_block_service_locked (UVM/uvm_va_block.c:12307) */
    NV_STATUS (*service)(uvmtier_ctx_t *, uint32_t n_batches, const uvmtier_ordered_fault_t *, const uvmtier_fault_group_t *,
                         uvmtier_fault_batch_t *, const uvmtier_va_layout_t *, const uvmtier_block_state_t *,
                         uint16_t *d_entry_flags, uvmtier_service_group_t *, uvmtier_copy_run_t *, uint32_t run_cap,
                         uint32_t *d_counters);
    /* uvm_perf_prefetch_get_hint_va_block (UVM/uvm_perf_prefetch.h:92-98), host buffers */
    NV_STATUS (*prefetch_get_hint)(uvmtier_ctx_t *, size_t n_blocks, const uvmtier_prefetch_block_t *,
                                   const uvmtier_page_mask_t *resident, const uvmtier_page_mask_t *faulted,
                                   const uvmtier_page_mask_t *thrashing, const uvmtier_page_mask_t *cpu_read_mapped,
                                   uvmtier_prefetch_info_t *, uvmtier_page_mask_t *out_prefetch_mask, uint32_t *out_residency);
    /* uvm_va_block_make_resident (UVM/uvm_va_block.h:770-777), batched */
    NV_STATUS (*make_resident)(uvmtier_ctx_t *, uint32_t n_requests, const uvmtier_make_resident_req_t *,
                               const uvmtier_page_mask_t *d_page_masks, const uvmtier_page_mask_t *d_prefetch_masks,
                               const uvmtier_va_layout_t *, const uvmtier_block_state_t *, uvmtier_make_resident_result_t *,
                               uvmtier_copy_run_t *, uint32_t run_cap, uint32_t *d_counters);
    /* the CE pushes of block_copy_resident_pages_between (UVM/uvm_va_block.c:4069-4377) for a list of runs */
    NV_STATUS (*copy_runs)(uvmtier_ctx_t *, const uvmtier_copy_run_t *d_runs, const uint32_t *d_n_runs, uint32_t run_cap,
                           int method);
    /* ce_hal->memcopy / block_copy_pages (UVM/uvm_va_block.c:3863-3991) for a list of pages */
    NV_STATUS (*copy_page_list)(uvmtier_ctx_t *, uint32_t dst_tier, uint32_t src_tier, size_t n, const uint64_t *h_dst_off,
                                const uint64_t *h_src_off, uint32_t page_size, int method);
    /* uvm_parent_gpu_service_replayable_faults (UVM/uvm_gpu_replayable_faults.c:2906-2960): the whole loop, from a
     * ring in pinned host memory to the serviced batches, chunk by chunk */
    NV_STATUS (*service_replayable_faults)(uvmtier_ctx_t *, const uvmtier_fault_packet_t *h_ring_pinned,
                                           uvmtier_fault_packet_t *d_ring, uint32_t ring_entries, uint32_t get, uint32_t put,
                                           int clear_valid, uint32_t chunk_batches, uint32_t *d_disposition,
                                           uvmtier_ordered_fault_t *, uvmtier_fault_group_t *, uvmtier_fault_batch_t *,
                                           const uvmtier_va_layout_t *, const uvmtier_block_state_t *, uint16_t *d_entry_flags,
                                           uvmtier_service_group_t *, uvmtier_copy_run_t *, uint32_t run_cap,
                                           uint32_t *d_counters, const uvmtier_service_readback_t *, uint32_t *out_get);
    /* uvm_access_counter_buffer_hal_struct::parse_entry + fetch_access_counter_buffer_entries + preprocess_notifications
     * (UVM/uvm_hal.h:866-885, UVM/uvm_gpu_access_counters.c:862-992) */
    NV_STATUS (*access_counters_fetch_preprocess)(uvmtier_ctx_t *, uvmtier_access_counter_packet_t *d_ring, uint32_t ring_entries,
                                                  uint32_t get, uint32_t put, int clear_valid, uvmtier_ordered_notification_t *,
                                                  uvmtier_fault_group_t *, uvmtier_fault_batch_t *,
                                                  uvmtier_access_counter_entry_t *d_entries, uint32_t *out_get);
    /* service_notifications_batch (UVM/uvm_gpu_access_counters.c:1513-1615) */
    NV_STATUS (*access_counters_service)(uvmtier_ctx_t *, uint32_t n_batches, const uvmtier_ordered_notification_t *,
                                         const uvmtier_fault_group_t *, uvmtier_fault_batch_t *, const uvmtier_va_layout_t *,
                                         const uvmtier_block_state_t *, uvmtier_service_group_t *, uvmtier_copy_run_t *,
                                         uint32_t run_cap, uint32_t *d_counters);
    /* uvm_parent_gpu_service_non_replayable_fault_buffer (UVM/uvm_gpu_non_replayable_faults.c:777-802) */
    NV_STATUS (*service_non_replayable_faults)(uvmtier_ctx_t *, const uvmtier_fault_packet_t *d_shadow_buffer, uint32_t n_faults,
                                               const uvmtier_va_layout_t *, const uvmtier_block_state_t *,
                                               uvmtier_fault_buffer_entry_t *d_entries, uint16_t *d_entry_flags,
                                               uvmtier_service_group_t *, uvmtier_copy_run_t *, uint32_t run_cap, uint32_t *d_counters);
} uvmtier_ops_t;

/* The B200 implementation of every slot. */
extern const uvmtier_ops_t uvmtier_gpu_ops;

#ifdef __cplusplus
}
#endif
#endif /* UVMTIER_H */
