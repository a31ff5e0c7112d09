/*
 * ref_prefetch_glue.c -- TEST INFRASTRUCTURE ONLY (oracle/_ref build).
 *
 * Compiles the reference's kernel-open/nvidia-uvm/uvm_perf_prefetch.c UNMODIFIED, from where it
 * lies under /root/reference, into oracle/_ref/libref_prefetch.so.  The reference's page-mask and
 * region helpers (uvm_va_block.h) and its va_block type definitions (uvm_va_block_types.h) are
 * pulled in as build-time extracts (oracle/build_ref.sh writes them to a temp dir; they never
 * enter the repo).  Everything else the file drags in through the UVM header tree is replaced by
 * include-guard pre-emption plus the minimal stand-ins below.
 *
 * Exported symbols (ref_*) are a flat C harness API so tests/ can drive the reference code with
 * the same SoA arrays the CUDA path takes.  Only tests/ and bench.py's cpu_baseline leg load this.
 */
#include "shim/kshim.h"

/* Pre-empt the heavy headers uvm_perf_prefetch.c includes. */
#define _UVM_LINUX_H
#define __UVM_COMMON_H__
#define __UVM_VA_BLOCK_H__
#define __UVM_VA_RANGE_H__
#define __UVM_PERF_EVENTS_H__
#define __UVM_PERF_MODULE_H__
#define __UVM_KVMALLOC_H__
#define __UVM_TEST_H__
#define __UVM_PROCESSORS_H__
#define __UVM_VA_BLOCK_TYPES_H__

/* uvm_va_block_types.h:36-80 verbatim (UVM_VA_BLOCK_BITS .. uvm_page_mask_t). */
#include "x_va_block_types.h"

/* ---- minimal processor ids (restated from uvm_processors.h:280-349,450-473) ---- */
typedef struct { NvU32 val; } uvm_processor_id_t;
typedef uvm_processor_id_t uvm_gpu_id_t;
#define UVM_ID_CPU_VALUE 0
#define UVM_ID_GPU0_VALUE 1
#define UVM_ID_MAX_GPUS 256
#define UVM_ID_MAX_PROCESSORS (UVM_ID_MAX_GPUS + 1)
#define UVM_ID_CPU ((uvm_processor_id_t){ .val = UVM_ID_CPU_VALUE })
#define UVM_ID_INVALID ((uvm_processor_id_t){ .val = UVM_ID_MAX_PROCESSORS })
static bool uvm_id_equal(uvm_processor_id_t a, uvm_processor_id_t b) { return a.val == b.val; }
#define UVM_ID_IS_CPU(id) uvm_id_equal(id, UVM_ID_CPU)
#define UVM_ID_IS_INVALID(id) uvm_id_equal(id, UVM_ID_INVALID)
#define UVM_ID_IS_VALID(id) (!UVM_ID_IS_INVALID(id))
#define UVM_ID_IS_GPU(id) (!UVM_ID_IS_CPU(id) && !UVM_ID_IS_INVALID(id))
static NvU32 uvm_id_gpu_index(uvm_gpu_id_t id) { return id.val - UVM_ID_GPU0_VALUE; }
typedef struct { DECLARE_BITMAP(bitmap, UVM_ID_MAX_PROCESSORS); } uvm_processor_mask_t;
static bool uvm_processor_mask_empty(const uvm_processor_mask_t *m) { return bitmap_empty(m->bitmap, UVM_ID_MAX_PROCESSORS); }
static bool uvm_processor_mask_test(const uvm_processor_mask_t *m, uvm_processor_id_t id) { return test_bit(id.val, m->bitmap); }

/* ---- minimal va_space / va_block / policy / context (fields uvm_perf_prefetch.c touches) ---- */
typedef struct uvm_gpu_struct { uvm_gpu_id_t id; NvU64 big_page_size; } uvm_gpu_t;
typedef struct uvm_va_space_struct {
    int lock;
    uvm_processor_mask_t registered_gpu_va_spaces;
    struct { bool page_prefetch_enabled; } test;
} uvm_va_space_t;
typedef struct { uvm_processor_id_t preferred_location; } uvm_va_policy_t;
typedef enum { UVM_PTE_BITS_CPU_READ, UVM_PTE_BITS_CPU_WRITE, UVM_PTE_BITS_CPU_MAX } uvm_pte_bits_cpu_t;
typedef struct uvm_va_block_gpu_state_struct { uvm_page_mask_t resident; } uvm_va_block_gpu_state_t;
typedef struct uvm_va_block_struct {
    NvU64 start, end;
    int lock;
    uvm_va_space_t *va_space;
    uvm_va_policy_t policy;
    uvm_processor_mask_t resident;
    struct { uvm_page_mask_t resident; uvm_page_mask_t pte_bits[UVM_PTE_BITS_CPU_MAX]; } cpu;
    uvm_va_block_gpu_state_t *gpus[UVM_ID_MAX_GPUS];
    struct { uvm_processor_id_t last_migration_proc_id; NvU16 fault_migrations_to_last_proc; } prefetch_info;
    const uvm_page_mask_t *thrashing_pages; /* harness-supplied; NULL = no thrashing info */
    NvU64 big_page_size;
} uvm_va_block_t;
typedef struct { uvm_page_mask_t scratch_page_mask; struct { void *vma; } hmm; } uvm_va_block_context_t;
typedef struct { struct { NvU64 start, end; } node; } uvm_range_group_range_t;

static inline NvU64 uvm_va_block_size(uvm_va_block_t *b) { return b->end - b->start + 1; }
static inline size_t uvm_va_block_num_cpu_pages(uvm_va_block_t *b) { return uvm_va_block_size(b) / PAGE_SIZE; }
static uvm_page_index_t uvm_va_block_cpu_page_index(uvm_va_block_t *b, NvU64 addr) { return (addr - b->start) / PAGE_SIZE; }
static inline bool uvm_va_block_is_hmm(uvm_va_block_t *b) { (void)b; return false; }
static inline uvm_va_space_t *uvm_va_block_get_va_space(uvm_va_block_t *b) { return b->va_space; }
typedef struct { uvm_va_block_t *va_block; uvm_va_block_region_t region; } uvm_reverse_map_t;
static inline NvU64 uvm_va_block_cpu_page_address(uvm_va_block_t *b, uvm_page_index_t i) { return b->start + PAGE_SIZE * i; }

/* uvm_va_block.h region + page-mask helpers, verbatim extract (see build_ref.sh for line ranges). */
#pragma GCC diagnostic ignored "-Wunused-function"
#include "x_va_block_helpers.h"

/* ---- one-line stand-ins for callees outside the hot path ---- */
static const uvm_page_mask_t *uvm_perf_thrashing_get_thrashing_pages(uvm_va_block_t *b) { return b->thrashing_pages; }
static const uvm_va_policy_t *uvm_va_policy_get_region(uvm_va_block_t *b, uvm_va_block_region_t r) { (void)r; return &b->policy; }
static uvm_gpu_t g_gpu_stub;
static uvm_gpu_t *uvm_gpu_get(uvm_gpu_id_t id) { g_gpu_stub.id = id; return &g_gpu_stub; }
static NvU64 uvm_va_block_gpu_big_page_size(uvm_va_block_t *b, uvm_gpu_t *g) { (void)g; return b->big_page_size; }
static uvm_va_block_region_t uvm_hmm_get_prefetch_region(uvm_va_block_t *b, void *vma, const uvm_va_policy_t *p, NvU64 a)
{ (void)vma; (void)p; (void)a; return uvm_va_block_region_from_block(b); }
static bool uvm_hmm_check_context_vma_is_valid(uvm_va_block_t *b, void *vma, uvm_va_block_region_t r) { (void)b; (void)vma; (void)r; return true; }
static uvm_page_mask_t *uvm_va_block_resident_mask_get(uvm_va_block_t *b, uvm_processor_id_t id, int nid)
{ (void)nid; return UVM_ID_IS_CPU(id) ? &b->cpu.resident : &b->gpus[uvm_id_gpu_index(id)]->resident; }
/* harness has no range groups: the clipping loop at uvm_perf_prefetch.c:484-494 iterates zero times */
#define uvm_range_group_for_each_range_in(rgr, va_space, start, end) for ((rgr) = NULL; (rgr) != NULL; )
static uvm_va_space_t *uvm_va_space_get(struct file *f) { (void)f; return NULL; }
static void uvm_va_space_down_write(uvm_va_space_t *s) { (void)s; }
static void uvm_va_space_up_write(uvm_va_space_t *s) { (void)s; }
typedef struct { NvU32 policy; } UVM_TEST_SET_PAGE_PREFETCH_POLICY_PARAMS;
enum { UVM_TEST_PAGE_PREFETCH_POLICY_ENABLE, UVM_TEST_PAGE_PREFETCH_POLICY_DISABLE, UVM_TEST_PAGE_PREFETCH_POLICY_MAX };

/* uvm_va_block.c:2177-2220 verbatim: range_big_page_region_all .. uvm_va_block_big_page_region_subset */
#include "x_big_page_region.h"

/* The reference translation unit itself, unmodified. */
#include "uvm_perf_prefetch.c"

/* ======================= flat harness API ======================= */
#include "ref_api.h"

int ref_prefetch_init(unsigned enable, unsigned threshold, unsigned min_faults)
{
    uvm_perf_prefetch_enable = enable;
    uvm_perf_prefetch_threshold = threshold;
    uvm_perf_prefetch_min_faults = min_faults;
    return (int)uvm_perf_prefetch_init();
}

unsigned ref_prefetch_threshold(void) { return g_uvm_perf_prefetch_threshold; }
unsigned ref_prefetch_min_faults(void) { return g_uvm_perf_prefetch_min_faults; }

/* Reference KAT uvm_perf_utils_test.c:682-715 restated against the reference tree code: returns the
 * five counters seen walking leaf->root from `page`. */
int ref_prefetch_tree_traverse(const uint64_t *pages, unsigned leaf_count, unsigned level_count, unsigned offset,
                               unsigned page, uint16_t *counts, uint16_t *firsts, uint16_t *outers)
{
    uvm_perf_prefetch_bitmap_tree_t tree;
    uvm_perf_prefetch_bitmap_tree_iter_t iter;
    NvU16 counter;
    int n = 0;
    memcpy(tree.pages.bitmap, pages, sizeof(tree.pages.bitmap));
    tree.leaf_count = leaf_count;
    tree.level_count = level_count;
    tree.offset = offset;
    uvm_perf_prefetch_bitmap_tree_traverse_counters(counter, &tree, page, &iter) {
        uvm_va_block_region_t r = uvm_perf_prefetch_bitmap_tree_iter_get_range(&tree, &iter);
        counts[n] = counter;
        firsts[n] = r.first;
        outers[n] = r.outer;
        n++;
    }
    return n;
}

void ref_prefetch_compute_ats(size_t n, const uint64_t *faulted, const uint16_t *faulted_region,
                              const uint16_t *max_region, const uint64_t *residency /* nullable */,
                              uint64_t *out_mask)
{
    static uvm_va_space_t space;
    uvm_perf_prefetch_bitmap_tree_t tree;
    size_t i;
    space.test.page_prefetch_enabled = true;
    for (i = 0; i < n; i++) {
        uvm_page_mask_t f, r, o;
        memcpy(f.bitmap, faulted + 8 * i, 64);
        if (residency)
            memcpy(r.bitmap, residency + 8 * i, 64);
        uvm_perf_prefetch_compute_ats(&space, &f,
                                      uvm_va_block_region(faulted_region[2 * i], faulted_region[2 * i + 1]),
                                      uvm_va_block_region(max_region[2 * i], max_region[2 * i + 1]),
                                      residency ? &r : NULL, &tree, &o);
        memcpy(out_mask + 8 * i, o.bitmap, 64);
    }
}

void ref_prefetch_get_hint_va_block(size_t n, const ref_prefetch_block_t *blk, const uint64_t *resident,
                                    const uint64_t *faulted, const uint64_t *thrashing /* nullable */,
                                    const uint64_t *cpu_read_mapped /* nullable */, ref_prefetch_info_t *info,
                                    uint64_t *out_mask, uint32_t *out_residency)
{
    static uvm_va_space_t space;
    static uvm_va_block_t block;
    static uvm_va_block_gpu_state_t gstate;
    static uvm_va_block_context_t ctx;
    uvm_perf_prefetch_bitmap_tree_t tree;
    size_t i;

    space.test.page_prefetch_enabled = true;
    for (i = 0; i < n; i++) {
        const ref_prefetch_block_t *b = &blk[i];
        uvm_processor_id_t dst = { .val = b->new_residency };
        uvm_page_mask_t f, t;
        uvm_perf_prefetch_hint_t hint;

        memset(&block, 0, sizeof(block));
        memset(&space.registered_gpu_va_spaces, 0, sizeof(space.registered_gpu_va_spaces));
        block.va_space = &space;
        block.start = b->start;
        block.end = b->start + (NvU64)b->num_pages * PAGE_SIZE - 1;
        block.policy.preferred_location.val = b->preferred_location;
        block.big_page_size = b->big_page_size;
        block.prefetch_info.last_migration_proc_id.val = info[i].last_migration_proc_id;
        block.prefetch_info.fault_migrations_to_last_proc = info[i].fault_migrations_to_last_proc;
        if (!(b->flags & REF_PF_BLOCK_RESIDENT_EMPTY))
            __set_bit(0, block.resident.bitmap); /* only emptiness is observed (uvm_perf_prefetch.c:363) */
        if (UVM_ID_IS_GPU(dst)) {
            if (b->flags & REF_PF_GPU_VA_SPACE_REGISTERED)
                __set_bit(dst.val, space.registered_gpu_va_spaces.bitmap);
            if (b->flags & REF_PF_HAS_RESIDENT_MASK) {
                memcpy(gstate.resident.bitmap, resident + 8 * i, 64);
                block.gpus[uvm_id_gpu_index(dst)] = &gstate;
            }
        }
        else {
            memcpy(block.cpu.resident.bitmap, resident + 8 * i, 64);
            if (cpu_read_mapped)
                memcpy(block.cpu.pte_bits[UVM_PTE_BITS_CPU_READ].bitmap, cpu_read_mapped + 8 * i, 64);
        }
        memcpy(f.bitmap, faulted + 8 * i, 64);
        if (thrashing && (b->flags & REF_PF_HAS_THRASHING)) {
            memcpy(t.bitmap, thrashing + 8 * i, 64);
            block.thrashing_pages = &t;
        }
        uvm_perf_prefetch_get_hint_va_block(&block, &ctx, dst, &f,
                                            uvm_va_block_region(b->faulted_first, b->faulted_outer), &tree, &hint);
        memcpy(out_mask + 8 * i, hint.prefetch_pages_mask.bitmap, 64);
        out_residency[i] = hint.residency.val;
        info[i].last_migration_proc_id = block.prefetch_info.last_migration_proc_id.val;
        info[i].fault_migrations_to_last_proc = block.prefetch_info.fault_migrations_to_last_proc;
    }
}
