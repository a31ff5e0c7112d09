/*
 * ref_faults_glue.c -- TEST INFRASTRUCTURE ONLY (oracle/_ref build, TU 1 of 3).
 *
 * Compiles the reference's fault-fetch / coalesce / sort code -- line-range extracts of
 * kernel-open/nvidia-uvm/uvm_gpu_replayable_faults.c (:695-740 comparators, :742-986 fetch+merge,
 * :988-1020 sort comparators, :1118-1212 preprocess + duplicate check) and the Volta+ packet
 * parser of uvm_volta_fault_buffer.c (:105-152, :199-231, :312-391) -- verbatim, against minimal
 * stand-ins for uvm_parent_gpu_t & co (restated from uvm_gpu.h:62-84,286-339,347-418).
 * The Blackwell/Hopper HAL overrides live in their own TUs (their dev_fault.h macros clash).
 * Extracts are produced by oracle/build_ref.sh in a temp dir and never enter the repo.
 */
#include "shim/kshim.h"
#include "nvmisc.h"
#include "uvm_types.h"

#include "x_push_macros.h"

/* ---- forward decls the extracts need ---- */
typedef struct uvm_va_space_struct uvm_va_space_t;
typedef struct uvm_gpu_struct uvm_gpu_t;
typedef struct uvm_parent_gpu_struct uvm_parent_gpu_t;
typedef struct uvm_fault_buffer_entry_struct uvm_fault_buffer_entry_t;
typedef struct uvm_fault_service_batch_context_struct uvm_fault_service_batch_context_t;
typedef struct uvm_va_block_struct uvm_va_block_t;

#include "x_hal_types.h"

typedef struct { NvU32 val; } uvm_processor_id_t;
typedef uvm_processor_id_t uvm_gpu_id_t;
static NvU32 uvm_id_value(uvm_processor_id_t id) { return id.val; }

struct uvm_va_space_struct { int index; int perf_events; };
struct uvm_gpu_struct { uvm_gpu_id_t id; uvm_parent_gpu_t *parent; };

typedef struct {
    NvU32 num_pending_faults;
    bool has_fatal_faults;
    uvm_fault_buffer_entry_t *last_fault;
} uvm_fault_utlb_info_t;

struct uvm_fault_service_batch_context_struct {
    uvm_fault_buffer_entry_t *fault_cache;
    uvm_fault_buffer_entry_t **ordered_fault_cache;
    uvm_fault_utlb_info_t *utlbs;
    NvU32 max_utlb_id;
    NvU32 num_cached_faults;
    NvU32 num_coalesced_faults;
    uvm_va_space_t *fatal_va_space;
    uvm_gpu_t *fatal_gpu;
    bool has_throttled_faults;
    NvU32 num_invalid_prefetch_faults;
    NvU32 num_duplicate_faults;
    NvU32 num_replays;
    NvU32 batch_id;
    bool is_single_instance_ptr;
    uvm_fault_buffer_entry_t *last_fault;
};

typedef struct { NvU32 batch_id; } uvm_non_replayable_fault_buffer_t;      /* the one field the non-replayable leg touches */
typedef struct uvm_replayable_fault_buffer_struct {
    NvU32 max_faults;
    NvU32 cached_get;
    NvU32 cached_put;
    NvU32 utlb_count;
    struct uvm_service_block_context_struct *block_service_context_ptr;   /* used by ref_service_glue.c only */
    uvm_non_replayable_fault_buffer_t *non_replayable_ptr;                /* used by ref_nonreplayable_glue.c only */
} uvm_replayable_fault_buffer_t;

typedef struct {
    NvU32 (*read_put)(uvm_parent_gpu_t *);
    bool (*entry_is_valid)(uvm_parent_gpu_t *, NvU32);
    void (*entry_clear_valid)(uvm_parent_gpu_t *, NvU32);
    NV_STATUS (*parse_replayable_entry)(uvm_parent_gpu_t *, NvU32, uvm_fault_buffer_entry_t *);
    uvm_fault_type_t (*get_fault_type)(const NvU32 *);
    uvm_mmu_engine_type_t (*get_mmu_engine_type)(NvU16, uvm_fault_client_type_t, NvU16);
    NvU8 (*get_ve_id)(NvU16, uvm_mmu_engine_type_t);
} uvm_fault_buffer_hal_t;
typedef struct { NvU16 (*mmu_client_id_to_utlb_id)(NvU16); } uvm_arch_hal_t;

struct uvm_parent_gpu_struct {
    bool replayable_faults_supported;
    bool fault_cancel_va_supported;
    NvU32 utlb_per_gpc_count;
    unsigned va_bits;                      /* harness: width for uvm_parent_gpu_canonical_address */
    struct {
        NvU32 max_batch_size;
        uvm_replayable_fault_buffer_t replayable;
        struct { struct { void *bufferAddress; } replayable; } rm_info;
    } fault_buffer;
    struct { struct { int service_lock; } replayable_faults; } isr;
    uvm_fault_buffer_hal_t *fault_buffer_hal;
    uvm_arch_hal_t *arch_hal;
    struct uvm_access_counter_buffer_hal_struct *access_counter_buffer_hal;   /* used by ref_access_counters_front_glue.c only */
    NvU32 hw_put;                          /* harness: what read_put returns */
    NvU32 written_get;                     /* harness: last write_get */
};

static struct { bool conf_computing_enabled; } g_uvm_global;
static unsigned uvm_perf_fault_coalesce = 1;   /* uvm_gpu_replayable_faults.c:115 */
static NV_STATUS uvm_global_get_status(void) { return NV_OK; }
static void write_get(uvm_parent_gpu_t *p, NvU32 get) { p->fault_buffer.replayable.cached_get = get; p->written_get = get; }

/* uvm_parent_gpu_canonical_address (uvm_gpu.c:428-459) with the GPU/CPU width decision folded into
 * va_bits (57 on Hopper+/Blackwell); the tail check there is a no-op once the widths agree. */
static NvU64 uvm_parent_gpu_canonical_address(uvm_parent_gpu_t *p, NvU64 addr)
{
    unsigned s = p->va_bits;
    return (NvU64)((NvS64)(addr << (64 - s)) >> (64 - s));
}

/* Linux lib/sort.c is an unstable heapsort whose source is outside the reference tree, so the order
 * of comparator-equal elements is unpinned.  Canonical tie-break: the elements are
 * pointers into fault_cache, ties go by ascending pointer = ascending fault_cache index. */
static int (*g_sort_cmp)(const void *, const void *);
static int sort_with_tiebreak(const void *a, const void *b)
{
    int r = g_sort_cmp(a, b);
    if (r)
        return r;
    {
        const void *pa = *(void *const *)a, *pb = *(void *const *)b;
        return pa < pb ? -1 : (pa > pb ? 1 : 0);
    }
}
static void sort(void *base, size_t num, size_t size, int (*cmp)(const void *, const void *), void *swap)
{
    (void)swap;
    g_sort_cmp = cmp;
    qsort(base, num, size, sort_with_tiebreak);
}

/* instance pointer -> (va_space, gpu): flat table instead of uvm_parent_gpu_fault_entry_to_va_space */
#include "ref_api.h"
static const ref_instance_ptr_entry_t *g_tbl;
static size_t g_tbl_n;
static uvm_va_space_t g_spaces[65536];
static uvm_gpu_t g_gpus[UVM_MAX_GPUS + 2];
static NV_STATUS translate_instance_ptrs(uvm_parent_gpu_t *parent_gpu, uvm_fault_service_batch_context_t *batch_context);
static void uvm_perf_event_notify_gpu_fault(int *ev, uvm_va_block_t *b, uvm_gpu_id_t id, uvm_processor_id_t pref,
                                            uvm_fault_buffer_entry_t *e, NvU32 batch_id, bool dup)
{ (void)ev; (void)b; (void)id; (void)pref; (void)e; (void)batch_id; (void)dup; }

/* ---- the Volta+ packet parser, verbatim (volta dev_fault.h values, as the kernel builds it) ---- */
#include "hwref/volta/gv100/dev_fault.h"
#include "clc369.h"
typedef struct { NvU8 bufferEntry[NVC369_BUF_SIZE]; } fault_buffer_entry_c369_t;
#define MAX_SUBCONTEXTS 64 /* uvm_volta_fault_buffer.c:90 */
#include "x_volta_parse.h"

/* ---- the fault fetch / merge / sort code, verbatim ---- */
#include "x_faults.h"

/* translate_instance_ptrs (:1028-1116) reduced to its table walk: same "copy from previous if the
 * instance pointer matches" shortcut (:1041-1045); an unknown pointer is NV_ERR_INVALID_CHANNEL ->
 * flush -> NV_WARN_MORE_PROCESSING_REQUIRED (:1097-1107). */
static NV_STATUS translate_instance_ptrs(uvm_parent_gpu_t *parent_gpu, uvm_fault_service_batch_context_t *batch_context)
{
    NvU32 i;
    (void)parent_gpu;
    for (i = 0; i < batch_context->num_coalesced_faults; ++i) {
        uvm_fault_buffer_entry_t *e = batch_context->ordered_fault_cache[i];
        size_t k;
        if (i != 0 && cmp_fault_instance_ptr(e, batch_context->ordered_fault_cache[i - 1]) == 0) {
            e->va_space = batch_context->ordered_fault_cache[i - 1]->va_space;
            e->gpu = batch_context->ordered_fault_cache[i - 1]->gpu;
            continue;
        }
        for (k = 0; k < g_tbl_n; k++) {
            if (g_tbl[k].instance_ptr == e->instance_ptr.address &&
                g_tbl[k].aperture == (NvU8)e->instance_ptr.aperture &&
                (g_tbl[k].ve_id == REF_VE_ID_ANY || g_tbl[k].ve_id == e->fault_source.ve_id))
                break;
        }
        if (k == g_tbl_n)
            return NV_WARN_MORE_PROCESSING_REQUIRED;
        e->va_space = &g_spaces[g_tbl[k].va_space];
        e->gpu = &g_gpus[g_tbl[k].gpu_id];
    }
    return NV_OK;
}

/* ======================= flat harness API ======================= */
extern uvm_fault_type_t uvm_hal_blackwell_fault_buffer_get_fault_type(const NvU32 *fault_entry);
extern uvm_mmu_engine_type_t uvm_hal_blackwell_fault_buffer_get_mmu_engine_type(NvU16, uvm_fault_client_type_t, NvU16);
extern NvU8 uvm_hal_hopper_fault_buffer_get_ve_id(NvU16 mmu_engine_id, uvm_mmu_engine_type_t mmu_engine_type);
extern NvU16 uvm_hal_blackwell_mmu_client_id_to_utlb_id(NvU16 client_id);

static NvU32 hal_read_put(uvm_parent_gpu_t *p) { return p->hw_put; }
/* uvm_hal_pascal_fault_buffer_entry_is_valid / _clear_valid (uvm_pascal_fault_buffer.c:287-318) on
 * the C369 packet: VALID is bit 31 of word 7 in both B069 and C369 layouts. */
static bool hal_entry_is_valid(uvm_parent_gpu_t *p, NvU32 index)
{
    NvU32 *e = get_fault_buffer_entry(p, index);
    return READ_HWVALUE_MW(e, C369, BUF_ENTRY, VALID);
}
static void hal_entry_clear_valid(uvm_parent_gpu_t *p, NvU32 index)
{
    NvU32 *e = get_fault_buffer_entry(p, index);
    WRITE_HWCONST_MW(e, C369, BUF_ENTRY, VALID, FALSE);
}
/* uvm_hal_volta_fault_buffer_parse_replayable_entry (:393-434) without the CC branch */
static NV_STATUS hal_parse_replayable_entry(uvm_parent_gpu_t *p, NvU32 index, uvm_fault_buffer_entry_t *out)
{
    NvU32 *e = get_fault_buffer_entry(p, index);
    parse_fault_entry_common(p, e, out);
    p->fault_buffer_hal->entry_clear_valid(p, index);
    return NV_OK;
}

static uvm_fault_buffer_hal_t g_hal = {
    hal_read_put, hal_entry_is_valid, hal_entry_clear_valid, hal_parse_replayable_entry,
    uvm_hal_blackwell_fault_buffer_get_fault_type, uvm_hal_blackwell_fault_buffer_get_mmu_engine_type,
    uvm_hal_hopper_fault_buffer_get_ve_id,
};
static uvm_arch_hal_t g_arch = { uvm_hal_blackwell_mmu_client_id_to_utlb_id };

static void export_entry(const uvm_fault_buffer_entry_t *e, ref_fault_entry_t *o)
{
    memset(o, 0, sizeof(*o));
    o->fault_address = e->fault_address;
    o->timestamp = e->timestamp;
    o->instance_ptr_address = e->instance_ptr.address;
    o->client_id = e->fault_source.client_id;
    o->mmu_engine_id = e->fault_source.mmu_engine_id;
    o->utlb_id = e->fault_source.utlb_id;
    o->instance_ptr_aperture = e->instance_ptr.aperture;
    o->fault_type = e->fault_type;
    o->fault_access_type = e->fault_access_type;
    o->client_type = e->fault_source.client_type;
    o->mmu_engine_type = e->fault_source.mmu_engine_type;
    o->gpc_id = e->fault_source.gpc_id;
    o->ve_id = e->fault_source.ve_id;
    o->flags = (e->is_fatal ? REF_FE_FATAL : 0) | (e->is_replayable ? REF_FE_REPLAYABLE : 0) |
               (e->is_virtual ? REF_FE_VIRTUAL : 0) | (e->in_protected_mode ? REF_FE_PROTECTED : 0) |
               (e->filtered ? REF_FE_FILTERED : 0);
    o->num_instances = e->num_instances;
    o->access_type_mask = (NvU8)e->access_type_mask;
}

/*
 * Runs fetch_fault_buffer_entries + preprocess_fault_batch over a ring of n_entries packets in
 * batches of max_batch_size, exactly as uvm_parent_gpu_service_replayable_faults (:2906-2960)
 * drives them, minus servicing.  Outputs, per batch b (base = b * max_batch_size):
 *   entries[base + i]   parsed + merge state of fault_cache[i]            (i < num_cached)
 *   ordered[base + j]   fault_cache index of ordered_fault_cache[j]       (j < num_coalesced)
 *   ordered_ctx[base+j] translation-table row the entry resolved to (va_space/gpu)
 *   batch[b]            counters
 * The ring is modified (valid bits cleared), like the reference does.
 */
/* Optional extra output: batch_context->utlbs[u].num_pending_faults after the fetch of batch b, at
 * out[b * utlb_count + u].  NULL switches it off. */
static uint16_t *g_utlb_out;
void ref_faults_set_utlb_output(uint16_t *out) { g_utlb_out = out; }

int ref_faults_fetch_preprocess(uint32_t *ring, uint32_t n_entries, const ref_fault_gpu_params_t *gp,
                                const ref_instance_ptr_entry_t *tbl, size_t tbl_n, ref_fault_entry_t *entries,
                                uint32_t *ordered, uint32_t *ordered_va_space, uint32_t *ordered_gpu,
                                ref_fault_batch_t *batch)
{
    uvm_parent_gpu_t gpu;
    uvm_fault_service_batch_context_t bc;
    NvU32 B = gp->max_batch_size, utlb_count = gp->gpc_count * gp->utlb_per_gpc_count, b = 0;
    size_t i;

    memset(&gpu, 0, sizeof(gpu));
    memset(&bc, 0, sizeof(bc));
    gpu.replayable_faults_supported = true;
    gpu.fault_cancel_va_supported = true;
    gpu.utlb_per_gpc_count = gp->utlb_per_gpc_count;
    gpu.va_bits = gp->va_bits;
    gpu.fault_buffer.max_batch_size = B;
    gpu.fault_buffer.replayable.max_faults = n_entries + 1;   /* ring never wraps in the harness */
    gpu.fault_buffer.replayable.utlb_count = utlb_count;
    gpu.fault_buffer.rm_info.replayable.bufferAddress = ring;
    gpu.fault_buffer_hal = &g_hal;
    gpu.arch_hal = &g_arch;
    gpu.hw_put = n_entries;
    uvm_perf_fault_coalesce = gp->coalesce;
    g_tbl = tbl;
    g_tbl_n = tbl_n;
    for (i = 0; i < 65536; i++) g_spaces[i].index = (int)i;
    for (i = 0; i < UVM_MAX_GPUS + 2; i++) g_gpus[i].id.val = (NvU32)i;

    bc.fault_cache = calloc(B, sizeof(*bc.fault_cache));
    bc.ordered_fault_cache = calloc(B, sizeof(*bc.ordered_fault_cache));
    bc.utlbs = calloc(utlb_count, sizeof(*bc.utlbs));
    bc.max_utlb_id = utlb_count - 1;  /* first fetch clears every uTLB, like a fresh context */

    while (gpu.fault_buffer.replayable.cached_get != n_entries) {
        NvU32 base = gpu.fault_buffer.replayable.cached_get, j;
        NV_STATUS st;
        ref_fault_batch_t *ob = &batch[b];
        memset(ob, 0, sizeof(*ob));
        st = fetch_fault_buffer_entries(&gpu, &bc, FAULT_FETCH_MODE_BATCH_READY);
        if (st != NV_OK || bc.num_cached_faults == 0)
            break;
        ob->num_cached_faults = bc.num_cached_faults;
        ob->num_coalesced_faults = bc.num_coalesced_faults;
        ob->is_single_instance_ptr = bc.is_single_instance_ptr;
        ob->max_utlb_id = bc.max_utlb_id;
        if (g_utlb_out)
            for (j = 0; j < utlb_count; j++)
                g_utlb_out[(size_t)b * utlb_count + j] = (uint16_t)bc.utlbs[j].num_pending_faults;
        st = preprocess_fault_batch(&gpu, &bc);
        ob->status = st;
        for (j = 0; j < bc.num_cached_faults; j++)
            export_entry(&bc.fault_cache[j], &entries[base + j]);
        if (st == NV_OK) {
            for (j = 0; j < bc.num_coalesced_faults; j++) {
                uvm_fault_buffer_entry_t *e = bc.ordered_fault_cache[j];
                ordered[base + j] = (uint32_t)(e - bc.fault_cache);
                ordered_va_space[base + j] = (uint32_t)e->va_space->index;
                ordered_gpu[base + j] = e->gpu->id.val;
            }
        }
        b++;
    }
    free(bc.fault_cache);
    free(bc.ordered_fault_cache);
    free(bc.utlbs);
    return (int)b;
}

/* parse only: F1 decode of n packets (ring is NOT modified) */
void ref_faults_parse(const uint32_t *ring, uint32_t n_entries, const ref_fault_gpu_params_t *gp, ref_fault_entry_t *entries)
{
    uvm_parent_gpu_t gpu;
    NvU32 i;
    memset(&gpu, 0, sizeof(gpu));
    gpu.utlb_per_gpc_count = gp->utlb_per_gpc_count;
    gpu.va_bits = gp->va_bits;
    gpu.fault_buffer.replayable.max_faults = n_entries + 1;
    gpu.fault_buffer.replayable.utlb_count = gp->gpc_count * gp->utlb_per_gpc_count;
    gpu.fault_buffer_hal = &g_hal;
    gpu.arch_hal = &g_arch;
    for (i = 0; i < n_entries; i++) {
        uvm_fault_buffer_entry_t e;
        NvU32 pkt[8];
        memset(&e, 0, sizeof(e));
        memcpy(pkt, ring + 8 * i, 32);
        parse_fault_entry_common(&gpu, pkt, &e);
        e.fault_address = UVM_PAGE_ALIGN_DOWN(e.fault_address);           /* :919 */
        e.is_fatal = (e.fault_type >= UVM_FAULT_TYPE_FATAL);              /* :922 */
        export_entry(&e, &entries[i]);
        entries[i].num_instances = 0;
        entries[i].access_type_mask = 0;
    }
}
