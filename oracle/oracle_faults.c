/*
 * oracle_faults.c -- TEST INFRASTRUCTURE ONLY: CPU restatement of the replayable-fault batch
 * preprocessing of kernel-open/nvidia-uvm/uvm_gpu_replayable_faults.c (UVM = that directory):
 *   F1  packet decode             UVM/uvm_volta_fault_buffer.c:312-391 + Blackwell/Hopper overrides
 *   F2  fetch + sequential merge  uvm_gpu_replayable_faults.c:742-986
 *   F3  preprocess (two sorts)    :988-1020, :1134-1178
 *
 * Pinning: nothing in the reference tree holds known-answer vectors for this stage.
 * The restatement is pinned against oracle/_ref/libref_faults.so -- the reference's own functions
 * compiled verbatim -- on seeded random, adversarial and directed traces; golden vectors made by
 * that library are committed under tests/golden/.  The one unpinned degree of freedom is the order
 * of comparator-equal entries (Linux sort() is an unstable heapsort outside the tree): canonical
 * tie-break = ascending fault_cache index, here, in the shimmed reference and in the CUDA kernel.
 */
#include <stdlib.h>

#include "oracle_common.h"
#include "ref_api.h"

enum { ACC_PREFETCH = 0, ACC_READ, ACC_WRITE, ACC_ATOMIC_WEAK, ACC_ATOMIC_STRONG, ACC_COUNT }; /* uvm_hal_types.h:204-212 */
enum { APERTURE_SYS = 9, APERTURE_VID = 10, APERTURE_MAX = 12 };                               /* uvm_hal_types.h:33-51  */
enum { FT_FATAL = 5, FT_COUNT = 17 };                                                          /* uvm_hal_types.h:275-305 */
enum { ENG_GRAPHICS = 0, ENG_HOST = 1, ENG_CE = 2 };                                           /* uvm_hal_types.h:319-325 */
#define NVSTATUS_WARN_MORE_PROCESSING_REQUIRED 0x10005u

static unsigned bits(const uint32_t *w, unsigned word, unsigned hi, unsigned lo)
{
    return (w[word] >> lo) & (hi - lo == 31 ? 0xffffffffu : ((1u << (hi - lo + 1)) - 1));
}

/* Blackwell get_fault_type, UVM/uvm_blackwell_fault_buffer.c:34-76 (hw values: hwref/blackwell/gb100/dev_fault.h:269-285) */
static unsigned fault_type_blackwell(unsigned hw)
{
    static const uint8_t map[16] = { 0, 5, 1, 6, 7, 8, 3, 4, 9, 10, 11, 16, 13, 14, 15, 2 };
    return hw < 16 ? map[hw] : FT_COUNT;
}

/* get_fault_access_type / is_fault_address_virtual, UVM/uvm_volta_fault_buffer.c:105-152 */
static unsigned access_type(unsigned hw, bool *is_virtual)
{
    *is_virtual = hw < 8;
    switch (hw) {
        case 0: case 8: return ACC_READ;
        case 1: case 9: return ACC_WRITE;
        case 2: case 10: return ACC_ATOMIC_STRONG;
        case 3: case 11: return ACC_PREFETCH;
        case 4: return ACC_ATOMIC_WEAK;
    }
    *is_virtual = true; /* the reference returns UVM_FAULT_ACCESS_TYPE_COUNT (5) as a bool here: true */
    return ACC_COUNT;
}

/* Blackwell client id -> uTLB inside the GPC, UVM/uvm_blackwell_mmu.c:128-201 */
static unsigned client_to_utlb_blackwell(unsigned c)
{
    if (c <= 0x07) return 1 + c;                      /* T1_0..7   -> LTP0..7   */
    if (c >= 0x08 && c <= 0x0f) return 2 + 2 * (c - 0x08); /* PE_0..7 -> LTP1,3,..15 */
    if (c >= 0x10 && c <= 0x12) return 0;             /* RAST, GCC, GPCCS -> RGG */
    if (c >= 0x21 && c <= 0x28) return 9 + (c - 0x21);     /* T1_8..15  -> LTP8..15  */
    if (c >= 0x29 && c <= 0x30) return 2 + 2 * (c - 0x29); /* TPCCS_0..7 -> LTP1,3,..15 */
    if (c == 0x31 || c == 0x32) return 18 + 2 * (c - 0x31); /* PE_8,9   -> LTP17,19  */
    if (c == 0x33 || c == 0x34) return 18 + 2 * (c - 0x33); /* TPCCS_8,9 -> LTP17,19 */
    if (c >= 0x35 && c <= 0x38) return 17 + (c - 0x35);    /* T1_16..19 -> LTP16..19 */
    return 0;                                          /* the reference asserts and returns 0 */
}

/* Blackwell get_mmu_engine_type, UVM/uvm_blackwell_fault_buffer.c:78-122 */
static unsigned engine_type_blackwell(unsigned client_type_hub, unsigned client)
{
    if (client_type_hub) {
        if ((client >= 0x20 && client <= 0x27) || client == 0x01 || client == 0x02 || client == 0x18 || client == 0x0c)
            return ENG_CE;
        if (client == 0x06 || (client >= 0x63 && client <= 0x6e))
            return ENG_HOST;
    }
    return ENG_GRAPHICS;
}

/* parse_fault_entry_common (:312-391) + fetch loop fix-ups (:919-922) */
static void decode(const uint32_t *w, const ref_fault_gpu_params_t *gp, ref_fault_entry_t *e)
{
    bool is_virtual;
    unsigned ap = bits(w, 0, 9, 8), hub, client, eng_id, eng_type;
    uint64_t addr;
    memset(e, 0, sizeof(*e));
    e->instance_ptr_address = (((uint64_t)w[1] << 20) | bits(w, 0, 31, 12)) << 12;
    e->instance_ptr_aperture = ap == 0 ? APERTURE_VID : (ap >= 2 ? APERTURE_SYS : APERTURE_MAX);
    addr = (((uint64_t)w[3] << 20) | bits(w, 2, 31, 12)) << 12;
    addr = (uint64_t)((int64_t)(addr << (64 - gp->va_bits)) >> (64 - gp->va_bits)); /* uvm_gpu.c:428-459 */
    e->fault_address = addr & ~0xfffull;
    e->timestamp = ((uint64_t)w[5] << 32) | w[4];
    e->fault_type = fault_type_blackwell(bits(w, 7, 4, 0));
    e->fault_access_type = access_type(bits(w, 7, 19, 16), &is_virtual);
    hub = bits(w, 7, 20, 20);
    e->client_type = hub;
    client = bits(w, 7, 14, 8);
    e->client_id = client;
    e->gpc_id = bits(w, 7, 28, 24);
    if (!hub)
        e->utlb_id = e->gpc_id * gp->utlb_per_gpc_count + client_to_utlb_blackwell(client);
    eng_id = bits(w, 6, 8, 0);
    e->mmu_engine_id = eng_id;
    eng_type = engine_type_blackwell(hub, client);
    e->mmu_engine_type = eng_type;
    e->ve_id = eng_type == ENG_GRAPHICS ? (uint8_t)(eng_id - 384) : 0;   /* uvm_hopper_fault_buffer.c:31-44 */
    e->flags = (bits(w, 7, 7, 7) ? REF_FE_REPLAYABLE : 0) | (is_virtual ? REF_FE_VIRTUAL : 0) |
               (bits(w, 7, 29, 29) ? REF_FE_PROTECTED : 0) | (e->fault_type >= FT_FATAL ? REF_FE_FATAL : 0);
}

void oracle_faults_parse(const uint32_t *ring, uint32_t n_entries, const ref_fault_gpu_params_t *gp, ref_fault_entry_t *entries)
{
    uint32_t i;
    for (i = 0; i < n_entries; i++)
        decode(ring + 8 * i, gp, &entries[i]);
}

/* cmp_fault_instance_ptr, :695-704: aperture, then address, then ve_id */
static int cmp_inst(const ref_fault_entry_t *a, const ref_fault_entry_t *b)
{
    if (a->instance_ptr_aperture != b->instance_ptr_aperture)
        return a->instance_ptr_aperture < b->instance_ptr_aperture ? -1 : 1;
    if (a->instance_ptr_address != b->instance_ptr_address)
        return a->instance_ptr_address < b->instance_ptr_address ? -1 : 1;
    if (a->ve_id != b->ve_id)
        return a->ve_id < b->ve_id ? -1 : 1;
    return 0;
}

/* fetch_fault_buffer_merge_entry, :753-780.  Returns true when `cur` took over as representative. */
static bool merge(ref_fault_entry_t *cur, ref_fault_entry_t *last)
{
    last->num_instances++;
    last->access_type_mask |= 1u << cur->fault_access_type;
    if (cur->fault_access_type > last->fault_access_type) {
        cur->access_type_mask = last->access_type_mask;
        cur->num_instances = last->num_instances;
        last->flags |= REF_FE_FILTERED;
        return true;
    }
    cur->flags |= REF_FE_FILTERED;
    return false;
}

typedef struct { const ref_fault_entry_t *e; uint32_t idx, va_space, gpu; } sort_item_t;

static int cmp_item_inst(const void *a, const void *b)
{
    const sort_item_t *x = a, *y = b;
    int r = cmp_inst(x->e, y->e);
    return r ? r : (x->idx < y->idx ? -1 : (x->idx > y->idx ? 1 : 0));
}
/* cmp_sort_fault_entry_by_va_space_gpu_address_access_type, :1000-1020; access type DESCENDING (:739) */
static int cmp_item_main(const void *a, const void *b)
{
    const sort_item_t *x = a, *y = b;
    if (x->va_space != y->va_space) return x->va_space < y->va_space ? -1 : 1;
    if (x->gpu != y->gpu) return x->gpu < y->gpu ? -1 : 1;
    if (x->e->fault_address != y->e->fault_address) return x->e->fault_address < y->e->fault_address ? -1 : 1;
    if (x->e->fault_access_type != y->e->fault_access_type) return x->e->fault_access_type > y->e->fault_access_type ? -1 : 1;
    return x->idx < y->idx ? -1 : (x->idx > y->idx ? 1 : 0);
}

/* Optional extra output of oracle_faults_fetch_preprocess: utlbs[u].num_pending_faults of batch b
 * (uvm_gpu.h:62-84; cleared :881, counted :966) at out[b * utlb_count + u].  NULL switches it off. */
static uint16_t *g_utlb_out;
void oracle_faults_set_utlb_output(uint16_t *out) { g_utlb_out = out; }

int oracle_faults_fetch_preprocess(uint32_t *ring, uint32_t n_entries, const ref_fault_gpu_params_t *gp,
                                   const ref_instance_ptr_entry_t *tbl, size_t tbl_n, ref_fault_entry_t *entries,
                                   uint32_t *ordered, uint32_t *ordered_va_space, uint32_t *ordered_gpu,
                                   ref_fault_batch_t *batch)
{
    const uint32_t B = gp->max_batch_size, utlb_count = gp->gpc_count * gp->utlb_per_gpc_count;
    int32_t *utlb_last = malloc(sizeof(int32_t) * utlb_count);
    uint32_t *utlb_pending = malloc(sizeof(uint32_t) * utlb_count);
    sort_item_t *items = malloc(sizeof(sort_item_t) * B);
    uint32_t get = 0, nb = 0;

    while (get != n_entries) {
        ref_fault_entry_t *fc = entries + get;          /* fault_cache of this batch */
        ref_fault_batch_t *ob = &batch[nb];
        uint32_t n = 0, coalesced = 0, max_utlb = 0, i, m;
        int32_t last = -1;
        bool single = true, translate_ok = true;

        memset(ob, 0, sizeof(*ob));
        memset(utlb_pending, 0, sizeof(uint32_t) * utlb_count);

        /* ---- F2: fetch_fault_buffer_entries :844-986 ---- */
        while (get + n != n_entries && n < B) {
            const uint32_t *pkt = ring + 8 * (size_t)(get + n);
            ref_fault_entry_t *cur = &fc[n];
            bool same_inst = true, merged = false;
            uint32_t u;
            if (!(pkt[7] >> 31))
                break;                                   /* BATCH_READY: stop at the first entry not yet valid (:899-903) */
            decode(pkt, gp, cur);
            ring[8 * (size_t)(get + n) + 7] &= 0x7fffffffu; /* entry_clear_valid (:431) */
            u = cur->utlb_id;
            if (u > max_utlb) max_utlb = u;
            if (n > 0) {
                same_inst = cmp_inst(cur, &fc[last]) == 0;                        /* :944 */
                if (gp->coalesce && !(cur->flags & REF_FE_FATAL)) {               /* :947-956 -> :782-820 */
                    ref_fault_entry_t *lt = utlb_pending[u] ? &fc[utlb_last[u]] : NULL;
                    ref_fault_entry_t *lg = &fc[last];
                    bool is_tlb = lt && cmp_inst(cur, lt) == 0 && cur->fault_address == lt->fault_address;
                    bool is_glob = same_inst && cur->fault_address == lg->fault_address &&
                                   cur->fault_access_type <= lg->fault_access_type;
                    if (is_tlb) {
                        if (merge(cur, lt))
                            utlb_last[u] = (int32_t)n;   /* note: the batch-wide `last` is NOT moved (:804-809) */
                        merged = true;
                    }
                    else if (is_glob) {
                        merge(cur, lg);                  /* cannot promote: access type is <= */
                        merged = true;
                    }
                }
            }
            if (!merged) {
                if (single && !same_inst) single = false;                          /* :959-960 */
                cur->num_instances = 1;
                cur->access_type_mask = 1u << cur->fault_access_type;
                utlb_pending[u]++;
                utlb_last[u] = (int32_t)n;
                last = (int32_t)n;
                coalesced++;
            }
            n++;
        }
        if (n == 0)
            break;
        ob->num_cached_faults = n;
        ob->num_coalesced_faults = coalesced;
        ob->is_single_instance_ptr = single;
        ob->max_utlb_id = max_utlb;
        if (g_utlb_out)
            for (i = 0; i < utlb_count; i++)
                g_utlb_out[(size_t)nb * utlb_count + i] = (uint16_t)utlb_pending[i];

        /* ---- F3: preprocess_fault_batch :1134-1178 ---- */
        for (i = 0, m = 0; i < n; i++)
            if (!(fc[i].flags & REF_FE_FILTERED)) { items[m].e = &fc[i]; items[m].idx = i; m++; }
        if (!single)
            qsort(items, m, sizeof(*items), cmp_item_inst);
        for (i = 0; i < m && translate_ok; i++) {        /* translate_instance_ptrs :1028-1116 */
            size_t k;
            if (i && cmp_inst(items[i].e, items[i - 1].e) == 0) {
                items[i].va_space = items[i - 1].va_space; items[i].gpu = items[i - 1].gpu;
                continue;
            }
            for (k = 0; k < tbl_n; k++)
                if (tbl[k].instance_ptr == items[i].e->instance_ptr_address && tbl[k].aperture == items[i].e->instance_ptr_aperture &&
                    (tbl[k].ve_id == REF_VE_ID_ANY || tbl[k].ve_id == items[i].e->ve_id))
                    break;
            if (k == tbl_n) { translate_ok = false; break; }
            items[i].va_space = tbl[k].va_space; items[i].gpu = tbl[k].gpu_id;
        }
        if (!translate_ok) {
            ob->status = NVSTATUS_WARN_MORE_PROCESSING_REQUIRED;   /* stale channel: flush + refetch (:1097-1107) */
        }
        else {
            qsort(items, m, sizeof(*items), cmp_item_main);
            for (i = 0; i < m; i++) {
                ordered[get + i] = items[i].idx;
                ordered_va_space[get + i] = items[i].va_space;
                ordered_gpu[get + i] = items[i].gpu;
            }
        }
        get += n;
        nb++;
    }
    free(utlb_last); free(utlb_pending); free(items);
    return (int)nb;
}
