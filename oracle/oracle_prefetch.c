/*
 * oracle_prefetch.c -- TEST INFRASTRUCTURE ONLY: CPU restatement of the uvm_perf_prefetch
 * bitmap-tree prefetch hint (reference: kernel-open/nvidia-uvm/uvm_perf_prefetch.c).
 *
 * Pinned by: the reference KAT uvm_perf_utils_test.c:682-715 (tests/test_oracle_prefetch.py) and
 * bit-exact agreement with oracle/_ref/libref_prefetch.so (the reference file compiled unmodified)
 * on seeded random + directed inputs; golden vectors from that run live in tests/golden/.
 *
 * It follows the reference literally (walk leaf->root per faulted page); the CUDA kernel uses a
 * different, closed-form formulation, which is exactly why this stays literal.
 */
#include "oracle_common.h"
#include "ref_api.h"

static unsigned g_threshold = 51;  /* uvm_perf_prefetch.c:42  UVM_PREFETCH_THRESHOLD_DEFAULT */
static unsigned g_min_faults = 1;  /* uvm_perf_prefetch.c:51  UVM_PREFETCH_MIN_FAULTS_DEFAULT */
static bool g_enable = true;       /* uvm_perf_prefetch.c:39 */

#define ID_INVALID 257u /* UVM_ID_INVALID: UVM_ID_MAX_PROCESSORS, uvm_processors.h:294,300 */

/* uvm_perf_prefetch_init, uvm_perf_prefetch.c:504-535: out-of-range tunables fall back to defaults */
int oracle_prefetch_init(unsigned enable, unsigned threshold, unsigned min_faults)
{
    g_enable = enable != 0;
    if (!g_enable)
        return 0;
    g_threshold = threshold <= 100 ? threshold : 51;
    g_min_faults = (min_faults >= 1 && min_faults <= 20) ? min_faults : 1;
    return 0;
}
unsigned oracle_prefetch_threshold(void) { return g_threshold; }
unsigned oracle_prefetch_min_faults(void) { return g_min_faults; }

typedef struct {
    or_mask_t pages;
    unsigned offset, leaf_count, level_count;
} tree_t; /* uvm_perf_prefetch_bitmap_tree_t, uvm_perf_prefetch.h:41-50 */

static unsigned ceil_log2(unsigned v) { unsigned l = 0; while ((1u << l) < v) l++; return l; }

/* Node of `leaf` at `level` (root = 0): [start, outer) clamped to leaf_count.
 * uvm_perf_utils.h:165-178 + uvm_perf_prefetch.c:81-100 */
static void node_range(const tree_t *t, unsigned level, unsigned leaf, unsigned *start, unsigned *outer)
{
    unsigned span = 1u << (t->level_count - 1 - level);
    *start = leaf & ~(span - 1);
    *outer = *start + span;
    if (*outer > t->leaf_count)
        *outer = t->leaf_count;
}

/* KAT hook: counters seen from leaf to root */
int oracle_prefetch_tree_traverse(const uint64_t *pages, unsigned leaf_count, unsigned level_count, unsigned offset,
                                  unsigned page, uint16_t *counts, uint16_t *firsts, uint16_t *outers)
{
    tree_t t;
    int n = 0, level;
    or_load(&t.pages, pages);
    t.leaf_count = leaf_count; t.level_count = level_count; t.offset = offset;
    for (level = (int)level_count - 1; level >= 0; level--) {
        unsigned s, o;
        node_range(&t, level, page, &s, &o);
        counts[n] = or_weight_range(&t.pages, s, o);
        firsts[n] = s; outers[n] = o; n++;
    }
    return n;
}

/* compute_prefetch_region, uvm_perf_prefetch.c:102-146.  Returns false for the empty region. */
static bool region_for_page(const tree_t *t, unsigned page, unsigned max_first, unsigned max_outer,
                            unsigned *first, unsigned *outer)
{
    unsigned leaf = page - max_first + t->offset;
    unsigned rf = 0, ro = 0;
    int level;
    for (level = (int)t->level_count - 1; level >= 0; level--) {
        unsigned s, o, cnt;
        node_range(t, level, leaf, &s, &o);
        cnt = or_weight_range(&t->pages, s, o);
        if (cnt * 100 > (o - s) * g_threshold) { rf = s; ro = o; } /* keeps the LAST passing level = largest */
    }
    if (!ro)
        return false;
    /* clamp to real pages (:123-143), quirk included: below-offset maps to `offset`, not to max.first */
    rf += max_first;
    if (rf < t->offset) rf = t->offset;
    else { rf -= t->offset; if (rf < max_first) rf = max_first; }
    ro += max_first;
    if (ro < t->offset) ro = t->offset;
    else { ro -= t->offset; if (ro > max_outer) ro = max_outer; }
    *first = rf; *outer = ro;
    return true;
}

/* compute_prefetch_mask, uvm_perf_prefetch.c:299-319 */
static void prefetch_mask(const tree_t *t, const or_mask_t *faulted, unsigned f_first, unsigned f_outer,
                          unsigned max_first, unsigned max_outer, or_mask_t *out)
{
    unsigned p;
    or_zero(out);
    for (p = f_first; p < f_outer; p++) {
        unsigned rf, ro;
        if (!or_test(faulted, p))
            continue;
        if (!region_for_page(t, p, max_first, max_outer, &rf, &ro))
            continue;              /* region (0,0): fill nothing; outer 0 != max.outer unless empty block */
        if (ro > rf)
            or_fill_range(out, rf, ro);
        if (ro == max_outer)
            break;
    }
}

/* init_bitmap_tree_from_region, uvm_perf_prefetch.c:222-238 */
static void tree_init(tree_t *t, unsigned max_first, unsigned max_outer, const or_mask_t *resident, const or_mask_t *faulted)
{
    or_mask_t m = *faulted;
    if (resident)
        or_or(&m, resident, faulted);
    or_shr(&t->pages, &m, max_first);
    t->offset = 0;
    t->leaf_count = max_outer - max_first;
    t->level_count = ceil_log2(t->leaf_count) + 1;
}

/* grow_fault_granularity_if_no_thrashing, uvm_perf_prefetch.c:148-162 */
static void grow_if_quiet(tree_t *t, unsigned first, unsigned outer, unsigned base,
                          const or_mask_t *faulted, const or_mask_t *thrashing)
{
    if (or_any_range(faulted, first, outer) && (!thrashing || !or_any_range(thrashing, first, outer)))
        or_fill_range(&t->pages, first - base + t->offset, outer - base + t->offset);
}

/* update_bitmap_tree_from_va_block + grow_fault_granularity, uvm_perf_prefetch.c:164-220,240-297;
 * big-page subset: uvm_va_block.c:2177-2189,2202-2220 */
static void tree_update(tree_t *t, uint64_t block_start, unsigned big_pages, unsigned max_first, unsigned max_outer,
                        const or_mask_t *faulted, const or_mask_t *thrashing)
{
    uint64_t big = (uint64_t)big_pages * 4096;
    uint64_t start = block_start + (uint64_t)max_first * 4096;
    uint64_t end1 = block_start + (uint64_t)max_outer * 4096; /* end + 1 */
    uint64_t fa = (start + big - 1) & ~(big - 1);
    uint64_t oa = end1 & ~(big - 1);
    unsigned bf = 0, bo = 0, p;
    if (oa > fa) {
        bf = (unsigned)((fa - start) / 4096) + max_first;
        bo = (unsigned)((oa - start) / 4096) + max_first;
    }
    if (bo && bf > max_first) {
        /* unaligned start: shift so that tree leaves are big-page aligned */
        t->offset = big_pages - (bf - max_first);
        t->leaf_count = (max_outer - max_first) + t->offset;
        or_shl(&t->pages, &t->pages, t->offset);
        t->level_count = ceil_log2(t->leaf_count) + 1;
    }
    if (!bo) {                                       /* no whole big page in the region */
        grow_if_quiet(t, max_first, max_outer, max_first, faulted, thrashing);
        return;
    }
    if (bf > max_first)
        grow_if_quiet(t, max_first, bf, max_first, faulted, thrashing);
    for (p = bf; p < bo; p += big_pages)
        grow_if_quiet(t, p, p + big_pages, max_first, faulted, thrashing);
    if (bo < max_outer)
        grow_if_quiet(t, bo, max_outer, max_first, faulted, thrashing);
}

/* uvm_perf_prefetch_compute_ats, uvm_perf_prefetch.c:425-445 (batched over n independent blocks) */
void oracle_prefetch_compute_ats(size_t n, const uint64_t *faulted, const uint16_t *faulted_region,
                                 const uint16_t *max_region, const uint64_t *residency, uint64_t *out_mask)
{
    size_t i;
    for (i = 0; i < n; i++) {
        or_mask_t f, r, o;
        tree_t t;
        or_load(&f, faulted + 8 * i);
        if (residency) or_load(&r, residency + 8 * i);
        or_zero(&o);
        if (g_enable) {
            tree_init(&t, max_region[2 * i], max_region[2 * i + 1], residency ? &r : NULL, &f);
            prefetch_mask(&t, &f, faulted_region[2 * i], faulted_region[2 * i + 1],
                          max_region[2 * i], max_region[2 * i + 1], &o);
        }
        or_store(out_mask + 8 * i, &o);
    }
}

/* uvm_perf_prefetch_get_hint_va_block (:447-502) over uvm_perf_prefetch_prenotify_fault_migrations
 * (:327-413), batched over n independent managed (non-HMM) blocks, no range groups. */
void oracle_prefetch_get_hint_va_block(size_t n, const ref_prefetch_block_t *blk, const uint64_t *resident,
                                       const uint64_t *faulted, const uint64_t *thrashing,
                                       const uint64_t *cpu_read_mapped, ref_prefetch_info_t *info,
                                       uint64_t *out_mask, uint32_t *out_residency)
{
    size_t i;
    for (i = 0; i < n; i++) {
        const ref_prefetch_block_t *b = &blk[i];
        const bool dst_cpu = b->new_residency == 0;
        const bool have_res = dst_cpu || (b->flags & REF_PF_HAS_RESIDENT_MASK);
        const bool have_thr = thrashing && (b->flags & REF_PF_HAS_THRASHING);
        or_mask_t res, f, thr, work, out;
        unsigned max_first = 0, max_outer = b->num_pages, pending;
        tree_t t;

        or_zero(&out);
        out_residency[i] = ID_INVALID;
        if (!g_enable) { or_store(out_mask + 8 * i, &out); continue; }

        or_load(&f, faulted + 8 * i);
        if (have_res) or_load(&res, resident + 8 * i);
        if (have_thr) or_load(&thr, thrashing + 8 * i);

        if (info[i].last_migration_proc_id != b->new_residency) {           /* :340-343 */
            info[i].last_migration_proc_id = b->new_residency;
            info[i].fault_migrations_to_last_proc = 0;
        }
        if ((b->flags & REF_PF_BLOCK_RESIDENT_EMPTY) && b->new_residency == b->preferred_location) {
            or_fill_range(&out, max_first, max_outer);                       /* first touch :363-366 */
        }
        else {
            /* big page size: the GPU's if it has a registered VA space, else 64K (:258-270) */
            unsigned big_pages = (!dst_cpu && (b->flags & REF_PF_GPU_VA_SPACE_REGISTERED)) ? b->big_page_size / 4096 : 16;
            tree_init(&t, max_first, max_outer, have_res ? &res : NULL, &f);
            tree_update(&t, b->start, big_pages, max_first, max_outer, &f, have_thr ? &thr : NULL);
            if (have_thr) or_andnot(&work, &f, &thr); else work = f;         /* :378-381 */
            prefetch_mask(&t, &work, b->faulted_first, b->faulted_outer, max_first, max_outer, &out);
        }
        or_andnot(&out, &out, &f);                                           /* :392 */
        if (dst_cpu) {                                                       /* :399-404 */
            or_mask_t mapped, hide;
            if (cpu_read_mapped) or_load(&mapped, cpu_read_mapped + 8 * i); else or_zero(&mapped);
            or_and(&hide, &res, &mapped);
            or_andnot(&out, &out, &hide);
        }
        if (have_thr) or_andnot(&out, &out, &thr);                           /* :407-408 */
        info[i].fault_migrations_to_last_proc =
            (uint16_t)(info[i].fault_migrations_to_last_proc + or_weight_range(&f, b->faulted_first, b->faulted_outer));
        pending = or_weight(&out);

        /* :477-500 -- the mask is returned even when no residency is hinted */
        if (info[i].fault_migrations_to_last_proc >= g_min_faults && pending > 0)
            out_residency[i] = info[i].last_migration_proc_id;
        or_store(out_mask + 8 * i, &out);
    }
}
