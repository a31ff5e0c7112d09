/*
 * oracle_common.h -- TEST INFRASTRUCTURE ONLY.
 * Small helpers for the CPU restatement oracle (oracle/liboracle.so).  The oracle is the checker
 * for the CUDA path: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load
 * it.  Nothing under open-gpu-kernel-modules_b200/ links, imports or calls it.
 *
 * 512-bit page masks are 8 little-endian u64 words, bit p of the mask = page p of the 2 MB
 * va_block (reference: uvm_page_mask_t, kernel-open/nvidia-uvm/uvm_va_block_types.h:77-80).
 * The helpers are word-wise (like the Linux bitmap library the reference uses) so that the oracle is
 * also a fair single-thread CPU baseline, not only a checker.
 */
#ifndef ORACLE_COMMON_H
#define ORACLE_COMMON_H

#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#define OR_PAGES 512
#define OR_WORDS 8

typedef struct { uint64_t w[OR_WORDS]; } or_mask_t;

static inline bool or_test(const or_mask_t *m, unsigned p) { return (m->w[p >> 6] >> (p & 63)) & 1; }
static inline void or_set(or_mask_t *m, unsigned p) { m->w[p >> 6] |= 1ULL << (p & 63); }
static inline void or_zero(or_mask_t *m) { memset(m, 0, sizeof(*m)); }
static inline void or_load(or_mask_t *m, const uint64_t *src) { memcpy(m->w, src, 64); }
static inline void or_store(uint64_t *dst, const or_mask_t *m) { memcpy(dst, m->w, 64); }

/* word i of the mask with bits [first, outer) set */
static inline uint64_t or_range_word(unsigned i, unsigned first, unsigned outer)
{
    int lo = (int)first - 64 * (int)i, hi = (int)outer - 64 * (int)i;
    uint64_t a, b;
    if (lo < 0) lo = 0;
    if (hi > 64) hi = 64;
    if (hi <= lo) return 0;
    a = hi == 64 ? ~0ULL : ((1ULL << hi) - 1);
    b = lo == 0 ? 0 : ((1ULL << lo) - 1);
    return a & ~b;
}
/* number of set bits in [first, outer) */
static inline unsigned or_weight_range(const or_mask_t *m, unsigned first, unsigned outer)
{
    unsigned i, n = 0;
    if (outer > OR_PAGES) outer = OR_PAGES;
    for (i = first >> 6; i < OR_WORDS && 64 * i < outer; i++)
        n += (unsigned)__builtin_popcountll(m->w[i] & or_range_word(i, first, outer));
    return n;
}
static inline unsigned or_weight(const or_mask_t *m)
{
    unsigned i, n = 0;
    for (i = 0; i < OR_WORDS; i++) n += (unsigned)__builtin_popcountll(m->w[i]);
    return n;
}
static inline bool or_any_range(const or_mask_t *m, unsigned first, unsigned outer)
{
    unsigned i;
    if (outer > OR_PAGES) outer = OR_PAGES;
    for (i = first >> 6; i < OR_WORDS && 64 * i < outer; i++)
        if (m->w[i] & or_range_word(i, first, outer)) return true;
    return false;
}
static inline void or_fill_range(or_mask_t *m, unsigned first, unsigned outer)
{
    unsigned i;
    if (outer > OR_PAGES) outer = OR_PAGES;
    for (i = first >> 6; i < OR_WORDS && 64 * i < outer; i++)
        m->w[i] |= or_range_word(i, first, outer);
}
static inline void or_clear_range(or_mask_t *m, unsigned first, unsigned outer)
{
    unsigned i;
    if (outer > OR_PAGES) outer = OR_PAGES;
    for (i = first >> 6; i < OR_WORDS && 64 * i < outer; i++)
        m->w[i] &= ~or_range_word(i, first, outer);
}
static inline void or_or(or_mask_t *d, const or_mask_t *a, const or_mask_t *b) { int i; for (i = 0; i < OR_WORDS; i++) d->w[i] = a->w[i] | b->w[i]; }
static inline void or_and(or_mask_t *d, const or_mask_t *a, const or_mask_t *b) { int i; for (i = 0; i < OR_WORDS; i++) d->w[i] = a->w[i] & b->w[i]; }
static inline void or_andnot(or_mask_t *d, const or_mask_t *a, const or_mask_t *b) { int i; for (i = 0; i < OR_WORDS; i++) d->w[i] = a->w[i] & ~b->w[i]; }
static inline bool or_empty(const or_mask_t *m) { int i; for (i = 0; i < OR_WORDS; i++) if (m->w[i]) return false; return true; }
/* logical shifts over the 512-bit value, bits shifted out are lost */
static inline void or_shr(or_mask_t *d, const or_mask_t *s, unsigned k)
{
    or_mask_t t;
    unsigned q = k >> 6, r = k & 63, i;
    for (i = 0; i < OR_WORDS; i++) {
        uint64_t lo = i + q < OR_WORDS ? s->w[i + q] : 0, hi = i + q + 1 < OR_WORDS ? s->w[i + q + 1] : 0;
        t.w[i] = r ? (lo >> r) | (hi << (64 - r)) : lo;
    }
    *d = t;
}
static inline void or_shl(or_mask_t *d, const or_mask_t *s, unsigned k)
{
    or_mask_t t;
    unsigned q = k >> 6, r = k & 63, i;
    for (i = 0; i < OR_WORDS; i++) {
        uint64_t hi = i >= q ? s->w[i - q] : 0, lo = i >= q + 1 ? s->w[i - q - 1] : 0;
        t.w[i] = r ? (hi << r) | (lo >> (64 - r)) : hi;
    }
    *d = t;
}

#endif
