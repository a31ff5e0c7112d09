/*
 * oracle_common.h -- TEST INFRASTRUCTURE ONLY.
 * Small helpers for the CPU restatement oracle (oracle/liboracle.so).  The oracle is the checker
 * for the CUDA path: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load
 * it.  Nothing under open-gpu-kernel-modules_b200/ links, imports or calls it.
 *
 * 512-bit page masks are 8 little-endian u64 words, bit p of the mask = page p of the 2 MB
 * va_block (reference: uvm_page_mask_t, kernel-open/nvidia-uvm/uvm_va_block_types.h:77-80).
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

/* number of set bits in [first, outer) */
static inline unsigned or_weight_range(const or_mask_t *m, unsigned first, unsigned outer)
{
    unsigned p, n = 0;
    for (p = first; p < outer; p++)
        n += or_test(m, p);
    return n;
}
static inline unsigned or_weight(const or_mask_t *m) { return or_weight_range(m, 0, OR_PAGES); }
static inline bool or_any_range(const or_mask_t *m, unsigned first, unsigned outer) { return or_weight_range(m, first, outer) != 0; }
static inline void or_fill_range(or_mask_t *m, unsigned first, unsigned outer)
{
    unsigned p;
    for (p = first; p < outer; p++)
        or_set(m, p);
}
static inline void or_clear_range(or_mask_t *m, unsigned first, unsigned outer)
{
    unsigned p;
    for (p = first; p < outer && p < OR_PAGES; p++)
        m->w[p >> 6] &= ~(1ULL << (p & 63));
}
static inline void or_or(or_mask_t *d, const or_mask_t *a, const or_mask_t *b) { int i; for (i = 0; i < OR_WORDS; i++) d->w[i] = a->w[i] | b->w[i]; }
static inline void or_and(or_mask_t *d, const or_mask_t *a, const or_mask_t *b) { int i; for (i = 0; i < OR_WORDS; i++) d->w[i] = a->w[i] & b->w[i]; }
static inline void or_andnot(or_mask_t *d, const or_mask_t *a, const or_mask_t *b) { int i; for (i = 0; i < OR_WORDS; i++) d->w[i] = a->w[i] & ~b->w[i]; }
static inline bool or_empty(const or_mask_t *m) { int i; for (i = 0; i < OR_WORDS; i++) if (m->w[i]) return false; return true; }
/* logical shifts over the 512-bit value, bits shifted out are lost */
static inline void or_shr(or_mask_t *d, const or_mask_t *s, unsigned k)
{
    or_mask_t t; unsigned p;
    or_zero(&t);
    for (p = 0; p + k < OR_PAGES; p++)
        if (or_test(s, p + k)) or_set(&t, p);
    *d = t;
}
static inline void or_shl(or_mask_t *d, const or_mask_t *s, unsigned k)
{
    or_mask_t t; unsigned p;
    or_zero(&t);
    for (p = 0; p + k < OR_PAGES; p++)
        if (or_test(s, p)) or_set(&t, p + k);
    *d = t;
}

#endif
