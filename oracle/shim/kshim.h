/*
 * kshim.h -- user-space stand-ins for the handful of Linux kernel primitives the
 * reference's hot-path C uses.  TEST INFRASTRUCTURE ONLY: this header is part of
 * the oracle build (oracle/_ref) that compiles the reference's own sources from
 * /root/reference against thin shims.  Nothing here is product code and nothing
 * under open-gpu-kernel-modules_b200/ includes it.
 *
 * The semantics restated here are those of the Linux bitmap/bitops/list API as
 * used through kernel-open/nvidia-uvm/uvm_va_block.h:1706-1881 and
 * uvm_perf_prefetch.c:237,284 (exact bit math).
 */
#ifndef ORACLE_KSHIM_H
#define ORACLE_KSHIM_H

#include <assert.h>
#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "nvtypes.h"
#include "nvstatus.h"

typedef int8_t s8;
typedef uint8_t u8;
typedef uint16_t u16;
typedef uint32_t u32;
typedef uint64_t u64;

#define PAGE_SHIFT 12
#define PAGE_SIZE (1UL << PAGE_SHIFT)
#define PAGE_MASK (~(PAGE_SIZE - 1))
#define PAGE_ALIGNED(x) ((((NvU64)(x)) & (PAGE_SIZE - 1)) == 0)
#define NUMA_NO_NODE (-1)

#define BITS_PER_LONG 64
#define BITS_TO_LONGS(n) (((n) + BITS_PER_LONG - 1) / BITS_PER_LONG)
#define DECLARE_BITMAP(name, bits) unsigned long name[BITS_TO_LONGS(bits)]
#define BITMAP_LAST_WORD_MASK(nbits) (~0UL >> (-(nbits) & (BITS_PER_LONG - 1)))
#define BITMAP_FIRST_WORD_MASK(start) (~0UL << ((start) & (BITS_PER_LONG - 1)))

#define BUILD_BUG_ON(c) _Static_assert(!(c), "BUILD_BUG_ON")
#define likely(x) (x)
#define unlikely(x) (x)
#define S_IRUGO 0444
#define module_param(name, type, perm) struct kshim_unused_##name
#define MODULE_PARM_DESC(name, desc) struct kshim_unused_desc_##name

#define min(a, b) ({ typeof(a) _mina = (a); typeof(b) _minb = (b); _mina < _minb ? _mina : _minb; })
#define max(a, b) ({ typeof(a) _maxa = (a); typeof(b) _maxb = (b); _maxa > _maxb ? _maxa : _maxb; })
#define min_t(t, a, b) ({ t _mina = (a); t _minb = (b); _mina < _minb ? _mina : _minb; })
#define max_t(t, a, b) ({ t _maxa = (a); t _maxb = (b); _maxa > _maxb ? _maxa : _maxb; })

#define UVM_ASSERT(e) assert(e)
#define UVM_ASSERT_MSG(e, ...) assert(e)
#define UVM_ASSERT_RELEASE(e) assert(e)
#define UVM_INFO_PRINT(...) ((void)0)
#define UVM_ERR_PRINT(...) ((void)0)
#define UVM_DBG_PRINT(...) ((void)0)

static inline bool is_power_of_2(unsigned long n) { return n != 0 && (n & (n - 1)) == 0; }

#define UVM_ALIGN_DOWN(x, a) ({ typeof(x) _a = a; UVM_ASSERT(is_power_of_2(_a)); (x) & ~(_a - 1); })
#define UVM_ALIGN_UP(x, a) ({ typeof(x) _a = a; UVM_ASSERT(is_power_of_2(_a)); ((x) + _a - 1) & ~(_a - 1); })
#define UVM_PAGE_ALIGN_DOWN(v) UVM_ALIGN_DOWN(v, PAGE_SIZE)
#define UVM_CMP_DEFAULT(a, b) ({ typeof(a) _a = a; typeof(b) _b = b; int _r; if (_a < _b) _r = -1; else if (_b < _a) _r = 1; else _r = 0; _r; })

#define UVM_PAGE_SIZE_4K (4ULL * 1024)
#define UVM_PAGE_SIZE_64K (64ULL * 1024)
#define UVM_PAGE_SIZE_128K (128ULL * 1024)
#define UVM_PAGE_SIZE_2M (2ULL * 1024 * 1024)

static inline int ilog2(unsigned long v) { return 63 - __builtin_clzl(v); }
static inline unsigned long roundup_pow_of_two(unsigned long v) { return v <= 1 ? 1UL : 1UL << (64 - __builtin_clzl(v - 1)); }
/* must be an integer constant expression: the reference sizes bit-fields with it */
#define order_base_2(n) ((n) <= 1 ? 0 : (n) <= 2 ? 1 : (n) <= 4 ? 2 : (n) <= 8 ? 3 : (n) <= 16 ? 4 : (n) <= 32 ? 5 : \
                         (n) <= 64 ? 6 : (n) <= 128 ? 7 : (n) <= 256 ? 8 : (n) <= 512 ? 9 : (n) <= 1024 ? 10 : \
                         (n) <= 2048 ? 11 : (n) <= 4096 ? 12 : (n) <= 8192 ? 13 : (n) <= 16384 ? 14 : (n) <= 32768 ? 15 : 16)
static inline unsigned long __fls(unsigned long w) { return 63 - __builtin_clzl(w); }
static inline unsigned long __ffs(unsigned long w) { return __builtin_ctzl(w); }
static inline NvU64 uvm_div_pow2_64(NvU64 n, NvU64 d) { return n >> ilog2(d); }

/* ---- single-bit ops ---- */
static inline bool test_bit(unsigned nr, const unsigned long *m) { return (m[nr / 64] >> (nr % 64)) & 1; }
static inline void __set_bit(unsigned nr, unsigned long *m) { m[nr / 64] |= 1UL << (nr % 64); }
static inline void __clear_bit(unsigned nr, unsigned long *m) { m[nr / 64] &= ~(1UL << (nr % 64)); }
static inline bool __test_and_set_bit(unsigned nr, unsigned long *m) { bool o = test_bit(nr, m); __set_bit(nr, m); return o; }
static inline bool __test_and_clear_bit(unsigned nr, unsigned long *m) { bool o = test_bit(nr, m); __clear_bit(nr, m); return o; }
#define set_bit __set_bit
#define clear_bit __clear_bit

/* ---- bitmap API: word-wise like Linux lib/bitmap.c, so that timing the reference through this shim is a fair
 * CPU baseline (the semantics are exact bit math either way) ---- */
static inline unsigned long kshim_range_word(unsigned i, long first, long outer)
{
    long lo = first - 64 * (long)i, hi = outer - 64 * (long)i;
    unsigned long a, b;
    if (lo < 0) lo = 0;
    if (hi > 64) hi = 64;
    if (hi <= lo) return 0;
    a = hi == 64 ? ~0UL : ((1UL << hi) - 1);
    b = lo == 0 ? 0 : ((1UL << lo) - 1);
    return a & ~b;
}
static inline void bitmap_zero(unsigned long *d, unsigned n) { memset(d, 0, BITS_TO_LONGS(n) * sizeof(long)); }
static inline void bitmap_fill(unsigned long *d, unsigned n) { unsigned i; for (i = 0; i < BITS_TO_LONGS(n); i++) d[i] = kshim_range_word(i, 0, n); }
static inline void bitmap_copy(unsigned long *d, const unsigned long *s, unsigned n) { memcpy(d, s, BITS_TO_LONGS(n) * sizeof(long)); }
static inline unsigned bitmap_weight(const unsigned long *s, unsigned n)
{ unsigned i, w = 0; for (i = 0; i < BITS_TO_LONGS(n); i++) w += (unsigned)__builtin_popcountl(s[i] & kshim_range_word(i, 0, n)); return w; }
static inline bool bitmap_empty(const unsigned long *s, unsigned n) { return bitmap_weight(s, n) == 0; }
static inline bool bitmap_full(const unsigned long *s, unsigned n) { return bitmap_weight(s, n) == n; }
static inline void bitmap_set(unsigned long *m, unsigned start, int len)
{ unsigned i; if (len <= 0) return; for (i = start / 64; 64 * i < start + (unsigned)len; i++) m[i] |= kshim_range_word(i, start, (long)start + len); }
static inline void bitmap_clear(unsigned long *m, unsigned start, int len)
{ unsigned i; if (len <= 0) return; for (i = start / 64; 64 * i < start + (unsigned)len; i++) m[i] &= ~kshim_range_word(i, start, (long)start + len); }
static inline bool bitmap_and(unsigned long *d, const unsigned long *a, const unsigned long *b, unsigned n)
{ unsigned i; unsigned long r = 0; for (i = 0; i < BITS_TO_LONGS(n); i++) r |= (d[i] = a[i] & b[i]); return r != 0; }
static inline bool bitmap_andnot(unsigned long *d, const unsigned long *a, const unsigned long *b, unsigned n)
{ unsigned i; unsigned long r = 0; for (i = 0; i < BITS_TO_LONGS(n); i++) r |= (d[i] = a[i] & ~b[i]); return r != 0; }
static inline void bitmap_or(unsigned long *d, const unsigned long *a, const unsigned long *b, unsigned n)
{ unsigned i; for (i = 0; i < BITS_TO_LONGS(n); i++) d[i] = a[i] | b[i]; }
static inline void bitmap_complement(unsigned long *d, const unsigned long *s, unsigned n)
{ unsigned i; for (i = 0; i < BITS_TO_LONGS(n); i++) d[i] = ~s[i]; if (n % 64) d[n / 64] &= BITMAP_LAST_WORD_MASK(n); }
static inline bool bitmap_equal(const unsigned long *a, const unsigned long *b, unsigned n)
{ unsigned i; for (i = 0; i < n; i++) if (test_bit(i, a) != test_bit(i, b)) return false; return true; }
static inline bool bitmap_subset(const unsigned long *a, const unsigned long *b, unsigned n)
{ unsigned i; for (i = 0; i < n; i++) if (test_bit(i, a) && !test_bit(i, b)) return false; return true; }
static inline bool bitmap_intersects(const unsigned long *a, const unsigned long *b, unsigned n)
{ unsigned i; for (i = 0; i < n; i++) if (test_bit(i, a) && test_bit(i, b)) return true; return false; }
/* dst may alias src, as the callers in uvm_perf_prefetch.c do */
static inline void bitmap_shift_right(unsigned long *d, const unsigned long *s, unsigned shift, unsigned n)
{
    const unsigned W = BITS_TO_LONGS(n), q = shift / 64, r = shift % 64;
    unsigned long t[BITS_TO_LONGS(n)];
    unsigned i;
    for (i = 0; i < W; i++) {
        unsigned long lo = i + q < W ? s[i + q] & kshim_range_word(i + q, 0, n) : 0;
        unsigned long hi = i + q + 1 < W ? s[i + q + 1] & kshim_range_word(i + q + 1, 0, n) : 0;
        t[i] = r ? (lo >> r) | (hi << (64 - r)) : lo;
    }
    memcpy(d, t, sizeof(t));
}
static inline void bitmap_shift_left(unsigned long *d, const unsigned long *s, unsigned shift, unsigned n)
{
    const unsigned W = BITS_TO_LONGS(n), q = shift / 64, r = shift % 64;
    unsigned long t[BITS_TO_LONGS(n)];
    unsigned i;
    for (i = 0; i < W; i++) {
        unsigned long hi = i >= q ? s[i - q] : 0, lo = i >= q + 1 ? s[i - q - 1] : 0;
        t[i] = (r ? (hi << r) | (lo >> (64 - r)) : hi) & kshim_range_word(i, 0, n);
    }
    memcpy(d, t, sizeof(t));
}

static inline unsigned long find_next_bit(const unsigned long *m, unsigned long size, unsigned long off)
{
    unsigned long i;
    for (i = off / 64; 64 * i < size; i++) {
        unsigned long w = m[i] & kshim_range_word((unsigned)i, (long)off, (long)size);
        if (w) return 64 * i + (unsigned long)__builtin_ctzl(w);
    }
    return size;
}
static inline unsigned long find_next_zero_bit(const unsigned long *m, unsigned long size, unsigned long off)
{
    unsigned long i;
    for (i = off / 64; 64 * i < size; i++) {
        unsigned long w = ~m[i] & kshim_range_word((unsigned)i, (long)off, (long)size);
        if (w) return 64 * i + (unsigned long)__builtin_ctzl(w);
    }
    return size;
}
static inline unsigned long find_first_bit(const unsigned long *m, unsigned long size) { return find_next_bit(m, size, 0); }
static inline unsigned long find_first_zero_bit(const unsigned long *m, unsigned long size) { return find_next_zero_bit(m, size, 0); }
static inline unsigned long find_last_bit(const unsigned long *m, unsigned long size)
{
    unsigned long i = BITS_TO_LONGS(size);
    while (i-- > 0) {
        unsigned long w = m[i] & kshim_range_word((unsigned)i, 0, (long)size);
        if (w) return 64 * i + 63 - (unsigned long)__builtin_clzl(w);
    }
    return size;
}

/* ---- list_head ---- */
struct list_head { struct list_head *next, *prev; };
static inline void INIT_LIST_HEAD(struct list_head *l) { l->next = l; l->prev = l; }
static inline void __list_add(struct list_head *n, struct list_head *prev, struct list_head *next) { next->prev = n; n->next = next; n->prev = prev; prev->next = n; }
static inline void list_add(struct list_head *n, struct list_head *head) { __list_add(n, head, head->next); }
static inline void list_add_tail(struct list_head *n, struct list_head *head) { __list_add(n, head->prev, head); }
static inline void list_replace(struct list_head *old, struct list_head *n) { n->next = old->next; n->next->prev = n; n->prev = old->prev; n->prev->next = n; }
static inline int list_empty(const struct list_head *h) { return h->next == h; }

/* ---- misc ---- */
#define smp_mb__after_atomic() __sync_synchronize()
typedef struct { int dummy; } uvm_spin_loop_t;
#define UVM_SPIN_WHILE(cond, spin) while (cond)
#define uvm_assert_rwsem_locked(l) ((void)0)
#define uvm_assert_rwsem_locked_read(l) ((void)0)
#define uvm_assert_mutex_locked(l) ((void)0)
#define uvm_sem_is_locked(l) true
struct file;

#endif /* ORACLE_KSHIM_H */
