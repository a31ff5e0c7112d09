/*
 * ref_copy_glue.c -- TEST INFRASTRUCTURE ONLY (oracle/_ref build -> libref_copy.so).
 *
 * Compiles the reference's CPU page-copy loop VERBATIM -- the line-range extract kernel-open/nvidia-uvm/uvm_va_block.c
 * :3946-3991 (block_copy_pages: for CPU-to-CPU moves, i.e. between two NUMA nodes such as host DRAM and a CXL node,
 * one kmap + memcpy(PAGE_SIZE) + kunmap per page, :3972-3984) -- against stand-ins for what it touches:
 *   struct page        one 4 KB page of the arena, so that page arithmetic is address arithmetic
 *   kmap / kunmap      identity / nothing ("kmap -> identity")
 *   uvm_cpu_chunk_*    a va_block is backed by one 2 MB chunk per NUMA node, contiguous in that node's arena
 *   dirty tracking     every page is dirty (what a migration of written data sees); the mark is a bit set
 *   block_copy_should_use_push   false: CPU source and CPU destination, uvm_block_cpu_to_cpu_copy_with_ce = 0
 *                                (the default, uvm_va_block.c:69-71)
 * bench.py times it between the DRAM and the CXL tier next to the GPU's copy of the same pages.
 */
#include "shim/kshim.h"
#include "ref_api.h"

#include <string.h>

#define PAGES_PER_BLOCK 512
struct page { unsigned char bytes[4096]; };
typedef uint16_t uvm_page_index_t;
typedef struct { uvm_page_index_t first, outer; } uvm_va_block_region_t;
typedef struct { int unused; } uvm_tracker_t;
typedef struct { int unused; } uvm_push_t;
typedef struct { struct page *pages; } uvm_cpu_chunk_t;
typedef struct {
    uvm_tracker_t tracker;
    uvm_cpu_chunk_t chunk[2];                   /* one 2 MB chunk per NUMA node */
    uint64_t dirty[2][PAGES_PER_BLOCK / 64];
} uvm_va_block_t;
typedef struct { unsigned val; } uvm_processor_id_t;
typedef struct { uvm_processor_id_t id; int nid; } block_copy_side_t;
typedef struct { block_copy_side_t src, dst; } block_copy_state_t;

static NV_STATUS uvm_tracker_wait(uvm_tracker_t *t) { (void)t; return NV_OK; }
static void *kmap(struct page *p) { return p; }
static void kunmap(void *addr) { (void)addr; }
static uvm_cpu_chunk_t *uvm_cpu_chunk_get_chunk_for_page(uvm_va_block_t *b, int nid, uvm_page_index_t page) { (void)page; return &b->chunk[nid]; }
static uvm_va_block_region_t uvm_cpu_chunk_block_region(uvm_va_block_t *b, uvm_cpu_chunk_t *c, uvm_page_index_t page)
{ uvm_va_block_region_t r = { 0, PAGES_PER_BLOCK }; (void)b; (void)c; (void)page; return r; }
static struct page *uvm_cpu_chunk_get_cpu_page(uvm_va_block_t *b, uvm_cpu_chunk_t *c, uvm_page_index_t page) { (void)b; return c->pages + page; }
static size_t uvm_cpu_chunk_get_size(uvm_cpu_chunk_t *c) { (void)c; return (size_t)PAGES_PER_BLOCK * 4096; }
static size_t uvm_va_block_region_size(uvm_va_block_region_t r) { return (size_t)(r.outer - r.first) * 4096; }
static bool block_cpu_page_is_dirty(uvm_va_block_t *b, uvm_page_index_t page, int nid) { (void)b; (void)page; (void)nid; return true; }
static void block_mark_cpu_page_dirty(uvm_va_block_t *b, uvm_page_index_t page, int nid) { b->dirty[nid][page / 64] |= 1ull << (page % 64); }
static bool block_copy_should_use_push(uvm_va_block_t *b, block_copy_state_t *s) { (void)b; (void)s; return false; }
static void block_copy_push(uvm_va_block_t *b, block_copy_state_t *s, uvm_va_block_region_t r, uvm_push_t *p) { (void)b; (void)s; (void)r; (void)p; }
#define for_each_va_block_page_in_region(page_index, region) \
    for ((page_index) = (region).first; (page_index) != (region).outer; (page_index)++)

#include "x_copy_pages.h"       /* uvm_va_block.c:3946-3991, verbatim */

/*
 * Copies pages [first, outer) of va_blocks [block_first, block_outer) from the arena at src_base (NUMA node 0 in the
 * reference's terms) to the one at dst_base (node 1); block b of an arena is its bytes [b * 2 MB, (b + 1) * 2 MB).
 * Returns the number of pages the reference's loop marked dirty at the destination (= pages copied).
 */
uint64_t ref_copy_pages_cpu(void *dst_base, const void *src_base, uint32_t block_first, uint32_t block_outer, uint32_t first,
                            uint32_t outer)
{
    uint64_t marked = 0;
    uint32_t b;
    for (b = block_first; b < block_outer; b++) {
        uvm_va_block_t blk;
        block_copy_state_t cs;
        uvm_va_block_region_t region = { (uvm_page_index_t)first, (uvm_page_index_t)outer };
        unsigned w;
        memset(&blk, 0, sizeof(blk));
        blk.chunk[0].pages = (struct page *)((unsigned char *)(uintptr_t)src_base + ((size_t)b << 21));
        blk.chunk[1].pages = (struct page *)((unsigned char *)dst_base + ((size_t)b << 21));
        cs.src.id.val = 0; cs.src.nid = 0;
        cs.dst.id.val = 0; cs.dst.nid = 1;
        if (block_copy_pages(&blk, &cs, region, NULL) != NV_OK)
            return ~0ull;
        for (w = 0; w < PAGES_PER_BLOCK / 64; w++) marked += (uint64_t)__builtin_popcountll(blk.dirty[1][w]);
    }
    return marked;
}
