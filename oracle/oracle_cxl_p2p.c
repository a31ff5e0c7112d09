/*
 * oracle_cxl_p2p.c -- TEST INFRASTRUCTURE ONLY: CPU restatement of the argument checks and page-array arithmetic
 * of the fork's CXL P2P DMA surface (src/nvidia/src/kernel/mem_mgr/p2p_cxl.c; page pinning in
 * kernel-open/nvidia/nv-p2p.c:1140-1260):
 *   RmP2PRegisterCxlBuffer   :225-352  (checks :241-267, huge-page rule _cxlP2PCanUseHugePages :149-160)
 *   RmP2PGetCxlPages         :423-481
 *   RmP2PCxlDmaRequest       :516-677  (checks :535-565, 4 GB clamp :617-622)
 * Status codes are the reference's (kernel-open/common/inc/nvstatuscodes.h).
 *
 * Pinning: by the reference's own file compiled UNMODIFIED (oracle/_ref/libref_cxl_p2p.so, ref_cxl_p2p_glue.c: RM
 * objects replaced by stand-ins, the kernel's page pinning by its page arithmetic): tests/test_oracle_cxl_p2p.py
 * compares every status code, the page geometry of a registration, the window of RmP2PGetCxlPages and the bytes of a
 * DMA request (4 GB clamp) on directed and random arguments.  The byte pattern and the register -> DMA -> verify ->
 * unregister flow of tests/cxl_p2p_test.c:467-513, 532-632 are replayed on the GPU by tests/test_gpu_cxl_p2p.py.
 */
#include <stdint.h>

enum { OK = 0, ERR_INVALID_ARGUMENT = 0x1F, ERR_INVALID_STATE = 0x40, ERR_INSUFFICIENT_RESOURCES = 0x1A };
#define PAGE_4K 4096ull
#define PAGE_2M (2ull << 20)

/* RmP2PRegisterCxlBuffer: status, and the page geometry of the buffer when it is accepted */
uint32_t oracle_cxl_register_check(uint64_t base, uint64_t size, uint32_t version, int has_out_handle, uint32_t n_registered,
                                   uint32_t *page_size, uint64_t *page_count, uint64_t *first_page_addr)
{
    int huge;
    if (size == 0 || !has_out_handle || version < 1 || version > 3) return ERR_INVALID_ARGUMENT;   /* :241-246 */
    if (size > (1ull << 40)) return ERR_INVALID_ARGUMENT;                                         /* :248-253 */
    if (base + size < base) return ERR_INVALID_ARGUMENT;                                          /* :255-260 */
    if (n_registered >= 256) return ERR_INSUFFICIENT_RESOURCES;                                   /* :262-267 */
    huge = (base & (PAGE_2M - 1)) == 0 && (size & (PAGE_2M - 1)) == 0 && size >= PAGE_2M;         /* :149-160 */
    if (huge) {
        *page_size = (uint32_t)PAGE_2M;
        *page_count = size / PAGE_2M;
        *first_page_addr = base;
    }
    else {
        const uint64_t start = base & ~(PAGE_4K - 1);                                             /* nv-p2p.c:1166 */
        *page_size = (uint32_t)PAGE_4K;
        *page_count = ((base + size - 1) >> 12) - (start >> 12) + 1;                              /* nv-p2p.c:1167 */
        *first_page_addr = start;
    }
    return OK;
}

/* RmP2PGetCxlPages: which pages of the array [offset, offset + size) names */
uint32_t oracle_cxl_get_pages(uint64_t buf_size, uint32_t buf_page_size, uint32_t buf_page_count, int registered, uint64_t offset,
                              uint64_t size, uint32_t *start_page, uint32_t *num_pages)
{
    uint32_t start, n;
    if (!registered) return ERR_INVALID_STATE;                                                    /* :445-448 */
    if (offset + size > buf_size) return ERR_INVALID_ARGUMENT;                                    /* :450-453 */
    start = (uint32_t)(offset / buf_page_size);                                                   /* :455-456 */
    n = (uint32_t)((offset + size + buf_page_size - 1) / buf_page_size) - start;
    if (start + n > buf_page_count) return ERR_INVALID_ARGUMENT;                                  /* :458-461 */
    *start_page = start;
    *num_pages = n;
    return OK;
}

/* RmP2PCxlDmaRequest: status and the number of bytes one request moves */
uint32_t oracle_cxl_dma_check(int has_gpu, int has_handle, int registered, uint64_t buf_size, uint64_t cxl_offset, uint64_t size,
                              uint64_t *bytes)
{
    if (!has_gpu || !has_handle || size == 0) return ERR_INVALID_ARGUMENT;                        /* :535-551 */
    if (!registered) return ERR_INVALID_STATE;                                                    /* :553-557 */
    if (cxl_offset > buf_size || size > buf_size - cxl_offset) return ERR_INVALID_ARGUMENT;       /* :559-565 */
    *bytes = size > 0xFFFFFFFFull ? 0xFFFFF000ull : size;                                         /* :617-622 */
    return OK;
}
