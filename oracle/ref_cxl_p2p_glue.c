/*
 * ref_cxl_p2p_glue.c -- TEST INFRASTRUCTURE ONLY (oracle/_ref build -> libref_cxl_p2p.so).
 *
 * Compiles the fork's CXL P2P DMA file src/nvidia/src/kernel/mem_mgr/p2p_cxl.c UNMODIFIED (the whole file is
 * #included below; its nine RM headers are satisfied by empty files the build recipe creates in a temp dir) against
 * stand-ins for the RM objects it touches:
 *   OBJGPU / MemoryManager / KernelBus   plain structs; the GPU is never "lost"
 *   MEMORY_DESCRIPTOR, memdesc*          a record of what was described; memmgrMemCopy records the transfer size
 *   portMem*                             malloc / memset / free
 *   nv_pin_cxl_buffer*, nv_get_cxl_*     the kernel glue of kernel-open/nvidia/nv-p2p.c:1140-1260 reduced to its
 *                                        page arithmetic (4 KB pages from the page-aligned base :1166-1167, or one
 *                                        entry per 2 MB page), addresses instead of pinned struct pages
 *   nv_enumerate_cxl_devices             fails, so RmP2PGetCxlSystemInfo reports its defaults (:117-123)
 * The wrappers at the end have the contract of the three functions of oracle/oracle_cxl_p2p.c, so that the
 * restatement (and through it the library's uvmtier_p2p_* entry points) is compared with the reference's own code:
 * status codes of every argument check, page size / count / first page of a registration, the page window of
 * RmP2PGetCxlPages, the bytes one RmP2PCxlDmaRequest moves (4 GB clamp included).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "nvtypes.h"
#include "nvstatus.h"

typedef NvU64 RmPhysAddr;
typedef struct MEMORY_DESCRIPTOR { NvU64 size, offset; NvU32 pages, page_size, addr_space; } MEMORY_DESCRIPTOR;
typedef struct MemoryManager { int unused; } MemoryManager;
typedef struct KernelBus { int unused; } KernelBus;
typedef struct OBJGPU OBJGPU;
struct OBJGPU {
    NvBool bIsSOC;
    NvBool (*getProperty)(OBJGPU *, int);
    MemoryManager mm;
    KernelBus bus;
};
typedef struct { MEMORY_DESCRIPTOR *pMemDesc; NvU64 offset; } TRANSFER_SURFACE;
enum { ADDR_SYSMEM = 1, ADDR_FBMEM = 2, NV_MEMORY_UNCACHED = 0, MEMDESC_FLAGS_NONE = 0, MEMDESC_FLAGS_SKIP_IOMMU_MAPPING = 1,
       TRANSFER_FLAGS_PREFER_CE = 1, PDB_PROP_GPU_IS_LOST = 1, AT_GPU = 0 };
#define GPU_GET_MEMORY_MANAGER(g) (&(g)->mm)
#define GPU_GET_KERNEL_BUS(g) (&(g)->bus)
#define NV_PRINTF(...) ((void)0)
#define portMemAllocNonPaged(n) malloc(n)
#define portMemSet(p, v, n) memset(p, v, n)
#define portMemFree(p) free(p)

static NV_STATUS memdescCreate(MEMORY_DESCRIPTOR **pp, OBJGPU *g, NvU64 size, NvU64 align, NvBool contig, int space, int attr, int flags)
{
    (void)g; (void)align; (void)contig; (void)attr; (void)flags;
    *pp = calloc(1, sizeof(**pp));
    if (!*pp) return NV_ERR_NO_MEMORY;
    (*pp)->size = size; (*pp)->addr_space = (NvU32)space;
    return NV_OK;
}
static void memdescDestroy(MEMORY_DESCRIPTOR *p) { free(p); }
static void memdescFillPages(MEMORY_DESCRIPTOR *p, NvU32 first, RmPhysAddr *pte, NvU32 n, NvU64 page_size)
{ (void)first; (void)pte; p->pages = n; p->page_size = (NvU32)page_size; }
static void memdescSetPageSize(MEMORY_DESCRIPTOR *p, int at, NvU64 page_size) { (void)at; p->page_size = (NvU32)page_size; }
static void memdescDescribe(MEMORY_DESCRIPTOR *p, int space, NvU64 offset, NvU64 size) { p->addr_space = (NvU32)space; p->offset = offset; p->size = size; }
static NvU64 g_last_copy_bytes;
static NV_STATUS memmgrMemCopy(MemoryManager *mm, TRANSFER_SURFACE *dst, TRANSFER_SURFACE *src, NvU32 size, int flags)
{ (void)mm; (void)dst; (void)src; (void)flags; g_last_copy_bytes = size; return NV_OK; }

/* ---- the kernel glue's page arithmetic (nv-p2p.c:1140-1260) ---- */
typedef struct { NvU64 *pages; NvU32 count, page_size; } pinned_t;
#define GLUE_MAX_ARRAY (1u << 16)      /* entries actually materialised: the callers here read only the first ones */
static NV_STATUS pin_common(NvU64 base, NvU64 size, NvU32 page_size, void **out)
{
    pinned_t *p = calloc(1, sizeof(*p));
    const NvU64 start = base & ~((NvU64)page_size - 1);
    NvU64 n = ((base + size - 1) / page_size) - (start / page_size) + 1, i;
    if (!p) return NV_ERR_NO_MEMORY;
    p->count = (NvU32)n; p->page_size = page_size;
    p->pages = malloc(sizeof(NvU64) * (n < GLUE_MAX_ARRAY ? n : GLUE_MAX_ARRAY));
    for (i = 0; i < n && i < GLUE_MAX_ARRAY; i++) p->pages[i] = start + i * page_size;
    *out = p;
    return NV_OK;
}
NV_STATUS nv_pin_cxl_buffer(NvU64 base, NvU64 size, void **out) { return pin_common(base, size, 4096, out); }
NV_STATUS nv_pin_cxl_buffer_hugepages(NvU64 base, NvU64 size, NvU32 page_size, void **out) { return pin_common(base, size, page_size, out); }
NV_STATUS nv_unpin_cxl_buffer(void *h) { pinned_t *p = h; free(p->pages); free(p); return NV_OK; }
NV_STATUS nv_get_cxl_buffer_pages(void *h, NvU64 **pages, NvU32 *count) { pinned_t *p = h; *pages = p->pages; *count = p->count; return NV_OK; }
NV_STATUS nv_get_cxl_buffer_hugepages(void *h, NvU64 **pages, NvU32 *count, NvU32 *page_size)
{ pinned_t *p = h; *pages = p->pages; *count = p->count; *page_size = p->page_size; return NV_OK; }
struct CXL_SYSTEM_INFO;
NV_STATUS nv_enumerate_cxl_devices(struct CXL_SYSTEM_INFO *info) { (void)info; return NV_ERR_NOT_SUPPORTED; }

#include "p2p_cxl.c"            /* the reference file, unmodified */

/* ======================= same contracts as oracle_cxl_p2p.c ======================= */
uint32_t ref_cxl_register_check(uint64_t base, uint64_t size, uint32_t version, int has_out_handle, uint32_t n_registered,
                                uint32_t *page_size, uint64_t *page_count, uint64_t *first_page_addr)
{
    void *h = NULL;
    NV_STATUS st;
    const NvU32 saved = g_cxlRegisteredBufferCount;
    g_cxlRegisteredBufferCount = n_registered;
    st = RmP2PRegisterCxlBuffer(NULL, base, size, version, has_out_handle ? &h : NULL);
    if (st == NV_OK) {
        CXL_P2P_BUFFER_HANDLE *b = h;
        *page_size = b->pageSize; *page_count = b->pageCount; *first_page_addr = b->pPageArray[0];
        RmP2PUnregisterCxlBuffer(h);
    }
    g_cxlRegisteredBufferCount = saved;
    return st;
}

uint32_t ref_cxl_get_pages(uint64_t buf_size, uint32_t buf_page_size, uint32_t buf_page_count, int registered, uint64_t offset,
                           uint64_t size, uint32_t *start_page, uint32_t *num_pages)
{
    CXL_P2P_BUFFER_HANDLE b;
    NvU64 *arr = malloc(sizeof(NvU64) * (buf_page_count ? buf_page_count : 1)), *out = malloc(sizeof(NvU64) * (buf_page_count ? buf_page_count : 1));
    NvU32 n = 0, ps = 0, i;
    NV_STATUS st;
    memset(&b, 0, sizeof(b));
    for (i = 0; i < buf_page_count; i++) arr[i] = i;            /* the entry is its own index: out[0] names the window's start */
    b.size = buf_size; b.pageSize = buf_page_size; b.pageCount = buf_page_count; b.pPageArray = arr; b.bRegistered = registered != 0;
    st = RmP2PGetCxlPages(&b, offset, size, out, &n, &ps);
    if (st == NV_OK) { *start_page = n ? (uint32_t)out[0] : (uint32_t)(offset / buf_page_size); *num_pages = n; }
    free(arr); free(out);
    return st;
}

static NvBool never(OBJGPU *g, int p) { (void)g; (void)p; return NV_FALSE; }
uint32_t ref_cxl_dma_check(int has_gpu, int has_handle, int registered, uint64_t buf_size, uint64_t cxl_offset, uint64_t size,
                           uint64_t *bytes)
{
    OBJGPU gpu;
    CXL_P2P_BUFFER_HANDLE b;
    NvU64 one_page = 0;
    NV_STATUS st;
    memset(&gpu, 0, sizeof(gpu));
    gpu.getProperty = never;
    memset(&b, 0, sizeof(b));
    b.size = buf_size; b.pageSize = 2u << 20; b.pageCount = 1; b.pPageArray = &one_page; b.bRegistered = registered != 0;
    g_last_copy_bytes = 0;
    st = RmP2PCxlDmaRequest(has_gpu ? &gpu : NULL, has_handle ? &b : NULL, 0, cxl_offset, size, 0);
    if (b.pCxlMemDesc) memdescDestroy(b.pCxlMemDesc);
    if (st == NV_OK) *bytes = g_last_copy_bytes;
    return st;
}

/* RmP2PGetCxlSystemInfo with enumeration failing: the defaults of :117-123 */
void ref_cxl_system_info_defaults(uint32_t *devices, uint32_t *mem_devices, uint32_t *link_up, uint32_t *version)
{
    NvU32 d = 9, m = 9, v = 9; NvBool l = NV_TRUE;
    RmP2PGetCxlSystemInfo(&d, &m, &l, &v);
    *devices = d; *mem_devices = m; *link_up = l; *version = v;
}
