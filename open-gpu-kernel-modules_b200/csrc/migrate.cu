// migrate.cu -- K3: memory tiers (HBM, host DRAM, CXL NUMA node, peer HBM) and the page-list copy entry points.
// Reference: the copy leg of uvm_va_block_make_resident (UVM/uvm_va_block.c:3863-3991, 4069-4377), CPU chunks
// allocated per NUMA node (UVM/uvm_pmm_sysmem.c:829-853), peer apertures (UVM/uvm_gpu.h:611-640).
//
// Layout.  Every tier is one contiguous arena; va_block slot s lives at byte offset s * 2 MB in each of them.
// Host tiers are anonymous mappings bound to a NUMA node with mbind(2) (raw syscall, no libnuma) and
// registered with CUDA as portable mapped memory, so the SMs read/write them in place over PCIe.
// Bound: the host link (PCIe Gen5 x16) for DRAM/CXL <-> HBM, NVLink for peer; payload bytes = pages * page_size.
#include "common.cuh"

#include <dirent.h>
#include <errno.h>
#include <stdlib.h>
#include <sys/mman.h>
#include <sys/syscall.h>
#include <unistd.h>

#include "migrate_kernels.cuh"

using namespace uvmtier_k;

struct uvmtier_tier {
    uvmtier_tier_info_t info;
    bool used;
    bool registered;      // cudaHostRegister done by us
    bool mmapped;         // mmap done by us
    bool cuda_alloc;      // cudaMalloc done by us
};

struct uvmtier_tiers {
    uvmtier_tier t[UVMTIER_MAX_TIERS];
    uint64_t *d_lists;    // scratch for uvmtier_copy_runs_device: the page list (all pairs) and the per-pair counters
    uint32_t *d_list_count;
    uint32_t list_cap;
    uint32_t *h_list_count;   // pinned
    bool smem_attr_set;
    const uint32_t *chunk_of_slot;   // slot -> 2 MB chunk of the HBM arena (owned by a uvmtier_pmm_t); NULL = arena is VA-sized
    uint64_t *d_lists_dst;           // destination offsets when they differ from the source offsets (chunk table); list_cap entries
};

static uvmtier_tiers *tiers_get(uvmtier_ctx_t *ctx)
{
    if (!ctx->tiers)
        ctx->tiers = (uvmtier_tiers *)calloc(1, sizeof(uvmtier_tiers));
    return ctx->tiers;
}

static void tier_release(uvmtier_ctx_t *ctx, uvmtier_tier *t)
{
    if (!t->used)
        return;
    cudaSetDevice(ctx->device);
    if (t->registered)
        cudaHostUnregister(t->info.host_ptr);
    if (t->mmapped)
        munmap(t->info.host_ptr, t->info.bytes);
    if (t->cuda_alloc)
        cudaFree(t->info.device_ptr);
    if (t->info.flags & UVMTIER_TIERF_IPC)
        cudaIpcCloseMemHandle(t->info.device_ptr);
    memset(t, 0, sizeof(*t));
}

void uvmtier_tiers_destroy(uvmtier_ctx_t *ctx)
{
    if (!ctx->tiers)
        return;
    for (uint32_t i = 0; i < UVMTIER_MAX_TIERS; i++)
        tier_release(ctx, &ctx->tiers->t[i]);
    if (ctx->tiers->d_lists)
        cudaFree(ctx->tiers->d_lists);
    if (ctx->tiers->d_lists_dst)
        cudaFree(ctx->tiers->d_lists_dst);
    if (ctx->tiers->d_list_count)
        cudaFree(ctx->tiers->d_list_count);
    if (ctx->tiers->h_list_count)
        cudaFreeHost(ctx->tiers->h_list_count);
    free(ctx->tiers);
    ctx->tiers = NULL;
}

static int read_small_file(const char *path, char *buf, size_t cap)
{
    FILE *f = fopen(path, "r");
    if (!f)
        return -1;
    size_t n = fread(buf, 1, cap - 1, f);
    fclose(f);
    buf[n] = 0;
    while (n && (buf[n - 1] == '\n' || buf[n - 1] == ' '))
        buf[--n] = 0;
    return (int)n;
}

extern "C" {

int uvmtier_numa_probe(int *out_dram_node, int *out_cxl_node)
{
    int dram = -1, cxl = -1, count = 0;
    DIR *d = opendir("/sys/devices/system/node");
    if (d) {
        struct dirent *e;
        while ((e = readdir(d))) {
            int id;
            if (sscanf(e->d_name, "node%d", &id) != 1)
                continue;
            char path[256], buf[4096];
            count++;
            snprintf(path, sizeof(path), "/sys/devices/system/node/node%d/meminfo", id);
            unsigned long long kb = 0;
            if (read_small_file(path, buf, sizeof(buf)) > 0) {
                const char *p = strstr(buf, "MemTotal:");
                if (p)
                    sscanf(p, "MemTotal: %llu", &kb);
            }
            snprintf(path, sizeof(path), "/sys/devices/system/node/node%d/cpulist", id);
            const int n = read_small_file(path, buf, sizeof(buf));
            const bool has_cpu = n > 0;
            if (kb > 0 && has_cpu && (dram < 0 || id < dram))
                dram = id;
            if (kb > 0 && !has_cpu && (cxl < 0 || id < cxl))
                cxl = id;            // memory-only node: how a CXL type-3 expander shows up
        }
        closedir(d);
    }
    if (out_dram_node)
        *out_dram_node = dram;
    if (out_cxl_node)
        *out_cxl_node = cxl;
    return count;
}

NV_STATUS uvmtier_tier_create(uvmtier_ctx_t *ctx, uint32_t tier, uint64_t bytes, int numa_node, uvmtier_tier_info_t *out)
{
    if (!ctx || tier >= UVMTIER_MAX_TIERS || !bytes || tier >= UVMTIER_TIER_PEER0)
        return NV_ERR_INVALID_ARGUMENT;
    uvmtier_tiers *ts = tiers_get(ctx);
    if (!ts)
        return NV_ERR_NO_MEMORY;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    uvmtier_tier *t = &ts->t[tier];
    tier_release(ctx, t);
    bytes = uvmtier_align_up(bytes, UVMTIER_VA_BLOCK_SIZE);
    t->info.bytes = bytes;
    t->info.numa_node = -1;
    if (tier == UVMTIER_TIER_HBM) {
        UVMTIER_CUDA(cudaMalloc(&t->info.device_ptr, bytes));
        t->cuda_alloc = true;
    }
    else {
        int dram = -1, cxl = -1;
        uvmtier_numa_probe(&dram, &cxl);
        if (numa_node < 0) {
            if (tier == UVMTIER_TIER_CXL) {
                numa_node = cxl >= 0 ? cxl : dram;
                if (cxl < 0)
                    t->info.flags |= UVMTIER_TIERF_CXL_EMULATED;     // said loudly, never silently
            }
            else {
                numa_node = dram;
            }
        }
        else if (tier == UVMTIER_TIER_CXL && numa_node != cxl) {
            t->info.flags |= UVMTIER_TIERF_CXL_EMULATED;
        }
        void *p = mmap(NULL, bytes, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
        if (p == MAP_FAILED)
            return NV_ERR_NO_MEMORY;
        t->mmapped = true;
        t->info.host_ptr = p;
        madvise(p, bytes, MADV_HUGEPAGE);
        if (numa_node >= 0 && numa_node < 64) {
            unsigned long mask = 1ul << numa_node;
            // MPOL_BIND = 2; strict node binding like alloc_pages_node(nid, __GFP_THISNODE)
            if (syscall(SYS_mbind, p, bytes, 2, &mask, sizeof(mask) * 8 + 1, 0) == 0)
                t->info.numa_node = numa_node;
        }
        cudaError_t e = cudaHostRegister(p, bytes, cudaHostRegisterPortable | cudaHostRegisterMapped);
        if (e != cudaSuccess) {
            uvmtier_set_error("cudaHostRegister", e, __FILE__, __LINE__);
            tier_release(ctx, t);
            return uvmtier_status_from_cuda(e);
        }
        t->registered = true;
        UVMTIER_CUDA(cudaHostGetDevicePointer(&t->info.device_ptr, p, 0));
    }
    t->used = true;
    if (out)
        *out = t->info;
    return NV_OK;
}

NV_STATUS uvmtier_tier_register(uvmtier_ctx_t *ctx, uint32_t tier, void *device_ptr, void *host_ptr, uint64_t bytes)
{
    if (!ctx || tier >= UVMTIER_MAX_TIERS || !device_ptr || !bytes)
        return NV_ERR_INVALID_ARGUMENT;
    uvmtier_tiers *ts = tiers_get(ctx);
    if (!ts)
        return NV_ERR_NO_MEMORY;
    uvmtier_tier *t = &ts->t[tier];
    tier_release(ctx, t);
    t->info.bytes = bytes;
    t->info.device_ptr = device_ptr;
    t->info.host_ptr = host_ptr;
    t->info.numa_node = -1;
    t->info.flags = UVMTIER_TIERF_EXTERNAL;
    t->used = true;
    return NV_OK;
}

NV_STATUS uvmtier_tier_map_peer(uvmtier_ctx_t *ctx, uint32_t peer_index, uvmtier_ctx_t *peer)
{
    if (!ctx || !peer || UVMTIER_TIER_PEER0 + peer_index >= UVMTIER_MAX_TIERS || !peer->tiers ||
        !peer->tiers->t[UVMTIER_TIER_HBM].used || ctx->device == peer->device)
        return NV_ERR_INVALID_ARGUMENT;
    int can = 0;
    UVMTIER_CUDA(cudaDeviceCanAccessPeer(&can, ctx->device, peer->device));
    if (!can)
        return NV_ERR_NOT_SUPPORTED;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    cudaError_t e = cudaDeviceEnablePeerAccess(peer->device, 0);
    if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
        uvmtier_set_error("cudaDeviceEnablePeerAccess", e, __FILE__, __LINE__);
        return uvmtier_status_from_cuda(e);
    }
    cudaGetLastError();
    const uvmtier_tier_info_t *pi = &peer->tiers->t[UVMTIER_TIER_HBM].info;
    return uvmtier_tier_register(ctx, UVMTIER_TIER_PEER0 + peer_index, pi->device_ptr, NULL, pi->bytes);
}

// One process per GPU (torchrun): the peer's HBM arena arrives as a CUDA IPC handle.
NV_STATUS uvmtier_tier_ipc_export(uvmtier_ctx_t *ctx, uint32_t tier, uint8_t handle[64])
{
    if (!ctx || tier >= UVMTIER_MAX_TIERS || !handle || !ctx->tiers || !ctx->tiers->t[tier].used || !ctx->tiers->t[tier].cuda_alloc)
        return NV_ERR_INVALID_ARGUMENT;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    cudaIpcMemHandle_t h;
    UVMTIER_CUDA(cudaIpcGetMemHandle(&h, ctx->tiers->t[tier].info.device_ptr));
    static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
    memcpy(handle, &h, 64);
    return NV_OK;
}

NV_STATUS uvmtier_tier_ipc_import(uvmtier_ctx_t *ctx, uint32_t tier, const uint8_t handle[64], uint64_t bytes)
{
    if (!ctx || tier < UVMTIER_TIER_PEER0 || tier >= UVMTIER_MAX_TIERS || !handle || !bytes)
        return NV_ERR_INVALID_ARGUMENT;
    uvmtier_tiers *ts = tiers_get(ctx);
    if (!ts)
        return NV_ERR_NO_MEMORY;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, 64);
    void *p = NULL;
    UVMTIER_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    NV_STATUS st = uvmtier_tier_register(ctx, tier, p, NULL, bytes);
    if (st == NV_OK)
        ts->t[tier].info.flags |= UVMTIER_TIERF_IPC;
    return st;
}

NV_STATUS uvmtier_tier_info(uvmtier_ctx_t *ctx, uint32_t tier, uvmtier_tier_info_t *out)
{
    if (!ctx || tier >= UVMTIER_MAX_TIERS || !out || !ctx->tiers || !ctx->tiers->t[tier].used)
        return NV_ERR_INVALID_ARGUMENT;
    *out = ctx->tiers->t[tier].info;
    return NV_OK;
}

NV_STATUS uvmtier_tier_destroy(uvmtier_ctx_t *ctx, uint32_t tier)
{
    if (!ctx || tier >= UVMTIER_MAX_TIERS || !ctx->tiers)
        return NV_ERR_INVALID_ARGUMENT;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    UVMTIER_CUDA(cudaStreamSynchronize(ctx->stream));
    tier_release(ctx, &ctx->tiers->t[tier]);
    return NV_OK;
}

static NV_STATUS launch_copy(uvmtier_ctx_t *ctx, const copy_list_params &p, int method)
{
    const uint64_t n_pieces = p.n_pages * p.pieces_per_page;
    if (!n_pieces)
        return NV_OK;
    if (method == 1) {
        uint64_t grid = n_pieces < (uint64_t)ctx->sm_count * 16 ? n_pieces : (uint64_t)ctx->sm_count * 16;
        UVMTIER_LAUNCH(k_copy_pages_ldst, (int)grid, 256, 0, ctx->stream, p);
        return NV_OK;
    }
    const size_t smem = (size_t)kStages * kStageBytes;
    if (!ctx->tiers->smem_attr_set) {
        UVMTIER_CUDA(cudaFuncSetAttribute(k_copy_pages_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        ctx->tiers->smem_attr_set = true;
    }
    const uint64_t items = (n_pieces + kPiecesPerStage - 1) / kPiecesPerStage;
    const uint64_t full = (uint64_t)ctx->sm_count * kCtasPerSm;        // persistent: every SM holds kCtasPerSm CTAs
    const uint64_t grid = items < full ? items : full;
    UVMTIER_LAUNCH(k_copy_pages_tma, (int)grid, kCopyThreads, smem, ctx->stream, p);
    return NV_OK;
}

static bool page_size_ok(uint32_t ps) { return ps == 4096u || ps == 65536u || ps == (2u << 20); }

NV_STATUS uvmtier_copy_page_list_device(uvmtier_ctx_t *ctx, uint32_t dst_tier, uint32_t src_tier, size_t n,
                                        const uint64_t *d_dst_off, const uint64_t *d_src_off, uint32_t page_size, int method)
{
    if (!ctx || dst_tier >= UVMTIER_MAX_TIERS || src_tier >= UVMTIER_MAX_TIERS || !page_size_ok(page_size))
        return NV_ERR_INVALID_ARGUMENT;
    if (!ctx->tiers || !ctx->tiers->t[dst_tier].used || !ctx->tiers->t[src_tier].used)
        return NV_ERR_INVALID_STATE;
    if (!n)
        return NV_OK;
    if (!d_dst_off || !d_src_off)
        return NV_ERR_INVALID_ARGUMENT;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    copy_list_params p;
    memset(&p, 0, sizeof(p));
    p.src_base = (const uint8_t *)ctx->tiers->t[src_tier].info.device_ptr;
    p.dst_base = (uint8_t *)ctx->tiers->t[dst_tier].info.device_ptr;
    p.src_off = d_src_off;
    p.dst_off = d_dst_off;
    p.n_pages = n;
    p.pieces_per_page = page_size / kPiece;
    return launch_copy(ctx, p, method);
}

NV_STATUS uvmtier_copy_page_list(uvmtier_ctx_t *ctx, uint32_t dst_tier, uint32_t src_tier, size_t n, const uint64_t *h_dst_off,
                                 const uint64_t *h_src_off, uint32_t page_size, int method)
{
    if (!ctx || (n && (!h_dst_off || !h_src_off)))
        return NV_ERR_INVALID_ARGUMENT;
    if (!n)
        return NV_OK;
    const size_t lb = uvmtier_align_up(n * 8, 256);
    NV_STATUS st = uvmtier_stage_reserve(ctx, 2 * lb, 0);
    if (st != NV_OK)
        return st;
    uint64_t *d_dst = (uint64_t *)ctx->d_stage, *d_src = (uint64_t *)((char *)ctx->d_stage + lb);
    UVMTIER_CUDA(cudaMemcpyAsync(d_dst, h_dst_off, n * 8, cudaMemcpyHostToDevice, ctx->stream));
    UVMTIER_CUDA(cudaMemcpyAsync(d_src, h_src_off, n * 8, cudaMemcpyHostToDevice, ctx->stream));
    st = uvmtier_copy_page_list_device(ctx, dst_tier, src_tier, n, d_dst, d_src, page_size, method);
    if (st != NV_OK)
        return st;
    UVMTIER_CUDA(cudaStreamSynchronize(ctx->stream));
    return NV_OK;
}

// The (source, destination) pairs in execution order: the peer GPUs first -- a page staged through DRAM (peer -> DRAM,
// then DRAM -> HBM, UVM/uvm_va_block.c:4754-4829) needs its first hop done before the DRAM -> HBM pair runs -- then
// DRAM, CXL, HBM, zero fill.
static const uint32_t kSrcTierOf[8] = {UVMTIER_TIER_DRAM, UVMTIER_TIER_CXL, UVMTIER_TIER_HBM, UVMTIER_TIER_PEER0,
                                       UVMTIER_TIER_PEER0 + 1, UVMTIER_TIER_PEER0 + 2, UVMTIER_TIER_PEER0 + 3, 0};
static const uint32_t kDstTierOf[3] = {UVMTIER_TIER_HBM, UVMTIER_TIER_DRAM, UVMTIER_TIER_CXL};
static const uint32_t kSrcOrder[8] = {3, 4, 5, 6, 0, 1, 2, 7};

static NV_STATUS copy_lists_reserve(uvmtier_ctx_t *ctx, uint64_t pages)
{
    uvmtier_tiers *ts = ctx->tiers;
    if (!ts->d_list_count)
        UVMTIER_CUDA(cudaMalloc(&ts->d_list_count, (3 * kRunPairs + 1) * sizeof(uint32_t)));
    if (!ts->h_list_count)
        UVMTIER_CUDA(cudaHostAlloc(&ts->h_list_count, kRunPairs * sizeof(uint32_t), cudaHostAllocPortable));
    if (pages > 0xffffffffull)
        return NV_ERR_INVALID_ARGUMENT;
    const bool two = ts->chunk_of_slot != NULL;
    if (pages > ts->list_cap || (two && !ts->d_lists_dst)) {
        UVMTIER_CUDA(cudaStreamSynchronize(ctx->stream));
        if (ts->d_lists) UVMTIER_CUDA(cudaFree(ts->d_lists));
        if (ts->d_lists_dst) UVMTIER_CUDA(cudaFree(ts->d_lists_dst));
        ts->d_lists = ts->d_lists_dst = NULL;
        const uint64_t keep = pages > ts->list_cap ? pages + pages / 4 + 4096 : ts->list_cap;
        ts->list_cap = 0;
        UVMTIER_CUDA(cudaMalloc(&ts->d_lists, keep * sizeof(uint64_t)));
        if (two)
            UVMTIER_CUDA(cudaMalloc(&ts->d_lists_dst, keep * sizeof(uint64_t)));
        ts->list_cap = (uint32_t)(keep < 0xffffffffull ? keep : 0xffffffffull);
    }
    return NV_OK;
}

} // extern "C"

void uvmtier_tiers_set_chunk_table(uvmtier_ctx_t *ctx, const uint32_t *chunk_of_slot)
{
    if (tiers_get(ctx))
        ctx->tiers->chunk_of_slot = chunk_of_slot;
}

extern "C" {

NV_STATUS uvmtier_copy_runs_device(uvmtier_ctx_t *ctx, const uvmtier_copy_run_t *d_runs, const uint32_t *d_n_runs,
                                   uint32_t run_cap, int method)
{
    if (!ctx || !d_runs || !d_n_runs)
        return NV_ERR_INVALID_ARGUMENT;
    if (!ctx->tiers)
        return NV_ERR_INVALID_STATE;
    if (!run_cap)
        return NV_OK;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    uvmtier_tiers *ts = ctx->tiers;
    NV_STATUS st = copy_lists_reserve(ctx, 0);
    if (st != NV_OK)
        return st;
    uint32_t *d_pair_count = ts->d_list_count, *d_pair_cursor = ts->d_list_count + kRunPairs;
    UVMTIER_CUDA(cudaMemsetAsync(ts->d_list_count, 0, 2 * kRunPairs * sizeof(uint32_t), ctx->stream));
    size_t grid = ((size_t)run_cap + 255) / 256, cap = (size_t)ctx->sm_count * 8;
    if (grid > cap)
        grid = cap;
    UVMTIER_LAUNCH(k_count_run_pages, (int)grid, 256, 0, ctx->stream, d_runs, d_n_runs, run_cap, d_pair_count);
    // the per-pair page counts size the list and the copy grids: one small D2H read (the reference also walks its
    // masks on the CPU)
    UVMTIER_CUDA(cudaMemcpyAsync(ts->h_list_count, d_pair_count, kRunPairs * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
    UVMTIER_CUDA(cudaStreamSynchronize(ctx->stream));
    run_pair_bases bases;
    uint64_t total = 0;
    for (uint32_t p = 0; p < kRunPairs; p++) {
        bases.base[p] = (uint32_t)total;
        total += ts->h_list_count[p];
    }
    if (!total)
        return NV_OK;
    st = copy_lists_reserve(ctx, total);
    if (st != NV_OK)
        return st;
    uint64_t *list_dst = ts->chunk_of_slot ? ts->d_lists_dst : ts->d_lists;
    UVMTIER_LAUNCH(k_expand_runs, (int)grid, 256, 0, ctx->stream, d_runs, d_n_runs, run_cap, bases, d_pair_cursor, ts->d_lists,
                   ts->list_cap, ts->chunk_of_slot ? ts->d_lists_dst : (uint64_t *)NULL, ts->chunk_of_slot, (const uint32_t *)NULL);
    for (uint32_t step = 0; step < kRunPairs; step++) {
        const uint32_t pair = kSrcOrder[step / 3] * 3 + step % 3;
        const uint32_t n = ts->h_list_count[pair];
        if (!n)
            continue;
        const uint32_t src = pair / 3, dst_tier = kDstTierOf[pair % 3];
        if (!ts->t[dst_tier].used)
            return NV_ERR_INVALID_STATE;
        uint8_t *dst_base = (uint8_t *)ts->t[dst_tier].info.device_ptr;
        if (src == UVMTIER_RUN_SRC_ZERO) {                // first touch: zero fill at the destination
            const uint32_t zg = n < (uint32_t)ctx->sm_count * 16 ? n : (uint32_t)ctx->sm_count * 16;
            UVMTIER_LAUNCH(k_zero_pages, (int)zg, 256, 0, ctx->stream, dst_base, (const uint64_t *)(list_dst + bases.base[pair]),
                           (uint64_t)n, 1u, (const uint32_t *)NULL, (const uint32_t *)NULL);
            continue;
        }
        const uint32_t src_tier = kSrcTierOf[src];
        if (src_tier == dst_tier) {
            uvmtier_set_error("uvmtier_copy_runs_device: a run names the same tier as source and destination",
                              cudaErrorInvalidValue, __FILE__, __LINE__);
            return NV_ERR_INVALID_ARGUMENT;
        }
        if (!ts->t[src_tier].used)
            return NV_ERR_INVALID_STATE;
        copy_list_params p;
        memset(&p, 0, sizeof(p));
        p.src_base = (const uint8_t *)ts->t[src_tier].info.device_ptr;
        p.dst_base = dst_base;
        p.src_off = ts->d_lists + bases.base[pair];
        p.dst_off = list_dst + bases.base[pair];
        p.n_pages = n;
        p.pieces_per_page = 1;
        st = launch_copy(ctx, p, method);
        if (st != NV_OK)
            return st;
    }
    return NV_OK;
}

} // extern "C"

// The same without a round trip to the host, for callers that enqueue many small lists back to back (the tiered loop):
// the page counts stay on the device, the list is sized for max_pages pages (more are dropped and reported through
// *d_overflow), and every pair named in pair_mask (bit = src * 3 + dst) gets a launch whether it has pages or not.
NV_STATUS uvmtier_copy_runs_enqueue(uvmtier_ctx_t *ctx, const uvmtier_copy_run_t *d_runs, const uint32_t *d_n_runs, uint32_t run_cap,
                                    uint64_t max_pages, uint32_t pair_mask)
{
    if (!ctx || !d_runs || !d_n_runs)
        return NV_ERR_INVALID_ARGUMENT;
    if (!ctx->tiers)
        return NV_ERR_INVALID_STATE;
    if (!run_cap || !max_pages)
        return NV_OK;
    uvmtier_tiers *ts = ctx->tiers;
    NV_STATUS st = copy_lists_reserve(ctx, max_pages);
    if (st != NV_OK)
        return st;
    uint32_t *d_pair_count = ts->d_list_count, *d_pair_cursor = ts->d_list_count + kRunPairs, *d_bases = ts->d_list_count + 2 * kRunPairs;
    UVMTIER_CUDA(cudaMemsetAsync(ts->d_list_count, 0, 2 * kRunPairs * sizeof(uint32_t), ctx->stream));
    size_t grid = ((size_t)run_cap + 255) / 256, cap = (size_t)ctx->sm_count * 8;
    if (grid > cap)
        grid = cap;
    UVMTIER_LAUNCH(k_count_run_pages, (int)grid, 256, 0, ctx->stream, d_runs, d_n_runs, run_cap, d_pair_count);
    UVMTIER_LAUNCH(k_pair_bases, 1, 32, 0, ctx->stream, (const uint32_t *)d_pair_count, d_bases);
    run_pair_bases unused;
    memset(&unused, 0, sizeof(unused));
    uint64_t *list_dst = ts->chunk_of_slot ? ts->d_lists_dst : ts->d_lists;
    UVMTIER_LAUNCH(k_expand_runs, (int)grid, 256, 0, ctx->stream, d_runs, d_n_runs, run_cap, unused, d_pair_cursor, ts->d_lists,
                   ts->list_cap, ts->chunk_of_slot ? ts->d_lists_dst : (uint64_t *)NULL, ts->chunk_of_slot, (const uint32_t *)d_bases);
    const size_t smem = (size_t)kStages * kStageBytes;
    if (!ts->smem_attr_set) {
        UVMTIER_CUDA(cudaFuncSetAttribute(k_copy_pages_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        ts->smem_attr_set = true;
    }
    for (uint32_t step = 0; step < kRunPairs; step++) {
        const uint32_t pair = kSrcOrder[step / 3] * 3 + step % 3;
        if (!((pair_mask >> pair) & 1u))
            continue;
        const uint32_t src = pair / 3, dst_tier = kDstTierOf[pair % 3];
        if (!ts->t[dst_tier].used)
            return NV_ERR_INVALID_STATE;
        uint8_t *dst_base = (uint8_t *)ts->t[dst_tier].info.device_ptr;
        if (src == UVMTIER_RUN_SRC_ZERO) {
            UVMTIER_LAUNCH(k_zero_pages, ctx->sm_count * 8, 256, 0, ctx->stream, dst_base, (const uint64_t *)list_dst, (uint64_t)0, 1u,
                           (const uint32_t *)(d_pair_count + pair), (const uint32_t *)(d_bases + pair));
            continue;
        }
        const uint32_t src_tier = kSrcTierOf[src];
        if (src_tier == dst_tier || !ts->t[src_tier].used)
            return NV_ERR_INVALID_STATE;
        copy_list_params p;
        memset(&p, 0, sizeof(p));
        p.src_base = (const uint8_t *)ts->t[src_tier].info.device_ptr;
        p.dst_base = dst_base;
        p.src_off = ts->d_lists;
        p.dst_off = list_dst;
        p.pieces_per_page = 1;
        p.n_pages_dev = d_pair_count + pair;
        p.list_base_dev = d_bases + pair;
        UVMTIER_LAUNCH(k_copy_pages_tma, ctx->sm_count * (int)kCtasPerSm, kCopyThreads, smem, ctx->stream, p);
    }
    return NV_OK;
}

extern "C" {

NV_STATUS uvmtier_checksum_pages_device(uvmtier_ctx_t *ctx, uint32_t tier, size_t n, const uint64_t *d_off, uint32_t page_size,
                                        uint64_t *d_out)
{
    if (!ctx || tier >= UVMTIER_MAX_TIERS || !page_size_ok(page_size))
        return NV_ERR_INVALID_ARGUMENT;
    if (!ctx->tiers || !ctx->tiers->t[tier].used)
        return NV_ERR_INVALID_STATE;
    if (!n)
        return NV_OK;
    if (!d_off || !d_out)
        return NV_ERR_INVALID_ARGUMENT;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    const size_t grid = n < (size_t)ctx->sm_count * 8 ? n : (size_t)ctx->sm_count * 8;
    UVMTIER_LAUNCH(k_checksum_pages, (int)grid, 256, 0, ctx->stream, (const uint8_t *)ctx->tiers->t[tier].info.device_ptr, d_off,
                   (uint64_t)n, page_size, d_out);
    return NV_OK;
}

NV_STATUS uvmtier_tier_memcpy_peak(uvmtier_ctx_t *ctx, uint32_t dst_tier, uint32_t src_tier, uint64_t bytes, int reps,
                                   float *out_best_ms)
{
    if (!ctx || dst_tier >= UVMTIER_MAX_TIERS || src_tier >= UVMTIER_MAX_TIERS || !out_best_ms || reps < 1)
        return NV_ERR_INVALID_ARGUMENT;
    if (!ctx->tiers || !ctx->tiers->t[dst_tier].used || !ctx->tiers->t[src_tier].used)
        return NV_ERR_INVALID_STATE;
    const uvmtier_tier_info_t *d = &ctx->tiers->t[dst_tier].info, *s = &ctx->tiers->t[src_tier].info;
    if (bytes > d->bytes || bytes > s->bytes)
        return NV_ERR_INVALID_ARGUMENT;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    cudaEvent_t a, b;
    UVMTIER_CUDA(cudaEventCreate(&a));
    UVMTIER_CUDA(cudaEventCreate(&b));
    float best = 1e30f;
    for (int r = 0; r < reps + 1; r++) {        // first one is a warm-up
        UVMTIER_CUDA(cudaEventRecord(a, ctx->stream));
        UVMTIER_CUDA(cudaMemcpyAsync(d->device_ptr, s->device_ptr, bytes, cudaMemcpyDefault, ctx->stream));
        UVMTIER_CUDA(cudaEventRecord(b, ctx->stream));
        UVMTIER_CUDA(cudaEventSynchronize(b));
        float ms = 0;
        UVMTIER_CUDA(cudaEventElapsedTime(&ms, a, b));
        if (r > 0 && ms < best)
            best = ms;
    }
    cudaEventDestroy(a);
    cudaEventDestroy(b);
    *out_best_ms = best;
    return NV_OK;
}

} // extern "C"

// ------------------------------------------------------------------------------------------------
// The fork's CXL P2P DMA surface (include/uvmtier.h section 3b; src/nvidia/src/kernel/mem_mgr/p2p_cxl.c).
// Registration pins the buffer and builds its page array; a DMA request is one contiguous range between the HBM
// tier and the buffer, moved by the page-copy pipeline above (4 KB pieces through TMA) with a small load/store
// kernel for what does not fill a piece or is not 16-byte aligned.
struct uvmtier_cxl_buffer {
    uint64_t magic;
    void *base;               // baseAddress
    uint8_t *device_base;     // the address the kernels use
    uint64_t size;
    uint32_t cxl_version;
    uint32_t page_count, page_size;
    uint64_t *pages;          // pPageArray
    bool registered, huge, pinned_here;
    void *pin_base;           // what cudaHostRegister was given: the whole pages around [base, base + size)
    int device;
};
static const uint64_t kCxlMagic = 0x43584C5032504255ull;   // "CXLP2PBU"
static std::atomic<uint32_t> g_cxl_count{0};               // g_cxlRegisteredBufferCount (p2p_cxl.c:143)
static std::atomic<uint64_t> g_cxl_bytes{0};               // g_cxlTotalRegisteredSize (:144)

// Bytes that the 4 KB pipeline does not take: any alignment, any length.
static __global__ void __launch_bounds__(256) k_copy_bytes(uint8_t *dst, const uint8_t *src, uint64_t n)
{
    const uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x, nth = (uint64_t)gridDim.x * blockDim.x;
    if ((((uintptr_t)dst | (uintptr_t)src) & 15u) == 0) {
        const uint64_t n16 = n / 16;
        for (uint64_t i = tid; i < n16; i += nth)
            ((uint4 *)dst)[i] = ((const uint4 *)src)[i];
        for (uint64_t i = n16 * 16 + tid; i < n; i += nth)
            dst[i] = src[i];
    }
    else {
        for (uint64_t i = tid; i < n; i += nth)
            dst[i] = src[i];
    }
}

extern "C" {

NV_STATUS uvmtier_p2p_get_cxl_system_info(uint32_t *num_devices, uint32_t *num_memory_devices, uint32_t *link_up,
                                          uint32_t *cxl_version)
{
    // nv_enumerate_cxl_devices walks the PCI bus for CXL DVSECs; what user space sees of a CXL memory expander is
    // a NUMA node with memory and no CPUs
    uint32_t n = 0;
    DIR *d = opendir("/sys/devices/system/node");
    if (d) {
        struct dirent *e;
        while ((e = readdir(d)) != NULL) {
            int node;
            if (sscanf(e->d_name, "node%d", &node) != 1)
                continue;
            char path[128], buf[256];
            snprintf(path, sizeof(path), "/sys/devices/system/node/node%d/cpulist", node);
            const int got = read_small_file(path, buf, sizeof(buf));
            bool has_cpus = false;
            for (int i = 0; i < got; i++)
                if (buf[i] >= '0' && buf[i] <= '9')
                    has_cpus = true;
            if (got >= 0 && !has_cpus)
                n++;
        }
        closedir(d);
    }
    if (num_devices) *num_devices = n;
    if (num_memory_devices) *num_memory_devices = n;
    if (link_up) *link_up = n > 0;
    if (cxl_version) *cxl_version = 2;                     // the reference's default (:118-126)
    return NV_OK;
}

NV_STATUS uvmtier_p2p_register_cxl_buffer(uvmtier_ctx_t *ctx, void *base, uint64_t size, uint32_t cxl_version,
                                          uvmtier_cxl_buffer_t **out_handle)
{
    const uint64_t addr = (uint64_t)(uintptr_t)base;
    if (!ctx || size == 0 || !out_handle || cxl_version < 1 || cxl_version > 3)       // :241-246
        return NV_ERR_INVALID_ARGUMENT;
    if (size > UVMTIER_CXL_P2P_MAX_BUFFER_SIZE)                                       // :248-253
        return NV_ERR_INVALID_ARGUMENT;
    if (addr + size < addr)                                                           // :255-260
        return NV_ERR_INVALID_ARGUMENT;
    if (g_cxl_count.load() >= UVMTIER_CXL_P2P_MAX_REGISTERED_BUFFERS)                 // :262-267
        return NV_ERR_INSUFFICIENT_RESOURCES;
    if (!base)
        return NV_ERR_INVALID_ADDRESS;                                                // pinning address 0 fails
    uvmtier_cxl_buffer *h = (uvmtier_cxl_buffer *)calloc(1, sizeof(*h));
    if (!h)
        return NV_ERR_NO_MEMORY;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    // pin + map: memory that is registered already (one of this library's host tiers, or the caller's own
    // cudaHostRegister) is used as it is
    void *dptr = NULL;
    if (cudaHostGetDevicePointer(&dptr, base, 0) != cudaSuccess) {
        cudaGetLastError();
        h->pin_base = (void *)(uintptr_t)(addr & ~(uint64_t)4095);
        const uint64_t pin_bytes = ((addr + size + 4095) & ~(uint64_t)4095) - (addr & ~(uint64_t)4095);
        const cudaError_t e = cudaHostRegister(h->pin_base, pin_bytes, cudaHostRegisterMapped | cudaHostRegisterPortable);
        if (e != cudaSuccess) {
            uvmtier_set_error("cudaHostRegister (CXL buffer)", e, __FILE__, __LINE__);
            cudaGetLastError();
            free(h);
            return e == cudaErrorMemoryAllocation ? NV_ERR_NO_MEMORY : NV_ERR_INVALID_ADDRESS;
        }
        h->pinned_here = true;
        if (cudaHostGetDevicePointer(&dptr, base, 0) != cudaSuccess) {
            cudaGetLastError();
            cudaHostUnregister(h->pin_base);
            free(h);
            return NV_ERR_INVALID_STATE;
        }
    }
    // _cxlP2PCanUseHugePages (:149-160), else 4 KB pages spanning the range (nv-p2p.c:1166-1167)
    h->huge = (addr & (UVMTIER_CXL_P2P_PAGE_SIZE_2M - 1)) == 0 && (size & (UVMTIER_CXL_P2P_PAGE_SIZE_2M - 1)) == 0 &&
              size >= UVMTIER_CXL_P2P_PAGE_SIZE_2M;
    h->page_size = h->huge ? UVMTIER_CXL_P2P_PAGE_SIZE_2M : UVMTIER_CXL_P2P_PAGE_SIZE_4K;
    const uint64_t start = h->huge ? addr : (addr & ~(uint64_t)4095);
    const uint64_t count = h->huge ? size / UVMTIER_CXL_P2P_PAGE_SIZE_2M : ((addr + size - 1) >> 12) - (start >> 12) + 1;
    if (count == 0 || count > (1ull << 28)) {                                         // nv-p2p.c:1169-1174
        if (h->pinned_here) cudaHostUnregister(h->pin_base);
        free(h);
        return NV_ERR_INVALID_ARGUMENT;
    }
    h->pages = (uint64_t *)malloc(count * sizeof(uint64_t));
    if (!h->pages) {
        if (h->pinned_here) cudaHostUnregister(h->pin_base);
        free(h);
        return NV_ERR_NO_MEMORY;
    }
    for (uint64_t i = 0; i < count; i++)
        h->pages[i] = start + i * h->page_size;
    h->magic = kCxlMagic;
    h->base = base;
    h->device_base = (uint8_t *)dptr;
    h->size = size;
    h->cxl_version = cxl_version;
    h->page_count = (uint32_t)count;
    h->registered = true;
    h->device = ctx->device;
    g_cxl_count.fetch_add(1);
    g_cxl_bytes.fetch_add(size);
    *out_handle = h;
    return NV_OK;
}

NV_STATUS uvmtier_p2p_unregister_cxl_buffer(uvmtier_ctx_t *ctx, uvmtier_cxl_buffer_t *h)
{
    if (!h || h->magic != kCxlMagic)
        return NV_ERR_INVALID_ARGUMENT;
    if (!h->registered)
        return NV_ERR_INVALID_STATE;
    h->registered = false;
    if (ctx) {
        cudaSetDevice(ctx->device);
        cudaStreamSynchronize(ctx->stream);                // a request may still be in flight (ASYNC)
    }
    if (h->pinned_here)
        cudaHostUnregister(h->pin_base);
    free(h->pages);
    h->pages = NULL;
    const uint64_t size = h->size;
    h->magic = 0;
    free(h);
    if (g_cxl_count.load() > 0)
        g_cxl_count.fetch_sub(1);
    if (g_cxl_bytes.load() >= size)
        g_cxl_bytes.fetch_sub(size);
    else
        g_cxl_bytes.store(0);
    return NV_OK;
}

NV_STATUS uvmtier_p2p_get_cxl_pages(uvmtier_cxl_buffer_t *h, uint64_t offset, uint64_t size, uint64_t *page_addrs,
                                    uint32_t *page_count, uint32_t *page_size)
{
    if (!h || h->magic != kCxlMagic || !page_addrs || !page_count || !page_size)
        return NV_ERR_INVALID_ARGUMENT;
    if (!h->registered)
        return NV_ERR_INVALID_STATE;
    if (offset + size > h->size)
        return NV_ERR_INVALID_ARGUMENT;
    const uint32_t start = (uint32_t)(offset / h->page_size);
    const uint32_t n = (uint32_t)((offset + size + h->page_size - 1) / h->page_size) - start;
    if (start + n > h->page_count)
        return NV_ERR_INVALID_ARGUMENT;
    for (uint32_t i = 0; i < n; i++)
        page_addrs[i] = h->pages[start + i];
    *page_count = n;
    *page_size = h->page_size;
    return NV_OK;
}

NV_STATUS uvmtier_p2p_put_cxl_pages(uvmtier_cxl_buffer_t *h, uint64_t offset, uint64_t size)
{
    (void)offset; (void)size;
    if (!h || h->magic != kCxlMagic)
        return NV_ERR_INVALID_ARGUMENT;
    if (!h->registered)
        return NV_ERR_INVALID_STATE;
    return NV_OK;                                          // pages stay pinned until unregistration (:506-513)
}

NV_STATUS uvmtier_p2p_cxl_dma_request(uvmtier_ctx_t *ctx, uvmtier_cxl_buffer_t *h, uint64_t gpu_offset, uint64_t cxl_offset,
                                      uint64_t size, uint32_t flags, uint64_t *out_bytes)
{
    if (out_bytes)
        *out_bytes = 0;
    if (!ctx || !h || h->magic != kCxlMagic || size == 0)                             // :535-551
        return NV_ERR_INVALID_ARGUMENT;
    if (!h->registered)                                                               // :553-557
        return NV_ERR_INVALID_STATE;
    if (cxl_offset > h->size || size > h->size - cxl_offset)                          // :559-565
        return NV_ERR_INVALID_ARGUMENT;
    if (!ctx->tiers || !ctx->tiers->t[UVMTIER_TIER_HBM].used)
        return NV_ERR_INVALID_STATE;
    const uvmtier_tier_info_t *hbm = &ctx->tiers->t[UVMTIER_TIER_HBM].info;
    uint64_t bytes = size;
    if (bytes > 0xFFFFFFFFull)                                                        // :617-622
        bytes = UVMTIER_CXL_P2P_DMA_MAX_TRANSFER;
    if (gpu_offset > hbm->bytes || bytes > hbm->bytes - gpu_offset)
        return NV_ERR_INVALID_ARGUMENT;
    UVMTIER_CUDA(cudaSetDevice(ctx->device));
    const bool to_gpu = (flags & UVMTIER_CXL_P2P_DMA_FLAG_CXL_TO_GPU) != 0;
    uint8_t *gpu = (uint8_t *)hbm->device_ptr + gpu_offset, *cxl = h->device_base + cxl_offset;
    uint8_t *dst = to_gpu ? gpu : cxl;
    const uint8_t *src = to_gpu ? cxl : gpu;
    uint64_t body = 0;
    if ((((uintptr_t)dst | (uintptr_t)src) & 15u) == 0) {
        // whole 4 KB pieces through the TMA pipeline: 2 MB "pages" first (one list-free linear range), then 4 KB ones
        copy_list_params p;
        memset(&p, 0, sizeof(p));
        const uint64_t big = bytes >> 21, small = (bytes - (big << 21)) >> 12;
        if (big) {
            p.src_base = src; p.dst_base = dst; p.n_pages = big; p.pieces_per_page = 512;
            const NV_STATUS st = launch_copy(ctx, p, 0);
            if (st != NV_OK)
                return st;
        }
        if (small) {
            p.src_base = src + (big << 21); p.dst_base = dst + (big << 21); p.n_pages = small; p.pieces_per_page = 1;
            const NV_STATUS st = launch_copy(ctx, p, 0);
            if (st != NV_OK)
                return st;
        }
        body = (big << 21) + (small << 12);
    }
    if (body < bytes) {
        const uint64_t rest = bytes - body;
        uint64_t grid = (rest / 16 + 255) / 256 + 1;
        if (grid > (uint64_t)ctx->sm_count * 8)
            grid = (uint64_t)ctx->sm_count * 8;
        UVMTIER_LAUNCH(k_copy_bytes, (int)grid, 256, 0, ctx->stream, dst + body, src + body, rest);
    }
    if (out_bytes)
        *out_bytes = bytes;
    if (!(flags & UVMTIER_CXL_P2P_DMA_FLAG_ASYNC))
        UVMTIER_CUDA(cudaStreamSynchronize(ctx->stream));
    return NV_OK;
}

void uvmtier_p2p_cxl_registered(uint32_t *count, uint64_t *bytes)
{
    if (count) *count = g_cxl_count.load();
    if (bytes) *bytes = g_cxl_bytes.load();
}

} // extern "C"
