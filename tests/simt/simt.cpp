// simt.cpp -- launch loop of the CPU SIMT emulator (see simt.h).  TEST INFRASTRUCTURE ONLY.
#include "simt.h"

thread_local dim3 threadIdx, blockIdx;
dim3 blockDim, gridDim;

namespace simt {
thread_local Warp *tl_warp;
thread_local Cta *tl_cta;
thread_local int tl_lane;
alignas(128) char g_dynamic_smem[256 * 1024];

struct ThreadArg {
    int tid;
    Cta *cta;
    Warp *warp;
    const std::function<void()> *fn;
};

static void *thread_main(void *p)
{
    ThreadArg *a = (ThreadArg *)p;
    tl_cta = a->cta;
    tl_warp = a->warp;
    tl_lane = a->tid & 31;
    threadIdx = dim3((unsigned)a->tid, 0, 0);
    for (unsigned b = 0; b < gridDim.x; b++) {
        blockIdx = dim3(b, 0, 0);
        (*a->fn)();
        pthread_barrier_wait(&a->cta->bar);   // CTAs run one after the other (shared memory is static)
    }
    return NULL;
}

void launch(dim3 grid, dim3 block, const std::function<void()> &fn)
{
    if (block.x % 32) {
        fprintf(stderr, "simt: blockDim.x must be a multiple of 32\n");
        abort();
    }
    gridDim = grid;
    blockDim = block;
    const int n = (int)block.x, nw = n / 32;
    Cta cta;
    pthread_barrier_init(&cta.bar, NULL, n);
    pthread_mutex_init(&cta.mu, NULL);
    std::vector<Warp> warps(nw);
    for (auto &w : warps)
        pthread_barrier_init(&w.bar, NULL, 32);
    std::vector<pthread_t> th(n);
    std::vector<ThreadArg> args(n);
    pthread_attr_t attr;
    pthread_attr_init(&attr);
    pthread_attr_setstacksize(&attr, 1 << 20);
    for (int t = 0; t < n; t++) {
        args[t] = ThreadArg{t, &cta, &warps[t / 32], &fn};
        if (pthread_create(&th[t], &attr, thread_main, &args[t])) {
            fprintf(stderr, "simt: pthread_create failed\n");
            abort();
        }
    }
    for (int t = 0; t < n; t++)
        pthread_join(th[t], NULL);
    for (auto &w : warps)
        pthread_barrier_destroy(&w.bar);
    pthread_barrier_destroy(&cta.bar);
}
} // namespace simt
