// migrate.cu -- K3 page-list migration (placeholder until the kernels land).
#include "common.cuh"
void uvmtier_tiers_destroy(uvmtier_ctx_t *ctx) { (void)ctx; }
