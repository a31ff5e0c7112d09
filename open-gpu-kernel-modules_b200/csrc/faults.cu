// faults.cu -- K0/K1 fault-batch preprocessing (placeholder until the kernels land).
#include "common.cuh"
void uvmtier_fault_state_destroy(uvmtier_ctx_t *ctx) { (void)ctx; }
