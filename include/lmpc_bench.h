/*
 * lmpc_bench.h — measurement helpers exported by liblmpc.so (not part of the drop-in ABI).
 * MEASURED_PEAKS.json holds HBM and bf16 peaks only; the convex-MPC path is FP64-compute
 * bound, so its roofline denominator is measured on the box by these micro-kernels.
 */
#ifndef LMPC_BENCH_H_
#define LMPC_BENCH_H_
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

enum {
  LMPC_MB_DFMA = 0,        /* scalar FP64 FMA chains */
  LMPC_MB_DMMA_M8N8K4 = 1, /* mma.sync.aligned.m8n8k4.f64 */
  LMPC_MB_DMMA_M16N8K4 = 2,
  LMPC_MB_DMMA_M16N8K8 = 3,
  LMPC_MB_DMMA_M16N8K16 = 4
};

/* Runs `reps` timed launches (after 2 warm-ups) of a register-resident FP64 throughput kernel
 * over all SMs of `device`; returns the best TFLOP/s (2 flops per FMA) in *tflops.
 * Returns 0 on success, non-zero on CUDA error. */
int32_t lmpc_microbench_fp64(int32_t device, int32_t kind, int32_t reps, double* tflops);

#ifdef __cplusplus
}
#endif
#endif
