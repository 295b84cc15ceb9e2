// lmpc_microbench.cu — FP64 peak micro-kernels (DFMA and DMMA shapes) for the roofline
// denominator (include/lmpc_bench.h).  Register-resident, no memory traffic.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/lmpc_bench.h"

namespace {

constexpr int kIters = 4096;

__global__ void __launch_bounds__(256) dfma_kernel(double* out, double a, double b) {
  double x[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) x[i] = (double)(threadIdx.x + i);
#pragma unroll 4
  for (int it = 0; it < kIters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) x[i] = fma(x[i], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += x[i];
  if (s == 123.456) out[0] = s;
}

template <int KIND>
__global__ void __launch_bounds__(256) dmma_kernel(double* out, double a, double b) {
  constexpr int NACC = 4;
  double c[NACC][4];
#pragma unroll
  for (int i = 0; i < NACC; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) c[i][j] = (double)(threadIdx.x + i + j);
  double a0 = a, a1 = a + 1, a2 = a + 2, a3 = a + 3, a4 = a + 4, a5 = a + 5, a6 = a + 6, a7 = a + 7;
  double b0 = b, b1 = b + 1, b2 = b + 2, b3 = b + 3;
#pragma unroll 2
  for (int it = 0; it < kIters; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i) {
      if (KIND == LMPC_MB_DMMA_M8N8K4) {
        asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                     : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a0), "d"(b0));
      } else if (KIND == LMPC_MB_DMMA_M16N8K4) {
        asm volatile("mma.sync.aligned.m16n8k4.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                     : "+d"(c[i][0]), "+d"(c[i][1]), "+d"(c[i][2]), "+d"(c[i][3]) : "d"(a0), "d"(a1), "d"(b0));
      } else if (KIND == LMPC_MB_DMMA_M16N8K8) {
        asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+d"(c[i][0]), "+d"(c[i][1]), "+d"(c[i][2]), "+d"(c[i][3])
                     : "d"(a0), "d"(a1), "d"(a2), "d"(a3), "d"(b0), "d"(b1));
      } else {
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};"
                     : "+d"(c[i][0]), "+d"(c[i][1]), "+d"(c[i][2]), "+d"(c[i][3])
                     : "d"(a0), "d"(a1), "d"(a2), "d"(a3), "d"(a4), "d"(a5), "d"(a6), "d"(a7),
                       "d"(b0), "d"(b1), "d"(b2), "d"(b3));
      }
    }
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < NACC; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) s += c[i][j];
  if (s == 123.456) out[0] = s;
}

}  // namespace

extern "C" int32_t lmpc_microbench_fp64(int32_t device, int32_t kind, int32_t reps, double* tflops) {
  if (!tflops || reps < 1) return 1;
  if (cudaSetDevice(device) != cudaSuccess) return 2;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return 2;
  double* d = nullptr;
  if (cudaMalloc(&d, 64) != cudaSuccess) return 2;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int blocks = prop.multiProcessorCount * 8, threads = 256;
  double fma_per_thread_iter = 0;
  switch (kind) {
    case LMPC_MB_DFMA: fma_per_thread_iter = 8; break;
    case LMPC_MB_DMMA_M8N8K4: fma_per_thread_iter = 4 * (8.0 * 8 * 4) / 32; break;
    case LMPC_MB_DMMA_M16N8K4: fma_per_thread_iter = 4 * (16.0 * 8 * 4) / 32; break;
    case LMPC_MB_DMMA_M16N8K8: fma_per_thread_iter = 4 * (16.0 * 8 * 8) / 32; break;
    case LMPC_MB_DMMA_M16N8K16: fma_per_thread_iter = 4 * (16.0 * 8 * 16) / 32; break;
    default: cudaFree(d); return 1;
  }
  double best = 0;
  for (int r = 0; r < reps + 2; ++r) {
    cudaEventRecord(e0);
    switch (kind) {
      case LMPC_MB_DFMA: dfma_kernel<<<blocks, threads>>>(d, 1.0000001, 1e-9); break;
      case LMPC_MB_DMMA_M8N8K4: dmma_kernel<LMPC_MB_DMMA_M8N8K4><<<blocks, threads>>>(d, 1e-3, 1e-3); break;
      case LMPC_MB_DMMA_M16N8K4: dmma_kernel<LMPC_MB_DMMA_M16N8K4><<<blocks, threads>>>(d, 1e-3, 1e-3); break;
      case LMPC_MB_DMMA_M16N8K8: dmma_kernel<LMPC_MB_DMMA_M16N8K8><<<blocks, threads>>>(d, 1e-3, 1e-3); break;
      default: dmma_kernel<LMPC_MB_DMMA_M16N8K16><<<blocks, threads>>>(d, 1e-3, 1e-3); break;
    }
    cudaEventRecord(e1);
    if (cudaEventSynchronize(e1) != cudaSuccess) { cudaFree(d); return 2; }
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    if (r >= 2 && ms > 0) {
      const double fl = 2.0 * fma_per_thread_iter * kIters * (double)blocks * threads;
      const double tf = fl / (ms * 1e-3) / 1e12;
      if (tf > best) best = tf;
    }
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  cudaFree(d);
  *tflops = best;
  return cudaGetLastError() == cudaSuccess ? 0 : 2;
}
