// Dependent-issue latencies (cycles) of the instructions the warp-per-QP solver is built from.
//   nvcc -gencode arch=compute_100a,code=sm_100a -o tools/probe/lat_probe tools/probe/lat_probe.cu
#include <cstdio>
#include <cuda_runtime.h>
#define N 512
__device__ __forceinline__ void dmma(double2& c, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c.x), "+d"(c.y) : "d"(a), "d"(b));
}
__global__ void k(long long* out, double* sink, const double* g, double a, double b) {
  __shared__ double sm[1024];
  const int lane = threadIdx.x;
  for (int i = lane; i < 1024; i += 32) sm[i] = (double)((i * 7 + 8) & 1023);
  __syncwarp();
  long long t0, t1;
  double x = a + lane;
  // DFMA chain
  t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < N; ++i) x = fma(x, a, b);
  t1 = clock64();
  if (lane == 0) out[0] = (t1 - t0);
  // DMMA chain through the accumulator
  double2 c = make_double2(x, x);
  t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < N; ++i) dmma(c, a, b);
  t1 = clock64();
  if (lane == 0) out[1] = (t1 - t0);
  // DMMA chain through the A operand (accumulator -> operand)
  t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < N; ++i) { double2 d = make_double2(0.0, 0.0); dmma(d, c.x, b); c = d; }
  t1 = clock64();
  if (lane == 0) out[2] = (t1 - t0);
  // 4 independent DMMA chains (throughput per warp)
  double2 c1 = c, c2 = c, c3 = c;
  t0 = clock64();
#pragma unroll 4
  for (int i = 0; i < N; ++i) { dmma(c, a, b); dmma(c1, a, b); dmma(c2, a, b); dmma(c3, a, b); }
  t1 = clock64();
  if (lane == 0) out[3] = (t1 - t0);
  x = c.x + c1.x + c2.x + c3.x;
  // rcp.approx.ftz.f64 chain
  t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < N; ++i) asm volatile("rcp.approx.ftz.f64 %0, %0;" : "+d"(x));
  t1 = clock64();
  if (lane == 0) out[4] = (t1 - t0);
  // LDS.64 pointer chase
  int idx = lane;
  t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < N; ++i) idx = (int)sm[idx & 1023];
  t1 = clock64();
  if (lane == 0) out[5] = (t1 - t0);
  // 64-bit shuffle chain
  double y = x + idx;
  t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < N; ++i) y = __shfl_xor_sync(0xffffffffu, y, 1);
  t1 = clock64();
  if (lane == 0) out[6] = (t1 - t0);
  // L2 pointer chase (ld.global.cg)
  long long p = lane;
  t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < N; ++i) p = (long long)__ldcg(g + (p & 65535));
  t1 = clock64();
  if (lane == 0) out[7] = (t1 - t0);
  // STS -> syncwarp -> LDS round trip
  double w = y;
  t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < N; ++i) { sm[lane] = w; __syncwarp(); w = sm[(lane + 1) & 31]; __syncwarp(); }
  t1 = clock64();
  if (lane == 0) out[8] = (t1 - t0);
  // redux.sync
  unsigned u = (unsigned)p + lane;
  t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < N; ++i) u = __reduce_max_sync(0xffffffffu, u) + lane;
  t1 = clock64();
  if (lane == 0) out[9] = (t1 - t0);
  sink[lane] = x + y + w + (double)p + (double)u;
}
int main() {
  long long* o; double* s; double* g;
  cudaMalloc(&o, 80); cudaMalloc(&s, 256); cudaMalloc(&g, 65536 * 8);
  double* hg = new double[65536];
  for (int i = 0; i < 65536; ++i) hg[i] = (double)((i * 4099 + 17) & 65535);
  cudaMemcpy(g, hg, 65536 * 8, cudaMemcpyHostToDevice);
  for (int r = 0; r < 2; ++r) k<<<1, 32>>>(o, s, g, 1.0000001, 1e-9);
  cudaDeviceSynchronize();
  long long h[10]; cudaMemcpy(h, o, 80, cudaMemcpyDeviceToHost);
  const char* nm[10] = {"DFMA dep", "DMMA dep (acc)", "DMMA dep (operand)", "DMMA x4 indep (per 4)", "rcp.approx.f64", "LDS.64 chase",
                        "SHFL 64-bit", "L2 ld.cg chase", "STS+sync+LDS+sync", "redux.sync+add"};
  for (int i = 0; i < 10; ++i) printf("%-24s %.1f cycles\n", nm[i], (double)h[i] / N);
  return 0;
}
