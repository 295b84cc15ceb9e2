// Probe of the FP64 mma.sync fragment layouts on sm_100a (m16n8k8, m16n8k16) against the
// "C-layout pair is directly an operand" convention of lmpc_wmma.cuh.  Build + run on the GPU box:
//   nvcc -gencode arch=compute_100a,code=sm_100a -o /tmp/mma_probe tools/probe/mma_probe.cu && /tmp/mma_probe
#include <cstdio>
#include <cmath>
#include <cuda_runtime.h>

__global__ void probe(const double* X, const double* Y, double* out, int hyp) {
  // X: 4 tiles 8x8 row-major (X1,X2 for k-chunk A; X3,X4 second k tile), Y: 2 tiles
  const int lane = threadIdx.x, g = lane >> 2, t = lane & 3;
  auto x = [&](int tile, int r, int c) { return X[tile * 64 + r * 8 + c]; };
  auto y = [&](int tile, int r, int c) { return Y[tile * 64 + r * 8 + c]; };
  // m16n8k8: [X0; X1] * Y0^T
  {
    double a0 = x(0, g, 2 * t), a2 = x(0, g, 2 * t + 1), a1 = x(1, g, 2 * t), a3 = x(1, g, 2 * t + 1);
    if (hyp == 1) { double tmp = a1; a1 = a2; a2 = tmp; }
    double b0 = y(0, g, 2 * t), b1 = y(0, g, 2 * t + 1);
    double c0 = 0, c1 = 0, c2 = 0, c3 = 0;
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+d"(c0), "+d"(c1), "+d"(c2), "+d"(c3) : "d"(a0), "d"(a1), "d"(a2), "d"(a3), "d"(b0), "d"(b1));
    out[0 * 64 + g * 8 + 2 * t] = c0; out[0 * 64 + g * 8 + 2 * t + 1] = c1;
    out[1 * 64 + g * 8 + 2 * t] = c2; out[1 * 64 + g * 8 + 2 * t + 1] = c3;
  }
  // m16n8k16: [X0 X2; X1 X3] * [Y0 Y1]^T
  {
    double a[8], b[4];
    a[0] = x(0, g, 2 * t); a[2] = x(0, g, 2 * t + 1); a[4] = x(2, g, 2 * t); a[6] = x(2, g, 2 * t + 1);
    a[1] = x(1, g, 2 * t); a[3] = x(1, g, 2 * t + 1); a[5] = x(3, g, 2 * t); a[7] = x(3, g, 2 * t + 1);
    b[0] = y(0, g, 2 * t); b[1] = y(0, g, 2 * t + 1); b[2] = y(1, g, 2 * t); b[3] = y(1, g, 2 * t + 1);
    double c0 = 0, c1 = 0, c2 = 0, c3 = 0;
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};"
                 : "+d"(c0), "+d"(c1), "+d"(c2), "+d"(c3)
                 : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]),
                   "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(b[3]));
    out[2 * 64 + g * 8 + 2 * t] = c0; out[2 * 64 + g * 8 + 2 * t + 1] = c1;
    out[3 * 64 + g * 8 + 2 * t] = c2; out[3 * 64 + g * 8 + 2 * t + 1] = c3;
  }
}

int main() {
  double hX[256], hY[128], hO[256];
  for (int i = 0; i < 256; ++i) hX[i] = sin(0.37 * i + 1.0);
  for (int i = 0; i < 128; ++i) hY[i] = cos(0.53 * i + 2.0);
  double *dX, *dY, *dO;
  cudaMalloc(&dX, sizeof hX); cudaMalloc(&dY, sizeof hY); cudaMalloc(&dO, sizeof hO);
  cudaMemcpy(dX, hX, sizeof hX, cudaMemcpyHostToDevice); cudaMemcpy(dY, hY, sizeof hY, cudaMemcpyHostToDevice);
  for (int hyp = 0; hyp < 2; ++hyp) {
    probe<<<1, 32>>>(dX, dY, dO, hyp);
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("cuda error\n"); return 1; }
    cudaMemcpy(hO, dO, sizeof hO, cudaMemcpyDeviceToHost);
    double e8 = 0, e16 = 0;
    for (int tile = 0; tile < 2; ++tile)
      for (int r = 0; r < 8; ++r)
        for (int c = 0; c < 8; ++c) {
          double s8 = 0, s16 = 0;
          for (int k = 0; k < 8; ++k) {
            s8 += hX[tile * 64 + r * 8 + k] * hY[c * 8 + k];
            s16 += hX[tile * 64 + r * 8 + k] * hY[c * 8 + k] + hX[(tile + 2) * 64 + r * 8 + k] * hY[64 + c * 8 + k];
          }
          e8 = fmax(e8, fabs(s8 - hO[tile * 64 + r * 8 + c]));
          e16 = fmax(e16, fabs(s16 - hO[(tile + 2) * 64 + r * 8 + c]));
        }
    printf("hyp %d: m16n8k8 max err %.3e   m16n8k16 max err %.3e\n", hyp, e8, e16);
  }
  return 0;
}
