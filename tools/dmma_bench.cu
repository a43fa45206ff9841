// Microbenchmark: FP64 tensor (DMMA) issue rate per shape vs plain DFMA on sm_100a.
// Used once to pick the mma.sync shape for the trailing update and to get the
// measured FP64 ceiling for the roofline denominator.  Not part of the product path.
#include <cstdio>
#include <cuda_runtime.h>

template <int SHAPE, int NACC>
__global__ void __launch_bounds__(256) k_dmma(double* out, int iters, double seed) {
  double c[NACC][4];
  double a[8], b[4];
#pragma unroll
  for (int i = 0; i < 8; ++i) a[i] = seed + threadIdx.x * 1e-9 + i;
#pragma unroll
  for (int i = 0; i < 4; ++i) b[i] = seed * 0.5 + i;
#pragma unroll
  for (int j = 0; j < NACC; ++j)
#pragma unroll
    for (int i = 0; i < 4; ++i) c[j][i] = 0.0;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int j = 0; j < NACC; ++j) {
      if (SHAPE == 884) {
        asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1},{%2},{%3},{%0,%1};\n"
                     : "+d"(c[j][0]), "+d"(c[j][1]) : "d"(a[0]), "d"(b[0]));
      } else if (SHAPE == 1684) {
        asm volatile("mma.sync.aligned.m16n8k4.row.col.f64.f64.f64.f64 {%0,%1,%2,%3},{%4,%5},{%6},{%0,%1,%2,%3};\n"
                     : "+d"(c[j][0]), "+d"(c[j][1]), "+d"(c[j][2]), "+d"(c[j][3])
                     : "d"(a[0]), "d"(a[1]), "d"(b[0]));
      } else if (SHAPE == 1688) {
        asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3},{%4,%5,%6,%7},{%8,%9},{%0,%1,%2,%3};\n"
                     : "+d"(c[j][0]), "+d"(c[j][1]), "+d"(c[j][2]), "+d"(c[j][3])
                     : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(b[0]), "d"(b[1]));
      } else {
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3},{%4,%5,%6,%7,%8,%9,%10,%11},{%12,%13,%14,%15},{%0,%1,%2,%3};\n"
                     : "+d"(c[j][0]), "+d"(c[j][1]), "+d"(c[j][2]), "+d"(c[j][3])
                     : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]),
                       "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(b[3]));
      }
    }
  }
  double s = 0;
#pragma unroll
  for (int j = 0; j < NACC; ++j) s += c[j][0] + c[j][1] + c[j][2] + c[j][3];
  if (s == 12345.678) out[0] = s;
}

template <int NACC>
__global__ void __launch_bounds__(256) k_dfma(double* out, int iters, double seed) {
  double c[NACC];
  double a = seed + threadIdx.x * 1e-9, b = seed * 0.5;
#pragma unroll
  for (int j = 0; j < NACC; ++j) c[j] = j;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int j = 0; j < NACC; ++j) c[j] = fma(a, b, c[j]);
  }
  double s = 0;
#pragma unroll
  for (int j = 0; j < NACC; ++j) s += c[j];
  if (s == 12345.678) out[0] = s;
}

template <typename F>
static double time_ms(F launch) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int i = 0; i < 3; ++i) launch();
  cudaEventRecord(e0);
  for (int i = 0; i < 5; ++i) launch();
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  return ms / 5.0;
}

int main() {
  double* out; cudaMalloc(&out, 8);
  int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const int iters = 20000;
  for (int wpb : {4, 8, 16}) {
    int grid = sms * (wpb == 16 ? 2 : 4), block = wpb * 32 > 256 ? 256 : wpb * 32;
    double warps = double(grid) * block / 32;
#define RUN(SHAPE, NACC, FLOP)                                                                      \
    {                                                                                              \
      double ms = time_ms([&] { k_dmma<SHAPE, NACC><<<grid, block>>>(out, iters, 1.0); });          \
      printf("dmma shape=%d nacc=%d grid=%d block=%d: %.3f ms  %.2f TFLOP/s\n", SHAPE, NACC, grid,  \
             block, ms, warps * iters * NACC * (FLOP) / ms * 1e-9);                                 \
    }
    RUN(884, 4, 512.0) RUN(884, 8, 512.0)
    RUN(1684, 4, 1024.0) RUN(1684, 8, 1024.0)
    RUN(1688, 4, 2048.0) RUN(1688, 8, 2048.0)
    RUN(16816, 4, 4096.0) RUN(16816, 8, 4096.0)
    {
      double ms = time_ms([&] { k_dfma<16><<<grid, block>>>(out, iters, 1.0); });
      printf("dfma nacc=16 grid=%d block=%d: %.3f ms  %.2f TFLOP/s\n", grid, block, ms,
             warps * 32 * iters * 16 * 2.0 / ms * 1e-9);
    }
  }
  cudaError_t e = cudaDeviceSynchronize();
  printf("status: %s\n", cudaGetErrorString(e));
  return e != cudaSuccess;
}
