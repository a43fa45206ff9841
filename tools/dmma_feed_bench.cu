// Microbenchmark: how close to the DMMA issue peak does ONE SM get with w warps per scheduler when
// the operands come (a) from registers, (b) from shared memory with the factorisation kernel's
// fragment pattern (8 LDS.64 + 16 DMMA per 4-wide k step)?  Not part of the product path.
#include <cstdio>
#include <cuda_runtime.h>
constexpr int LDT = 68;
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
template <int MODE>
__global__ void __launch_bounds__(256) k_feed(double* out, int iters) {
  extern __shared__ double sm[];   // [16 k][LDT] x 2 operands, replicated per 4-warp group
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, gq = lane >> 2, tq = lane & 3;
  const int grp = warp >> 2, w4 = warp & 3, wc = w4 >> 1, wr = w4 & 1;
  double* PA = sm + grp * (2 * 16 * LDT);
  double* PB = PA + 16 * LDT;
  for (int e = tid; e < 2 * 2 * 16 * LDT; e += blockDim.x) sm[e] = 1e-3 * (e % 97);
  __syncthreads();
  double acc[4][4][2];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
  double ra[4], rb[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) { ra[i] = 1.0 + i + lane * 1e-3; rb[i] = 0.5 + i; }
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) {
      double a[4], bb[4];
      if (MODE == 0) {
#pragma unroll
        for (int i = 0; i < 4; ++i) { a[i] = ra[i]; bb[i] = rb[i]; }
      } else {
        const int k = ks * 4 + tq;
#pragma unroll
        for (int mt = 0; mt < 4; ++mt) a[mt] = -PA[k * LDT + wc * 32 + mt * 8 + gq];
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) bb[nt] = PB[k * LDT + wr * 32 + nt * 8 + gq];
      }
#pragma unroll
      for (int mt = 0; mt < 4; ++mt)
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) dmma884(acc[mt][nt][0], acc[mt][nt][1], a[mt], bb[nt]);
    }
    if (MODE == 2) __syncwarp();
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) s += acc[i][j][0] + acc[i][j][1];
  if (s == 12345.678) out[0] = s;
}
template <int MODE>
static void run(int warps, int ctas_per_sm, int sms, double* out) {
  const int iters = 4000, block = warps * 32, grid = sms * ctas_per_sm;
  const size_t smem = 2 * 2 * 16 * LDT * sizeof(double);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int i = 0; i < 2; ++i) k_feed<MODE><<<grid, block, smem>>>(out, iters);
  cudaEventRecord(e0);
  for (int i = 0; i < 5; ++i) k_feed<MODE><<<grid, block, smem>>>(out, iters);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= 5;
  const double flop = double(grid) * warps * iters * 4 * 16 * 512.0;
  printf("mode=%d warps/CTA=%d CTAs/SM=%d: %.3f ms  %.2f TFLOP/s\n", MODE, warps, ctas_per_sm, ms, flop / ms * 1e-9);
}
int main() {
  double* out; cudaMalloc(&out, 8);
  int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  for (int w : {4, 8}) for (int c : {1, 2}) { if (w == 8 && c == 2) { run<0>(8, 2, sms, out); run<1>(8, 2, sms, out); continue; }
    run<0>(w, c, sms, out); run<1>(w, c, sms, out); run<2>(w, c, sms, out); }
  printf("status: %s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
