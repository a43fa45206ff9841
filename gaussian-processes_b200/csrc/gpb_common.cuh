// gpb_common.cuh -- device-side primitives shared by the sm_100a kernels of libgpb200.
// mbarrier / bulk-copy (TMA, SASS UBLKCP) / FP64 tensor-core (DMMA.8x8x4) wrappers and the
// acquire/release helpers of the tile scheduler.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/gpb200.h"

namespace gpb {

constexpr int TILE = 64;         // tile edge of the factorisation (block-column width)
constexpr int LDT = TILE + 4;    // padded smem leading dimension: (LDT mod 16) == 4 -> conflict-free
constexpr int MAXT = GPB_MAX_TERMS;

// Hyper-parameters of one proposal, as the kernels read them (device array [batch]).
struct Hyper {
  int nterms;
  int nbeta;
  int kind[MAXT];
  double p0[MAXT];
  double p1[MAXT];
  double beta[3];
};

// Geometry of one (batched) tall lower-trapezoidal system:
//   rows  0 .. npad-1           training points (identity-padded to a multiple of TILE)
//   rows  npad .. npad+apad-1   appended rows: m test points (K* rows), `design` explicit rows, then one rhs row (y - m(x))
//                               (identity mode: m == 0 and the `design` = npad rows are I, solved to L^-T)
// stored column-major, leading dimension ld = npad + apad, npad columns, one matrix per proposal.
struct Geom {
  int n, dim, m;        // training size, input dimension, number of test points
  int design;           // explicit appended rows (design-matrix columns) between the test rows and the rhs row
  int npad, apad, ld;   // padded sizes
  int nt, rt;           // tile columns (npad/TILE), tile rows (ld/TILE)
  int rhs_row;          // absolute row index of the rhs row (npad + m)
  int batch;
  int tgroup;           // proposals per ticket group of the factorisation kernel (<= batch)
  int ident0;           // first tile row of an appended identity block (npad rows -> L^-T), or -1
  int light_row;        // last tile row when at most 8 of its rows are in use (the rhs row, design rows), or -1:
                        // its tasks run one 8-row DMMA fragment instead of the 64-row tile (k_chol_tasks)
  size_t mat_stride;    // doubles between consecutive matrices
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// ---- mbarrier ---------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// Non-blocking probe (no hardware suspend): used where a spinning thread has other duties.
__device__ __forceinline__ bool mbar_test_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// try_wait with a suspend-time hint (ns): the warp sleeps in hardware for at most about that long.
__device__ __forceinline__ bool mbar_try_wait_ns(uint64_t* bar, uint32_t parity, uint32_t ns) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity), "r"(ns)
      : "memory");
  return ok != 0;
}
// Bounded wait: returns false if the watchdog budget (in try_wait rounds) ran out.
__device__ __forceinline__ bool mbar_wait(uint64_t* bar, uint32_t parity) {
  for (uint32_t it = 0; it < (1u << 24); ++it)
    if (mbar_try_wait(bar, parity)) return true;
  return false;
}

// ---- bulk async copy global -> shared (TMA 1-D; completes on an mbarrier) ----------------------
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes,
                                         uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
          "r"(smem_u32(dst_smem)),
      "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
// 3-D tiled TMA load (SASS UTMALDG): box {rows, cols, 1} of the batched column-major matrices.
__device__ __forceinline__ void tma_load_3d(void* dst_smem, const void* tmap, int c0, int c1, int c2,
                                            uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes "
      "[%0], [%1, {%2, %3, %4}], [%5];" ::"r"(smem_u32(dst_smem)),
      "l"(tmap), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(bar))
      : "memory");
}
// 2-D tiled TMA load: box {cols, rows} of the padded L_jj^-T tiles.
__device__ __forceinline__ void tma_load_2d(void* dst_smem, const void* tmap, int c0, int c1,
                                            uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes "
      "[%0], [%1, {%2, %3}], [%4];" ::"r"(smem_u32(dst_smem)),
      "l"(tmap), "r"(c0), "r"(c1), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const void* tmap) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(tmap) : "memory");
}
// Order generic-proxy accesses against async-proxy (bulk copy) accesses.
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async;" ::: "memory");
}

// ---- scheduler flags ---------------------------------------------------------------------------
__device__ __forceinline__ int ld_acquire(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release(int* p, int v) {
  asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ double ld_cg(const double* p) {
  double v;
  asm volatile("ld.global.cg.f64 %0, [%1];" : "=d"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ double2 ld_cg2(const double* p) {
  double2 v;
  asm volatile("ld.global.cg.v2.f64 {%0,%1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p));
  return v;
}

// ---- FP64 tensor core: D(8x8) += A(8x4) * B(4x8), SASS DMMA.8x8x4 ------------------------------
// lane = 4*g + t:  a = A[row g][k t],  b = B[k t][col g],  c0/c1 = C[row g][col 2t], [col 2t+1]
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
#ifndef GPB_DMMA_ASM
#define GPB_DMMA_ASM asm volatile
#endif
  GPB_DMMA_ASM("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}

__device__ __forceinline__ void bar_named(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
// Named barrier that also ANDs a predicate over the participating threads (bar.red.and).
__device__ __forceinline__ bool bar_named_and(int id, int nthreads, bool pred) {
  uint32_t r;
  asm volatile(
      "{\n\t.reg .pred p, q;\n\t"
      "setp.ne.u32 q, %3, 0;\n\t"
      "bar.red.and.pred p, %1, %2, q;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(r)
      : "r"(id), "r"(nthreads), "r"((uint32_t)pred)
      : "memory");
  return r != 0;
}

}  // namespace gpb
