// gpb_trsm.cu -- the matrix-level triangular solves, on k_chol_tasks in solve-only mode.
//
// Replaces (reference file:line under core/src/main/scala/gp/):
//   Predict.forwardSolve               Predict.scala:40-45   (dtrsv "L","N","N")       -> gpb_trsm_lower
//   Mcmc.backSolve (vector / matrix)   Mcmc.scala:35-41, 52-58 (dtrsv "U" / dtrsm "L","U") -> gpb_trsm_upper
#include "gpb_host.h"

using namespace gpb;
using namespace gpbh;

namespace {

// ---- matrix-level solve packing -----------------------------------------------------------------
// top square <- T (n x n, column-major, ld n), optionally index-reversed and transposed so that an
// upper-triangular U becomes the lower-triangular P U^T... see gpb_trsm_upper; appended rows <- B^T.
__global__ void k_pack_tri(Geom g, const double* __restrict__ T, int reverse, double* __restrict__ M) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x, c = blockIdx.y;
  if (r >= g.npad) return;
  double v = 0.0;
  if (r < g.n && c < g.n) {
    if (r >= c) {
      // lower solve: L(r,c).  upper solve: L'(r,c) = U(n-1-r, n-1-c) (P U P is lower triangular)
      v = reverse ? T[(size_t)(g.n - 1 - c) * g.n + (g.n - 1 - r)] : T[(size_t)c * g.n + r];
    }
  } else if (r == c) {
    v = 1.0;
  }
  M[(size_t)c * g.ld + r] = v;
}
__global__ void k_pack_rhs(Geom g, const double* __restrict__ B, int nrhs, int reverse,
                           double* __restrict__ M) {
  // appended row a (= rhs a), column c  <-  B(c or n-1-c, a)
  const int a = blockIdx.x * blockDim.x + threadIdx.x, c = blockIdx.y;
  if (a >= g.apad) return;
  double v = 0.0;
  if (a < nrhs && c < g.n) v = B[(size_t)a * g.n + (reverse ? g.n - 1 - c : c)];
  M[(size_t)c * g.ld + g.npad + a] = v;
}
__global__ void k_unpack_rhs(Geom g, const double* __restrict__ M, int nrhs, int reverse,
                             double* __restrict__ B) {
  const int a = blockIdx.x * blockDim.x + threadIdx.x, c = blockIdx.y;
  if (a >= nrhs || c >= g.n) return;
  B[(size_t)a * g.n + (reverse ? g.n - 1 - c : c)] = M[(size_t)c * g.ld + g.npad + a];
}

int trsm_common(gpb_ctx* ctx, int n, int nrhs, const double* T, double* B, int reverse) {
  if (!ctx || !T || !B || n < 1 || nrhs < 1) return GPB_E_INVALID;
  CK(cudaSetDevice(ctx->device));
  Geom g = make_geom(n, 1, nrhs, false, 1);
  int rc = ensure_workspace(ctx, g);
  if (rc) return rc;
  const size_t nT = (size_t)n * n, nB = (size_t)n * nrhs;
  if ((rc = grow(ctx, &ctx->aux, &ctx->aux_doubles, nT + nB))) return rc;
  double* dT = ctx->aux;
  double* dB = ctx->aux + nT;
  CK(cudaMemcpyAsync(dT, T, sizeof(double) * nT, cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaMemcpyAsync(dB, B, sizeof(double) * nB, cudaMemcpyHostToDevice, ctx->stream));
  k_pack_tri<<<dim3((g.npad + 127) / 128, g.npad), 128, 0, ctx->stream>>>(g, dT, reverse, ctx->mats);
  k_pack_rhs<<<dim3((g.apad + 63) / 64, g.npad), 64, 0, ctx->stream>>>(g, dB, nrhs, reverse, ctx->mats);
  CK(cudaEventRecord(ctx->ev[0], ctx->stream));
  CK(cudaEventRecord(ctx->ev[1], ctx->stream));
  ctx->launches += 2;
  if ((rc = do_chol(ctx, g, /*solve_only=*/1))) return rc;
  CK(cudaEventRecord(ctx->ev[2], ctx->stream));
  k_unpack_rhs<<<dim3((nrhs + 63) / 64, n), 64, 0, ctx->stream>>>(g, ctx->mats, nrhs, reverse, dB);
  ctx->launches += 1;
  CK(cudaMemcpyAsync(B, dB, sizeof(double) * nB, cudaMemcpyDeviceToHost, ctx->stream));
  return finish_timing(ctx);
}

}  // namespace

extern "C" {

int gpb_trsm_lower(gpb_ctx* ctx, int n, int nrhs, const double* L, double* B) {
  return trsm_common(ctx, n, nrhs, L, B, 0);
}

// U x = b with U upper triangular: with P the index reversal, (P U P) is lower triangular and
// (P U P)(P x) = P b -- the same forward substitution on reversed indices.
int gpb_trsm_upper(gpb_ctx* ctx, int n, int nrhs, const double* U, double* B) {
  return trsm_common(ctx, n, nrhs, U, B, 1);
}

}  // extern "C"
