// gpb_next.cu -- the "next" rows of the scope table, on the same kernels:
//   MeanParameters.samplePlane  MeanParameters.scala:74-103  -> gpb_plane_stats / gpb_sample_plane
//   GaussianProcess.draw        GaussianProcess.scala:67-83  -> gpb_draw (Cholesky root, see header)
#include "gpb_host.h"

using namespace gpb;
using namespace gpbh;

extern "C" {

int gpb_plane_stats(gpb_ctx* ctx, const gpb_dataset* ds, const gpb_term* terms, int nterms,
                    double nugget, int d, const double* X, double* G, double* w, int* info) {
  if (!ctx || !ds || ds->ctx != ctx || !X || !G || !w || d < 1 || d > 4) return GPB_E_INVALID;
  CK(cudaSetDevice(ctx->device));
  const Geom g = make_geom(ds->n, ds->dim, 0, true, 1, d);
  int rc = ensure_workspace(ctx, g);
  if (rc) return rc;
  // MeanParameters.scala:91: y = obs.map(_.y), no mean subtraction (nbeta = 0)
  if ((rc = fill_hyper(&ctx->h_hyper[0], terms, nterms, nullptr, 0, ds->dim))) return rc;
  const size_t nx = (size_t)d * ds->n, no = (size_t)d * d + d;
  if ((rc = grow(ctx, &ctx->aux, &ctx->aux_doubles, nx + no))) return rc;
  double* d_rows = ctx->aux;
  double* d_out = ctx->aux + nx;
  CK(cudaMemcpyAsync(ctx->hyper, ctx->h_hyper, sizeof(Hyper), cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaMemcpyAsync(d_rows, X, sizeof(double) * nx, cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaEventRecord(ctx->ev[0], ctx->stream));
  launch_build(ctx->stream, g, ds->x, ds->y, nullptr, ctx->hyper, nugget, ctx->mats, BUILD_FAST, 0, d_rows);
  ctx->launches += 1;
  CK(cudaEventRecord(ctx->ev[1], ctx->stream));
  if ((rc = do_chol(ctx, g, 0))) return rc;
  CK(cudaEventRecord(ctx->ev[2], ctx->stream));
  launch_plane_epilogue(ctx->stream, g, ctx->mats, d_rows, d_out);
  ctx->launches += 1;
  std::vector<double> h(no);
  int h_info = 0;
  CK(cudaMemcpyAsync(h.data(), d_out, sizeof(double) * no, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaMemcpyAsync(&h_info, ctx->info, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  if ((rc = finish_timing(ctx))) return rc;
  std::memcpy(G, h.data(), sizeof(double) * d * d);
  std::memcpy(w, h.data() + (size_t)d * d, sizeof(double) * d);
  if (info) *info = h_info;
  return GPB_OK;
}

// d x d host algebra of the Gibbs step (d = dim + 1 <= 4): Breeze `\` on a square matrix is an LU
// solve with partial pivoting, `cholesky` the lower factor.
static bool lu_solve(int d, const double* A, const double* b, double* x) {
  double M[4][5];
  for (int i = 0; i < d; ++i) { for (int j = 0; j < d; ++j) M[i][j] = A[i * d + j]; M[i][d] = b[i]; }
  for (int c = 0; c < d; ++c) {
    int p = c;
    for (int i = c + 1; i < d; ++i) if (std::fabs(M[i][c]) > std::fabs(M[p][c])) p = i;
    if (M[p][c] == 0.0) return false;
    if (p != c) for (int j = 0; j <= d; ++j) std::swap(M[p][j], M[c][j]);
    for (int i = c + 1; i < d; ++i) {
      const double f = M[i][c] / M[c][c];
      for (int j = c; j <= d; ++j) M[i][j] -= f * M[c][j];
    }
  }
  for (int i = d - 1; i >= 0; --i) {
    double s = M[i][d];
    for (int j = i + 1; j < d; ++j) s -= M[i][j] * x[j];
    x[i] = s / M[i][i];
  }
  return true;
}

int gpb_sample_plane(gpb_ctx* ctx, const gpb_dataset* ds, const gpb_term* terms, int nterms,
                     double nugget, int d, const double* X, double prior_mean, double prior_var,
                     const double* z, double* beta, int* info) {
  if (!z || !beta || !(prior_var > 0.0)) return GPB_E_INVALID;
  double G[16], w[4];
  int h_info = 0;
  int rc = gpb_plane_stats(ctx, ds, terms, nterms, nugget, d, X, G, w, &h_info);
  if (rc) return rc;
  if (info) *info = h_info;
  if (h_info != 0) return GPB_OK;  // kxx not positive definite: the reference throws from cholesky
  double prec[16], rhs[4], mean[4], root[16] = {0};
  for (int i = 0; i < d; ++i) {
    for (int j = 0; j < d; ++j) prec[i * d + j] = G[i * d + j] + (i == j ? 1.0 / prior_var : 0.0);
    rhs[i] = (1.0 / prior_var) * prior_mean + w[i];   // priorPrec * fill(mean) + x.t * u  (:94-95)
  }
  if (!lu_solve(d, prec, rhs, mean)) { if (info) *info = -1; return GPB_OK; }
  for (int j = 0; j < d; ++j) {                       // root = cholesky(prec)  (:97)
    double s = prec[j * d + j];
    for (int k = 0; k < j; ++k) s -= root[j * d + k] * root[j * d + k];
    if (!(s > 0.0)) { if (info) *info = j + 1; return GPB_OK; }
    root[j * d + j] = std::sqrt(s);
    for (int i = j + 1; i < d; ++i) {
      double t = prec[i * d + j];
      for (int k = 0; k < j; ++k) t -= root[i * d + k] * root[j * d + k];
      root[i * d + j] = t / root[j * d + j];
    }
  }
  for (int i = 0; i < d; ++i) {                       // beta = mean + root \ z  (:99)
    double s = z[i];
    for (int k = 0; k < i; ++k) s -= root[i * d + k] * (beta[k] - mean[k]);
    beta[i] = mean[i] + s / root[i * d + i];
  }
  return GPB_OK;
}

int gpb_draw(gpb_ctx* ctx, int n, int dim, const double* x, const gpb_term* terms, int nterms,
             double nugget, const double* mu, const double* z, double* out, int* info) {
  if (!ctx || !x || !z || !out || n < 1) return GPB_E_INVALID;
  if (dim != 1 && dim != 2) return GPB_E_UNSUPPORTED;
  CK(cudaSetDevice(ctx->device));
  const Geom g = make_geom(n, dim, 0, false, 1);
  int rc = ensure_workspace(ctx, g);
  if (rc) return rc;
  if ((rc = fill_hyper(&ctx->h_hyper[0], terms, nterms, nullptr, 0, dim))) return rc;
  const size_t nx = (size_t)n * dim;
  if ((rc = grow(ctx, &ctx->aux, &ctx->aux_doubles, nx + 3 * (size_t)n))) return rc;
  double* d_x = ctx->aux;
  double* d_z = d_x + nx;
  double* d_mu = d_z + n;
  double* d_out = d_mu + n;
  CK(cudaMemcpyAsync(ctx->hyper, ctx->h_hyper, sizeof(Hyper), cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaMemcpyAsync(d_x, x, sizeof(double) * nx, cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaMemcpyAsync(d_z, z, sizeof(double) * n, cudaMemcpyHostToDevice, ctx->stream));
  if (mu) CK(cudaMemcpyAsync(d_mu, mu, sizeof(double) * n, cudaMemcpyHostToDevice, ctx->stream));
  else CK(cudaMemsetAsync(d_mu, 0, sizeof(double) * n, ctx->stream));
  CK(cudaEventRecord(ctx->ev[0], ctx->stream));
  launch_build(ctx->stream, g, d_x, nullptr, nullptr, ctx->hyper, nugget, ctx->mats, BUILD_FAST, 0);
  ctx->launches += 1;
  CK(cudaEventRecord(ctx->ev[1], ctx->stream));
  if ((rc = do_chol(ctx, g, 0))) return rc;
  CK(cudaEventRecord(ctx->ev[2], ctx->stream));
  launch_draw(ctx->stream, g, ctx->mats, d_z, d_mu, d_out);
  ctx->launches += 1;
  int h_info = 0;
  CK(cudaMemcpyAsync(out, d_out, sizeof(double) * n, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaMemcpyAsync(&h_info, ctx->info, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  if ((rc = finish_timing(ctx))) return rc;
  if (info) *info = h_info;
  return GPB_OK;
}

}  // extern "C"
