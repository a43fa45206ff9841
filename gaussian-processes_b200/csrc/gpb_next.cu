// gpb_next.cu -- the "next" rows of the scope table, on the same kernels:
//   MeanParameters.samplePlane  MeanParameters.scala:74-103  -> gpb_plane_stats / gpb_sample_plane
//   GaussianProcess.draw        GaussianProcess.scala:67-83  -> gpb_draw (Cholesky root, see header)
#include "gpb_host.h"
#include "gpb_plane.h"

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
  const int pinfo = plane_draw(d, G, w, prior_mean, prior_var, z, beta);
  if (pinfo != 0 && info) *info = pinfo;
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
