// gpb_api.cu -- the C ABI of libgpb200.so (include/gpb200.h): contexts, datasets, workspace and
// the host-side sequencing of the sm_100a kernels.  No CPU fallback anywhere: every entry point
// either runs the CUDA path or returns an error status.
#include <cstdlib>

#include "gpb_host.h"

using namespace gpb;
using namespace gpbh;

namespace gpbh {
int cuda_fail(gpb_ctx* ctx, cudaError_t e, const char* where) {
  if (ctx) ctx->cuda_err = std::string(where) + ": " + cudaGetErrorString(e);
  return (e == cudaErrorMemoryAllocation) ? GPB_E_NOMEM : GPB_E_CUDA;
}
inline int round_up(int v, int m) { return (v + m - 1) / m * m; }

// m test points + (with_rhs ? 1 : 0) rhs row appended below the npad x npad square
Geom make_geom(int n, int dim, int m, bool with_rhs, int batch, int design, bool identity) {
  Geom g;
  g.npad = round_up(n, TILE);
  if (identity) { m = 0; design = g.npad; }
  g.n = n; g.dim = dim; g.m = m; g.design = design;
  g.apad = round_up(m + design + (with_rhs ? 1 : 0), TILE);
  g.ld = g.npad + g.apad;
  g.nt = g.npad / TILE;
  g.rt = g.ld / TILE;
  g.rhs_row = with_rhs ? g.npad + m + design : -1;
  g.batch = batch;
  g.tgroup = batch;
  g.ident0 = identity ? g.nt : -1;
  const int arows = m + design + (with_rhs ? 1 : 0);   // appended rows in use
  g.light_row = (arows > 0 && arows - (g.apad - TILE) <= 8) ? g.rt - 1 : -1;
  // GPB_LIGHT_ROWS=0: run that tile row like every other one (the results must not change: tests)
  static const bool light_on = [] { const char* e = getenv("GPB_LIGHT_ROWS"); return !(e && e[0] == '0'); }();
  if (!light_on) g.light_row = -1;
  g.mat_stride = (size_t)g.ld * g.npad;
  return g;
}

int ensure_workspace(gpb_ctx* ctx, const Geom& g) {
  int rc;
  if ((rc = grow(ctx, &ctx->mats, &ctx->mats_doubles, g.mat_stride * g.batch))) return rc;
  if ((rc = grow(ctx, &ctx->linvt, &ctx->linvt_doubles, chol_linvt_doubles(g)))) return rc;
  if ((rc = grow(ctx, &ctx->done, &ctx->done_ints, chol_done_ints(g)))) return rc;
  if ((rc = grow(ctx, &ctx->info, &ctx->info_ints, (size_t)g.batch))) return rc;
  if ((rc = grow(ctx, &ctx->hyper, &ctx->hyper_n, (size_t)g.batch))) return rc;
  if ((rc = grow(ctx, &ctx->ll, &ctx->ll_n, (size_t)g.batch))) return rc;
  if ((rc = grow_pinned(ctx, &ctx->h_hyper, &ctx->h_hyper_n, (size_t)g.batch))) return rc;
  if ((rc = grow_pinned(ctx, &ctx->h_ll, &ctx->h_ll_n, (size_t)g.batch))) return rc;
  return GPB_OK;
}

int fill_hyper(Hyper* h, const gpb_term* terms, int nterms, const double* beta, int nbeta, int dim) {
  if (nterms < 1 || nterms > GPB_MAX_TERMS || !terms) return GPB_E_INVALID;
  if (!(nbeta == 0 || nbeta == dim + 1)) return GPB_E_INVALID;
  if (nbeta > 0 && !beta) return GPB_E_INVALID;
  std::memset(h, 0, sizeof(Hyper));
  h->nterms = nterms;
  h->nbeta = nbeta;
  for (int t = 0; t < nterms; ++t) {
    if (terms[t].kind == GPB_KIND_MATERN) return GPB_E_UNSUPPORTED;
    if (terms[t].kind != GPB_KIND_SE && terms[t].kind != GPB_KIND_WHITE) return GPB_E_INVALID;
    h->kind[t] = terms[t].kind;
    h->p0[t] = terms[t].p0;
    h->p1[t] = terms[t].p1;
  }
  for (int i = 0; i < nbeta; ++i) h->beta[i] = beta[i];
  return GPB_OK;
}

// build -> factor/solve on ctx->stream, with phase events
int run_system(gpb_ctx* ctx, const Geom& g, const double* d_x, const double* d_y,
               const double* d_xstar, double nugget) {
  CK(cudaEventRecord(ctx->ev[0], ctx->stream));
  launch_build(ctx->stream, g, d_x, d_y, d_xstar, ctx->hyper, nugget, ctx->mats, BUILD_FAST, 0);
  ctx->launches += 1;
  CK(cudaEventRecord(ctx->ev[1], ctx->stream));
  int rc = do_chol(ctx, g, 0);
  if (rc) return rc;
  CK(cudaEventRecord(ctx->ev[2], ctx->stream));
  return GPB_OK;
}

// Proposals per ticket group: enough tasks in flight to fill the machine, few enough that the
// group's lower triangles (4 n^2 bytes each) stay L2-resident.  GPB_TICKET_GROUP overrides (tuning).
static int ticket_group(const gpb_ctx* ctx, const Geom& g) {
  static int env = -1;
  if (env < 0) { const char* e = getenv("GPB_TICKET_GROUP"); env = e ? atoi(e) : 0; }
  if (env > 0) return env < g.batch ? env : g.batch;
  return g.batch;
}

int do_chol(gpb_ctx* ctx, const Geom& g, int solve_only, int skip_reset) {
  unsigned long long* trace = nullptr;
  if (ctx->trace_on) {
    ctx->trace_tasks = chol_task_count(g);
    int rc = grow(ctx, &ctx->trace, &ctx->trace_n, (size_t)ctx->trace_tasks * CHOL_TRACE_SLOTS);
    if (rc) return rc;
    cudaMemsetAsync(ctx->trace, 0, sizeof(unsigned long long) * ctx->trace_tasks * CHOL_TRACE_SLOTS, ctx->stream);
    trace = ctx->trace;
  }
  Geom gg = g;
  gg.tgroup = ticket_group(ctx, g);
  const int nl = launch_chol(ctx->stream, gg, ctx->mats, ctx->linvt, ctx->done, ctx->info, ctx->sched,
                             ctx->sched_mode, ctx->num_sms, solve_only, trace, -1, skip_reset);
  if (nl == -2) return GPB_E_UNSUPPORTED;   // more than 2^31 tiles in one launch
  if (nl < 0) {
    ctx->cuda_err = "cuTensorMapEncodeTiled failed (TMA descriptor of the workspace)";
    return GPB_E_CUDA;
  }
  ctx->launches += nl;
  return GPB_OK;
}

int finish_timing(gpb_ctx* ctx) {
  CK(cudaEventRecord(ctx->ev[3], ctx->stream));
  CK(cudaMemcpyAsync(ctx->h_sched, ctx->sched, SCHED_HEAD_BYTES, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  CK(cudaGetLastError());
  for (int i = 0; i < 3; ++i) cudaEventElapsedTime(&ctx->timing[i], ctx->ev[i], ctx->ev[i + 1]);
  const int sched_error = ctx->h_sched->error;
  if (sched_error != 0) {
    ctx->cuda_err = "scheduler watchdog code " + std::to_string(sched_error);
    cudaMemset(ctx->sched, 0, SCHED_HEAD_BYTES);
    return GPB_E_WATCHDOG;
  }
  return GPB_OK;
}

}  // namespace gpbh

extern "C" {

const char* gpb_version(void) { return "gpb200 0.1 (sm_100a)"; }

const char* gpb_strerror(int status) {
  switch (status) {
    case GPB_OK: return "ok";
    case GPB_E_INVALID: return "invalid argument";
    case GPB_E_NODEVICE: return "no usable CUDA device (sm_100 required; there is no CPU fallback)";
    case GPB_E_CUDA: return "CUDA runtime error";
    case GPB_E_NOMEM: return "out of memory";
    case GPB_E_UNSUPPORTED: return "unsupported (Matern kernel / dim > 2 are out of scope)";
    case GPB_E_WATCHDOG: return "scheduler watchdog fired";
    case GPB_E_NOTPD: return "covariance matrix not positive definite";
    case GPB_E_NOCOMM: return "no communicator (NCCL not loadable, or gpb_comm_init_* not called on this context)";
    default: return "unknown status";
  }
}

const char* gpb_last_cuda_error(const gpb_ctx* ctx) { return ctx ? ctx->cuda_err.c_str() : ""; }
long long gpb_launch_count(const gpb_ctx* ctx) { return ctx ? ctx->launches : 0; }
int gpb_last_timing(const gpb_ctx* ctx, float* t3) {
  if (!ctx || !t3) return GPB_E_INVALID;
  for (int i = 0; i < 3; ++i) t3[i] = ctx->timing[i];
  return GPB_OK;
}
int gpb_set_trace(gpb_ctx* ctx, int enable) {
  if (!ctx) return GPB_E_INVALID;
  ctx->trace_on = enable ? 1 : 0;
  return GPB_OK;
}
long long gpb_get_trace(gpb_ctx* ctx, unsigned long long* out, long long max_tasks) {
  if (!ctx || !ctx->trace) return 0;
  long long nt = ctx->trace_tasks < max_tasks ? ctx->trace_tasks : max_tasks;
  if (out && nt > 0)
    cudaMemcpy(out, ctx->trace, sizeof(unsigned long long) * nt * CHOL_TRACE_SLOTS, cudaMemcpyDeviceToHost);
  return ctx->trace_tasks;
}
long long gpb_schedule_task(int n, int appended, int batch, long long ticket, int* out4) {
  if (n < 1 || appended < 0 || batch < 1) return GPB_E_INVALID;
  const Geom g = make_geom(n, 1, appended, false, batch);
  return chol_schedule_task(g, ticket, out4);
}
int gpb_set_sched(gpb_ctx* ctx, int mode) {
  if (!ctx || mode < 0 || mode > 1) return GPB_E_INVALID;
  ctx->sched_mode = mode;
  return GPB_OK;
}

int gpb_ctx_create(int device, gpb_ctx** out) {
  if (!out) return GPB_E_INVALID;
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return GPB_E_NODEVICE; }
  if (device < 0 || device >= ndev) return GPB_E_INVALID;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return GPB_E_NODEVICE;
  if (prop.major != 10) return GPB_E_NODEVICE;  // kernels are compiled for sm_100a only
  gpb_ctx* ctx = new (std::nothrow) gpb_ctx();
  if (!ctx) return GPB_E_NOMEM;
  ctx->device = device;
  ctx->num_sms = prop.multiProcessorCount;
  cudaError_t e = cudaSetDevice(device);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
  if (e == cudaSuccess) {
    // the flush / gather stream of the MH loop: highest priority, so that its (tiny) copies and the
    // NCCL kernel are scheduled in the gaps between the iterations' kernels
    int lo = 0, hi = 0;
    cudaDeviceGetStreamPriorityRange(&lo, &hi);
    e = cudaStreamCreateWithPriority(&ctx->copy_stream, cudaStreamNonBlocking, hi);
  }
  for (int i = 0; i < 4 && e == cudaSuccess; ++i) e = cudaEventCreate(&ctx->ev[i]);
  for (int i = 0; i < 2 && e == cudaSuccess; ++i) e = cudaEventCreateWithFlags(&ctx->ev_ready[i], cudaEventDisableTiming);
  for (int i = 0; i < 2 && e == cudaSuccess; ++i) e = cudaEventCreateWithFlags(&ctx->ev_flushed[i], cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaMalloc((void**)&ctx->sched, sizeof(SchedState));
  if (e == cudaSuccess) e = cudaMemset(ctx->sched, 0, sizeof(SchedState));
  if (e == cudaSuccess) e = cudaMallocHost((void**)&ctx->h_sched, sizeof(SchedState));
  if (e != cudaSuccess) { gpb_ctx_destroy(ctx); return GPB_E_CUDA; }
  *out = ctx;
  return GPB_OK;
}

void gpb_ctx_destroy(gpb_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  if (ctx->stream) cudaStreamSynchronize(ctx->stream);
  if (ctx->copy_stream) cudaStreamSynchronize(ctx->copy_stream);
  comm_release(ctx);
  cudaFree(ctx->mh); cudaFree(ctx->ring); cudaFreeHost(ctx->h_stage);
  for (int i = 0; i < 2; ++i) { if (ctx->ev_ready[i]) cudaEventDestroy(ctx->ev_ready[i]); if (ctx->ev_flushed[i]) cudaEventDestroy(ctx->ev_flushed[i]); }
  if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
  cudaFree(ctx->mats); cudaFree(ctx->linvt); cudaFree(ctx->done); cudaFree(ctx->info);
  cudaFree(ctx->trace); cudaFree(ctx->hyper); cudaFree(ctx->ll); cudaFree(ctx->aux); cudaFree(ctx->sched);
  cudaFreeHost(ctx->h_hyper); cudaFreeHost(ctx->h_ll); cudaFreeHost(ctx->h_sched);
  for (int i = 0; i < 4; ++i) if (ctx->ev[i]) cudaEventDestroy(ctx->ev[i]);
  if (ctx->stream) cudaStreamDestroy(ctx->stream);
  delete ctx;
}

int gpb_dataset_create(gpb_ctx* ctx, int n, int dim, const double* x, const double* y,
                       gpb_dataset** out) {
  if (!ctx || !out || !x || n < 1) return GPB_E_INVALID;
  if (dim != 1 && dim != 2) return GPB_E_UNSUPPORTED;
  *out = nullptr;
  CK(cudaSetDevice(ctx->device));
  gpb_dataset* ds = new (std::nothrow) gpb_dataset();
  if (!ds) return GPB_E_NOMEM;
  ds->ctx = ctx; ds->n = n; ds->dim = dim; ds->x = nullptr; ds->y = nullptr;
  cudaError_t e = cudaMalloc((void**)&ds->x, sizeof(double) * (size_t)n * dim);
  if (e == cudaSuccess) e = cudaMalloc((void**)&ds->y, sizeof(double) * (size_t)n);
  if (e == cudaSuccess) e = cudaMemcpy(ds->x, x, sizeof(double) * (size_t)n * dim, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) {
    if (y) e = cudaMemcpy(ds->y, y, sizeof(double) * (size_t)n, cudaMemcpyHostToDevice);
    else e = cudaMemset(ds->y, 0, sizeof(double) * (size_t)n);
  }
  if (e != cudaSuccess) { gpb_dataset_destroy(ds); return cuda_fail(ctx, e, "gpb_dataset_create"); }
  *out = ds;
  return GPB_OK;
}

void gpb_dataset_destroy(gpb_dataset* ds) {
  if (!ds) return;
  cudaFree(ds->x); cudaFree(ds->y);
  delete ds;
}

double gpb_ll_value(const gpb_ll* r, int formula) {
  if (!r || r->info != 0) return NAN;
  const double C = r->n * 0.5 * std::log(2 * M_PI);  // GaussianProcess.scala:126
  if (formula == GPB_FORMULA_GPML) return -0.5 * r->Q - r->S - C;
  return -0.5 * r->Q + 0.5 * (r->S + C) * r->U1;     // as parsed, SURVEY.md 0.1
}

int gpb_loglik_batch(gpb_ctx* ctx, const gpb_dataset* ds, int batch, const gpb_term* terms,
                     int nterms, const double* beta, int nbeta, double nugget, gpb_ll* out) {
  if (!ctx || !ds || ds->ctx != ctx || batch < 1 || !out) return GPB_E_INVALID;
  CK(cudaSetDevice(ctx->device));
  const Geom g = make_geom(ds->n, ds->dim, 0, true, batch);
  int rc = ensure_workspace(ctx, g);
  if (rc) return rc;
  for (int b = 0; b < batch; ++b) {
    rc = fill_hyper(&ctx->h_hyper[b], terms + (size_t)b * nterms, nterms,
                    nbeta ? beta + (size_t)b * nbeta : nullptr, nbeta, ds->dim);
    if (rc) return rc;
  }
  CK(cudaMemcpyAsync(ctx->hyper, ctx->h_hyper, sizeof(Hyper) * batch, cudaMemcpyHostToDevice, ctx->stream));
  rc = run_system(ctx, g, ds->x, ds->y, nullptr, nugget);
  if (rc) return rc;
  launch_ll_epilogue(ctx->stream, g, ctx->mats, ctx->info, ctx->ll);
  ctx->launches += 1;
  CK(cudaMemcpyAsync(ctx->h_ll, ctx->ll, sizeof(gpb_ll) * batch, cudaMemcpyDeviceToHost, ctx->stream));
  rc = finish_timing(ctx);
  if (rc) return rc;
  std::memcpy(out, ctx->h_ll, sizeof(gpb_ll) * batch);
  return GPB_OK;
}

int gpb_loglik(gpb_ctx* ctx, const gpb_dataset* ds, const gpb_term* terms, int nterms,
               const double* beta, int nbeta, double nugget, gpb_ll* out) {
  return gpb_loglik_batch(ctx, ds, 1, terms, nterms, beta, nbeta, nugget, out);
}

int gpb_predict(gpb_ctx* ctx, const gpb_dataset* ds, const gpb_term* terms, int nterms,
                double nugget, int m, const double* xstar, double* mean, double* var, int* info) {
  if (!ctx || !ds || ds->ctx != ctx || m < 1 || !xstar || !mean || !var) return GPB_E_INVALID;
  CK(cudaSetDevice(ctx->device));
  const Geom g = make_geom(ds->n, ds->dim, m, true, 1);
  int rc = ensure_workspace(ctx, g);
  if (rc) return rc;
  // Predict.scala:80 uses the raw y: no mean function here (nbeta = 0)
  rc = fill_hyper(&ctx->h_hyper[0], terms, nterms, nullptr, 0, ds->dim);
  if (rc) return rc;
  const size_t nx = (size_t)m * ds->dim;
  const size_t npart = 2 * (size_t)pred_slices(g) * m;
  if ((rc = grow(ctx, &ctx->aux, &ctx->aux_doubles, nx + 2 * (size_t)m + npart))) return rc;
  double* d_xstar = ctx->aux;
  double* d_mean = ctx->aux + nx;
  double* d_var = d_mean + m;
  double* d_part = d_var + m;
  CK(cudaMemcpyAsync(ctx->hyper, ctx->h_hyper, sizeof(Hyper), cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaMemcpyAsync(d_xstar, xstar, sizeof(double) * nx, cudaMemcpyHostToDevice, ctx->stream));
  rc = run_system(ctx, g, ds->x, ds->y, d_xstar, nugget);
  if (rc) return rc;
  launch_pred_epilogue(ctx->stream, g, ctx->mats, ctx->hyper, d_part, d_mean, d_var);
  ctx->launches += 2;
  CK(cudaMemcpyAsync(mean, d_mean, sizeof(double) * m, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaMemcpyAsync(var, d_var, sizeof(double) * m, cudaMemcpyDeviceToHost, ctx->stream));
  int h_info = 0;
  CK(cudaMemcpyAsync(&h_info, ctx->info, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  rc = finish_timing(ctx);
  if (rc) return rc;
  if (info) *info = h_info;
  return GPB_OK;
}


static int build_to_host(gpb_ctx* ctx, int n, int dim, const double* x, const gpb_term* terms,
                         int nterms, double nugget, double* out, int mode) {
  if (!ctx || !x || !out || n < 1) return GPB_E_INVALID;
  if (dim != 1 && dim != 2) return GPB_E_UNSUPPORTED;
  CK(cudaSetDevice(ctx->device));
  const Geom g = make_geom(n, dim, 0, false, 1);
  int rc = ensure_workspace(ctx, g);
  if (rc) return rc;
  if (mode != BUILD_DISTANCE) {
    rc = fill_hyper(&ctx->h_hyper[0], terms, nterms, nullptr, 0, dim);
    if (rc) return rc;
  } else {
    std::memset(&ctx->h_hyper[0], 0, sizeof(Hyper));
  }
  if ((rc = grow(ctx, &ctx->aux, &ctx->aux_doubles, (size_t)n * dim))) return rc;
  CK(cudaMemcpyAsync(ctx->aux, x, sizeof(double) * (size_t)n * dim, cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaMemcpyAsync(ctx->hyper, ctx->h_hyper, sizeof(Hyper), cudaMemcpyHostToDevice, ctx->stream));
  if (mode == BUILD_FAST)   // the lower trapezoid only, as the factorisation reads it; the rest reads as zero
    CK(cudaMemsetAsync(ctx->mats, 0, sizeof(double) * g.mat_stride, ctx->stream));
  launch_build(ctx->stream, g, ctx->aux, nullptr, nullptr, ctx->hyper, nugget, ctx->mats, mode, mode != BUILD_FAST);
  ctx->launches += 1;
  CK(cudaMemcpy2DAsync(out, sizeof(double) * n, ctx->mats, sizeof(double) * g.ld, sizeof(double) * n,
                       n, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  CK(cudaGetLastError());
  return GPB_OK;
}

int gpb_distance_matrix(gpb_ctx* ctx, int n, int dim, const double* x, double* D) {
  return build_to_host(ctx, n, dim, x, nullptr, 0, 0.0, D, 1);
}

int gpb_build_cov(gpb_ctx* ctx, int n, int dim, const double* x, const gpb_term* terms, int nterms,
                  double nugget, double* K) {
  return build_to_host(ctx, n, dim, x, terms, nterms, nugget, K, 0);
}

int gpb_build_cov_hot(gpb_ctx* ctx, int n, int dim, const double* x, const gpb_term* terms, int nterms,
                      double nugget, double* K) {
  int rc = build_to_host(ctx, n, dim, x, terms, nterms, nugget, K, BUILD_FAST);
  if (rc) return rc;
  // diagonal tiles are written whole: clear what lies above the diagonal
  for (int c = 1; c < n; ++c) std::memset(K + (size_t)c * n, 0, sizeof(double) * c);
  return GPB_OK;
}

int gpb_potrf(gpb_ctx* ctx, int n, double* A, int* info) {
  if (!ctx || !A || n < 1) return GPB_E_INVALID;
  CK(cudaSetDevice(ctx->device));
  const Geom g = make_geom(n, 1, 0, false, 1);
  int rc = ensure_workspace(ctx, g);
  if (rc) return rc;
  CK(cudaMemsetAsync(ctx->mats, 0, sizeof(double) * g.mat_stride, ctx->stream));
  CK(cudaMemcpy2DAsync(ctx->mats, sizeof(double) * g.ld, A, sizeof(double) * n, sizeof(double) * n, n,
                       cudaMemcpyHostToDevice, ctx->stream));
  launch_pad_identity(ctx->stream, g, ctx->mats);
  CK(cudaEventRecord(ctx->ev[0], ctx->stream));
  CK(cudaEventRecord(ctx->ev[1], ctx->stream));
  ctx->launches += 1;
  if ((rc = do_chol(ctx, g, 0))) return rc;
  CK(cudaEventRecord(ctx->ev[2], ctx->stream));
  launch_zero_upper(ctx->stream, g, ctx->mats);
  ctx->launches += 1;
  CK(cudaMemcpy2DAsync(A, sizeof(double) * n, ctx->mats, sizeof(double) * g.ld, sizeof(double) * n, n,
                       cudaMemcpyDeviceToHost, ctx->stream));
  int h_info = 0;
  CK(cudaMemcpyAsync(&h_info, ctx->info, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  rc = finish_timing(ctx);
  if (rc) return rc;
  if (info) *info = h_info;
  return GPB_OK;
}

}  // extern "C"
