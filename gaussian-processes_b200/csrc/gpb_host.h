// gpb_host.h -- host-side state behind the opaque handles of include/gpb200.h, shared by the
// translation units that implement the C ABI.
#pragma once
#include <cmath>
#include <cstdio>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "gpb_internal.h"

using gpb::Hyper;
using gpb::SchedState;

struct gpb_ctx {
  int device = 0;
  int num_sms = 0;
  int sched_mode = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
  // workspace (grown on demand, never shrunk)
  double* mats = nullptr;   size_t mats_doubles = 0;
  double* linvt = nullptr;  size_t linvt_doubles = 0;
  int* done = nullptr;      size_t done_ints = 0;
  int* info = nullptr;      size_t info_ints = 0;
  Hyper* hyper = nullptr;   size_t hyper_n = 0;
  gpb_ll* ll = nullptr;     size_t ll_n = 0;
  double* aux = nullptr;    size_t aux_doubles = 0;   // xstar / predictive outputs / MH state
  SchedState* sched = nullptr;
  unsigned long long* trace = nullptr; size_t trace_n = 0;  // optional per-task schedule trace
  int trace_on = 0; long long trace_tasks = 0;
  SchedState* h_sched = nullptr;  // pinned
  // pinned host staging
  Hyper* h_hyper = nullptr; size_t h_hyper_n = 0;
  gpb_ll* h_ll = nullptr;   size_t h_ll_n = 0;
  long long launches = 0;
  float timing[3] = {0, 0, 0};
  std::string cuda_err;
  // MH loop (gpb_mh.cu): chain state + stream + output ring; flush path on a second stream
  double* mh = nullptr;      size_t mh_doubles = 0;     // chain state, host-supplied streams
  double* ring = nullptr;    size_t ring_doubles = 0;   // 2 segments of T iterations (+ gathered copies)
  double* h_stage = nullptr; size_t h_stage_doubles = 0;  // pinned staging of the flushed segments
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t ev_ready[2] = {nullptr, nullptr}, ev_flushed[2] = {nullptr, nullptr};
  // communicator (gpb_comm.cu): ncclComm_t behind a void* so that only gpb_comm.cu needs nccl.h
  void* comm = nullptr;
  int comm_rank = 0, comm_nranks = 1;
};

struct gpb_dataset {
  gpb_ctx* ctx;
  int n, dim;
  double* x;  // SoA [dim][n]
  double* y;  // [n]
};


namespace gpbh {
using namespace gpb;

int cuda_fail(gpb_ctx* ctx, cudaError_t e, const char* where);
#define CK(call)                                                   \
  do {                                                             \
    cudaError_t e_ = (call);                                       \
    if (e_ != cudaSuccess) return gpbh::cuda_fail(ctx, e_, #call); \
  } while (0)

template <typename T>
int grow(gpb_ctx* ctx, T** p, size_t* have, size_t want) {
  if (*have >= want) return GPB_OK;
  if (*p) { cudaFree(*p); *p = nullptr; *have = 0; }
  cudaError_t e = cudaMalloc((void**)p, want * sizeof(T));
  if (e != cudaSuccess) { cudaGetLastError(); return cuda_fail(ctx, e, "cudaMalloc(workspace)"); }
  *have = want;
  return GPB_OK;
}
template <typename T>
int grow_pinned(gpb_ctx* ctx, T** p, size_t* have, size_t want) {
  if (*have >= want) return GPB_OK;
  if (*p) { cudaFreeHost(*p); *p = nullptr; *have = 0; }
  cudaError_t e = cudaMallocHost((void**)p, want * sizeof(T));
  if (e != cudaSuccess) { cudaGetLastError(); return cuda_fail(ctx, e, "cudaMallocHost"); }
  *have = want;
  return GPB_OK;
}

Geom make_geom(int n, int dim, int m, bool with_rhs, int batch, int design = 0, bool identity = false);
int ensure_workspace(gpb_ctx* ctx, const Geom& g);
int fill_hyper(Hyper* h, const gpb_term* terms, int nterms, const double* beta, int nbeta, int dim);
int run_system(gpb_ctx* ctx, const Geom& g, const double* d_x, const double* d_y,
               const double* d_xstar, double nugget);
int finish_timing(gpb_ctx* ctx);
// Launch the factorisation / panel-solve kernel(s) on ctx->stream; counts launches; status out.
int do_chol(gpb_ctx* ctx, const Geom& g, int solve_only, int skip_reset = 0);
// gpb_comm.cu: all-gather `count` doubles per rank, device to device, on `st` (recv: [nranks][count]).
int comm_allgather(gpb_ctx* ctx, const double* d_send, double* d_recv, size_t count, cudaStream_t st);
void comm_release(gpb_ctx* ctx);

}  // namespace gpbh
