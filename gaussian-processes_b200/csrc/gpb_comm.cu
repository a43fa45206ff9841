// gpb_comm.cu -- the multi-GPU side of the C ABI: chains shard across the devices of one box and the
// only exchange is an all-gather of accepted states, log-likelihoods and accept flags.
//
// Replaces the reference's parallel chain runner: Streaming.writeParallelChain(iters, 2, nIters, ...)
// (unvendored bayesian_dlms; call sites examples/src/main/scala/SimulatedGp.scala:163-165,
// TemperatureDlm.scala:353-358), whose "gather" is one CSV file per chain.
//
// NCCL is bound at run time (dlopen "libnccl.so.2"): libgpb200.so keeps loading where NCCL is absent,
// and inside a process that already carries an NCCL (torch) the same instance is reused.  One
// communicator per gpb_ctx; one ctx per process (torchrun) or per host thread (JVM).
#include <dlfcn.h>
#include <nccl.h>

#include <cstdlib>
#include <mutex>

#include "gpb_host.h"

using namespace gpb;
using namespace gpbh;

namespace {

struct Nccl {
  void* handle = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  bool ok = false;
};

Nccl& nccl() {
  static Nccl n;
  static std::once_flag once;
  std::call_once(once, [] {
    // Order: GPB_NCCL_LIB (explicit path), an NCCL the process already carries (e.g. the one bundled
    // with torch: a second, different libnccl.so.2 in the same process would shadow it by soname),
    // then the system library.  RTLD_LOCAL: our binding must not leak symbols into later loads.
    const char* env = getenv("GPB_NCCL_LIB");
    if (env && env[0]) n.handle = dlopen(env, RTLD_NOW | RTLD_LOCAL);
    if (!n.handle) n.handle = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL | RTLD_NOLOAD);
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* nm : names) {
      if (n.handle) break;
      n.handle = dlopen(nm, RTLD_NOW | RTLD_LOCAL);
    }
    if (!n.handle) return;
    auto sym = [&](const char* s) { return dlsym(n.handle, s); };
    n.GetUniqueId = reinterpret_cast<decltype(n.GetUniqueId)>(sym("ncclGetUniqueId"));
    n.CommInitRank = reinterpret_cast<decltype(n.CommInitRank)>(sym("ncclCommInitRank"));
    n.CommInitAll = reinterpret_cast<decltype(n.CommInitAll)>(sym("ncclCommInitAll"));
    n.CommDestroy = reinterpret_cast<decltype(n.CommDestroy)>(sym("ncclCommDestroy"));
    n.AllGather = reinterpret_cast<decltype(n.AllGather)>(sym("ncclAllGather"));
    n.GetErrorString = reinterpret_cast<decltype(n.GetErrorString)>(sym("ncclGetErrorString"));
    n.ok = n.GetUniqueId && n.CommInitRank && n.CommInitAll && n.CommDestroy && n.AllGather && n.GetErrorString;
  });
  return n;
}

int nccl_fail(gpb_ctx* ctx, ncclResult_t r, const char* where) {
  if (ctx) ctx->cuda_err = std::string(where) + ": " + nccl().GetErrorString(r);
  return GPB_E_CUDA;
}

static_assert(sizeof(ncclUniqueId) == GPB_COMM_ID_BYTES, "gpb200.h publishes the id size");

}  // namespace

namespace gpbh {

int comm_allgather(gpb_ctx* ctx, const double* d_send, double* d_recv, size_t count, cudaStream_t st) {
  if (!ctx->comm) return GPB_E_NOCOMM;
  const ncclResult_t r = nccl().AllGather(d_send, d_recv, count, ncclDouble, static_cast<ncclComm_t>(ctx->comm), st);
  if (r != ncclSuccess) return nccl_fail(ctx, r, "ncclAllGather");
  return GPB_OK;
}

void comm_release(gpb_ctx* ctx) {
  if (ctx->comm) {
    nccl().CommDestroy(static_cast<ncclComm_t>(ctx->comm));
    ctx->comm = nullptr;
  }
  ctx->comm_rank = 0; ctx->comm_nranks = 1;
}

}  // namespace gpbh

extern "C" {

int gpb_comm_unique_id(void* id128) {
  if (!id128) return GPB_E_INVALID;
  if (!nccl().ok) return GPB_E_NOCOMM;
  ncclUniqueId id;
  if (nccl().GetUniqueId(&id) != ncclSuccess) return GPB_E_CUDA;
  std::memcpy(id128, &id, sizeof(id));
  return GPB_OK;
}

int gpb_comm_init_rank(gpb_ctx* ctx, int nranks, int rank, const void* id128) {
  if (!ctx || !id128 || nranks < 1 || rank < 0 || rank >= nranks) return GPB_E_INVALID;
  if (!nccl().ok) return GPB_E_NOCOMM;
  CK(cudaSetDevice(ctx->device));
  comm_release(ctx);
  ncclUniqueId id;
  std::memcpy(&id, id128, sizeof(id));
  ncclComm_t comm = nullptr;
  const ncclResult_t r = nccl().CommInitRank(&comm, nranks, id, rank);
  if (r != ncclSuccess) return nccl_fail(ctx, r, "ncclCommInitRank");
  ctx->comm = comm; ctx->comm_rank = rank; ctx->comm_nranks = nranks;
  return GPB_OK;
}

int gpb_comm_init_all(gpb_ctx** ctxs, int n) {
  if (!ctxs || n < 1 || n > 64) return GPB_E_INVALID;
  if (!nccl().ok) return GPB_E_NOCOMM;
  int devs[64];
  ncclComm_t comms[64];
  for (int i = 0; i < n; ++i) {
    if (!ctxs[i]) return GPB_E_INVALID;
    for (int k = 0; k < i; ++k) if (ctxs[k]->device == ctxs[i]->device) return GPB_E_INVALID;  // one rank per device
    devs[i] = ctxs[i]->device;
    comm_release(ctxs[i]);
  }
  const ncclResult_t r = nccl().CommInitAll(comms, n, devs);
  if (r != ncclSuccess) return nccl_fail(ctxs[0], r, "ncclCommInitAll");
  for (int i = 0; i < n; ++i) { ctxs[i]->comm = comms[i]; ctxs[i]->comm_rank = i; ctxs[i]->comm_nranks = n; }
  return GPB_OK;
}

int gpb_comm_destroy(gpb_ctx* ctx) {
  if (!ctx) return GPB_E_INVALID;
  cudaSetDevice(ctx->device);
  if (ctx->copy_stream) cudaStreamSynchronize(ctx->copy_stream);
  comm_release(ctx);
  return GPB_OK;
}

int gpb_comm_info(const gpb_ctx* ctx, int* rank, int* nranks) {
  if (!ctx) return GPB_E_INVALID;
  if (rank) *rank = ctx->comm_rank;
  if (nranks) *nranks = ctx->comm_nranks;
  return ctx->comm ? GPB_OK : GPB_E_NOCOMM;
}

// Block partition of `total` chains: the first total % nranks ranks own one chain more.
int gpb_comm_shard(int total, int nranks, int rank, int* first, int* count) {
  if (total < 0 || nranks < 1 || rank < 0 || rank >= nranks) return GPB_E_INVALID;
  const int base = total / nranks, rem = total % nranks;
  if (count) *count = base + (rank < rem ? 1 : 0);
  if (first) *first = rank * base + (rank < rem ? rank : rem);
  return GPB_OK;
}

int gpb_gather(gpb_ctx* ctx, const double* send, int count, double* recv) {
  if (!ctx || !send || !recv || count < 1) return GPB_E_INVALID;
  if (!ctx->comm) return GPB_E_NOCOMM;
  CK(cudaSetDevice(ctx->device));
  const size_t total = (size_t)count * (ctx->comm_nranks + 1);
  int rc = grow(ctx, &ctx->ring, &ctx->ring_doubles, total);
  if (rc) return rc;
  double* d_send = ctx->ring;
  double* d_recv = ctx->ring + count;
  cudaStream_t st = ctx->copy_stream;
  CK(cudaMemcpyAsync(d_send, send, sizeof(double) * count, cudaMemcpyHostToDevice, st));
  if ((rc = comm_allgather(ctx, d_send, d_recv, (size_t)count, st))) return rc;
  CK(cudaMemcpyAsync(recv, d_recv, sizeof(double) * count * ctx->comm_nranks, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  ctx->launches += 1;
  return GPB_OK;
}

}  // extern "C"
