// gpb_noalloc.cu -- gpb_loglik_noalloc: one log marginal likelihood evaluation that allocates nothing.
//
// No counterpart in the reference; added for callers that
// cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
// it).  Same kernels as gpb_loglik (k_build -> k_chol_reset -> k_chol_tasks -> k_ll_epilogue); the
// workspace those kernels need -- the tall system, the L_jj^-T tiles, the progress counters, info, the
// scheduler state, the hyper-parameter block and the three reductions -- lives in ONE file-scope
// static __device__ array sized for n <= GPB_NOALLOC_MAX_N.  Every GPU that loads the library holds
// that array (2.7 MB) until the process exits.
//
// The array's name is a host-side stand-in in host code, not a device address: the address the
// kernels get comes from cudaGetSymbolAddress, taken once per device with that device current.
//
// Every call shares the array and finds the previous call's contents in it.  Nothing has to start at
// zero that the call does not write itself: k_build writes every tile of the lower trapezoid the
// factorisation reads (padding included), k_chol_reset the counters / info / ticket / identity tile,
// and the scheduler's error word is cleared here after a watchdog.  Calls that overlap on the GPU
// (two host threads, two streams) would race on the array: a process-wide mutex serialises the calls
// made through this entry point, and the call is synchronous on return.
#include <mutex>

#include "gpb_host.h"

using namespace gpb;

namespace {

constexpr int NA_MAX_N = GPB_NOALLOC_MAX_N;
constexpr int NA_NPAD = (NA_MAX_N + TILE - 1) / TILE * TILE;
constexpr int NA_LD = NA_NPAD + TILE;                       // one appended tile row: the rhs row
constexpr size_t align32(size_t doubles) { return (doubles + 31) / 32 * 32; }   // 256-byte sub-buffers
constexpr size_t NA_MATS = align32((size_t)NA_LD * NA_NPAD);
constexpr size_t NA_LINVT = align32(((size_t)NA_NPAD / TILE + 1) * TILE * LDT);
constexpr size_t NA_DONE = align32((3 * (size_t)NA_NPAD / TILE + 1 + 1) / 2);   // chol_done_ints: (rt + 2 nt) ints
constexpr size_t NA_INFO = 32;
constexpr size_t NA_SCHED = align32((sizeof(SchedState) + 7) / 8);
constexpr size_t NA_HYPER = align32((sizeof(Hyper) + 7) / 8);
constexpr size_t NA_LL = align32((sizeof(gpb_ll) + 7) / 8);
constexpr size_t NA_TOTAL = NA_MATS + NA_LINVT + NA_DONE + NA_INFO + NA_SCHED + NA_HYPER + NA_LL;

static __device__ __align__(256) double g_noalloc_scratch[NA_TOTAL];

constexpr int NA_MAXDEV = 64;
struct DeviceSlot { double* base = nullptr; int num_sms = 0; };
DeviceSlot g_slot[NA_MAXDEV];
std::mutex g_mu;

// once per device, that device current: the device address behind the symbol
int setup_device(int device, DeviceSlot** out) {
  DeviceSlot& s = g_slot[device];
  if (!s.base) {
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) { cudaGetLastError(); return GPB_E_NODEVICE; }
    if (prop.major != 10) return GPB_E_NODEVICE;   // sm_100a cubins only
    double* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p) { cudaGetLastError(); return GPB_E_CUDA; }
    s.num_sms = prop.multiProcessorCount;
    s.base = p;
  }
  *out = &s;
  return GPB_OK;
}

}  // namespace

extern "C" int gpb_loglik_noalloc(int device, void* stream, int n, int dim, const double* d_x,
                                  const double* d_y, const gpb_term* terms, int nterms, const double* beta,
                                  int nbeta, double nugget, gpb_ll* out) {
  if (!d_x || !d_y || !out || n < 1) return GPB_E_INVALID;
  if (dim != 1 && dim != 2) return GPB_E_UNSUPPORTED;
  if (n > NA_MAX_N) {
    std::fprintf(stderr, "gpb_loglik_noalloc: n = %d exceeds the static scratch (GPB_NOALLOC_MAX_N = %d); "
                         "use gpb_loglik\n", n, NA_MAX_N);
    return GPB_E_INVALID;
  }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return GPB_E_NODEVICE; }
  if (device < 0 || device >= ndev || device >= NA_MAXDEV) return GPB_E_INVALID;
  Hyper h;
  int rc = gpbh::fill_hyper(&h, terms, nterms, beta, nbeta, dim);
  if (rc) return rc;

  std::lock_guard<std::mutex> lock(g_mu);
  if (cudaSetDevice(device) != cudaSuccess) { cudaGetLastError(); return GPB_E_CUDA; }
  DeviceSlot* slot = nullptr;
  if ((rc = setup_device(device, &slot))) return rc;
  double* mats = slot->base;
  double* linvt = mats + NA_MATS;
  int* done = reinterpret_cast<int*>(linvt + NA_LINVT);
  int* info = reinterpret_cast<int*>(linvt + NA_LINVT + NA_DONE);
  SchedState* sched = reinterpret_cast<SchedState*>(linvt + NA_LINVT + NA_DONE + NA_INFO);
  Hyper* d_hyper = reinterpret_cast<Hyper*>(reinterpret_cast<double*>(sched) + NA_SCHED);
  gpb_ll* d_ll = reinterpret_cast<gpb_ll*>(reinterpret_cast<double*>(d_hyper) + NA_HYPER);

  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const Geom g = gpbh::make_geom(n, dim, 0, true, 1);
  cudaError_t e = cudaMemcpyAsync(d_hyper, &h, sizeof(Hyper), cudaMemcpyHostToDevice, st);   // pageable: staged before return
  if (e != cudaSuccess) { cudaGetLastError(); return GPB_E_CUDA; }
  launch_build(st, g, d_x, d_y, nullptr, d_hyper, nugget, mats, BUILD_FAST, 0);
  const int nl = launch_chol(st, g, mats, linvt, done, info, sched, 0, slot->num_sms);
  if (nl < 0) return nl == -2 ? GPB_E_UNSUPPORTED : GPB_E_CUDA;
  launch_ll_epilogue(st, g, mats, info, d_ll);
  int head[2] = {0, 0};   // SchedState: next_task, error
  e = cudaMemcpyAsync(out, d_ll, sizeof(gpb_ll), cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaMemcpyAsync(head, sched, SCHED_HEAD_BYTES, cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  if (e == cudaSuccess) e = cudaGetLastError();
  if (e != cudaSuccess) { cudaGetLastError(); return GPB_E_CUDA; }
  if (head[1] != 0) {
    cudaMemsetAsync(sched, 0, SCHED_HEAD_BYTES, st);
    cudaStreamSynchronize(st);
    return GPB_E_WATCHDOG;
  }
  return GPB_OK;
}
