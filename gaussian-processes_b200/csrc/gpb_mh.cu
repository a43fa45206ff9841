// gpb_mh.cu -- device-resident Metropolis-Hastings over the kernel hyper-parameters for many chains:
// the Mcmc.sample iteration (mean Gibbs step + kernel MH step), counter-based RNG streams, an output
// ring flushed (and all-gathered over NCCL) every T iterations on a second stream, and a one-launch
// kernel for the small systems of the shipped example.
//
// Replaces (reference file:line under core/src/main/scala/gp/ unless noted):
//   Mcmc.sample                        Mcmc.scala:63-76  (step = KernelParameters.sample compose
//                                      MeanParameters.sample: the mean Gibbs step runs FIRST)
//   KernelParameters.sample            KernelParameters.scala:231-246  (ll = prior + loglikelihood)
//   MeanParameters.sample / samplePlane  MeanParameters.scala:74-121
//   breeze MarkovChain.Kernels.metropolis (Breeze 0.13.2, unvendored): a = min(1, exp(new - old)),
//                                        u ~ U(0,1), u < a ? next : t   (NaN rejects, as on the JVM)
//   proposal / prior of the shipped example  examples/src/main/scala/SimulatedGp.scala:115-151
//   the parallel chain runner          Streaming.writeParallelChain, SimulatedGp.scala:163-165
//
// The loop never returns to the host inside an iteration: propose -> build -> factor/solve -> reduce
// -> accept are stream-ordered kernels; with a fixed mean ll(t) is cached (the reference re-evaluates).
#include <algorithm>

#include "gpb_cov.cuh"
#include "gpb_host.h"
#include "gpb_plane.h"

using namespace gpb;
using namespace gpbh;

namespace {

constexpr int MAXP = 2 * MAXT;
constexpr int TINY_N = 64;        // the one-launch kernel holds the whole system in shared memory
constexpr int TINY_LD = TINY_N + 2;
constexpr int TINY_THREADS = 128;

struct MhDev {
  int chains, P, nterms, nbeta, formula, n, dim;
  int d;          // coefficients the Gibbs step draws (nbeta with the mean step, else 0)
  int zdim, W;    // normals per (iteration, chain); state width P + d
  int RW;         // ring row: W states, ll, accept
  int rng, chain0, on_notpd;
  uint32_t k0, k1;
  double delta, nugget, prior_mean, prior_var;
  int kind[MAXT];
  double ig_shape[MAXP], ig_scale[MAXP], ig_const[MAXP];  // const = a ln b - lgamma(a)
};

// Device state of a run (all in ctx->mh).
struct MhState {
  double* cur;      // [chains][P]
  double* beta;     // [chains][3]
  double* prop;     // [chains][P]
  double* cur_ll;   // [chains]
  double* gstat;    // [chains][d*d + d]   Gibbs statistics
  double* z;        // host streams [iters][chains][zdim] (GPB_RNG_HOST)
  double* uu;       // [iters][chains]
  double* design;   // [d][n]
  int* dead;        // [chains] 0 alive, k + 1 = stopped at iteration k
  int* dinfo;       // [chains] LAPACK info that stopped the chain
  int* first_bad;   // first info > 0 seen on this device
};

// ---- Philox4x32-10 (Salmon et al., SC'11) ----------------------------------------------------------
__host__ __device__ inline void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
  for (int r = 0; r < 10; ++r) {
    const uint64_t p0 = (uint64_t)0xD2511F53u * c[0], p1 = (uint64_t)0xCD9E8D57u * c[2];
    const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0, n1 = (uint32_t)p1;
    const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1, n3 = (uint32_t)p0;
    c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
}
__host__ __device__ inline double u53(uint32_t hi, uint32_t lo) {  // (0, 1)
  const uint64_t v = (((uint64_t)hi << 32) | lo) >> 11;
  return ((double)v + 0.5) * (1.0 / 9007199254740992.0);
}
// Standard normal `slot` of (iteration, chain): Box-Muller pair slot / 2, counter (iter, pair, chain, 0).
__host__ __device__ inline double philox_normal(uint32_t k0, uint32_t k1, int it, int chain, int slot) {
  uint32_t ctr[4] = {(uint32_t)it, (uint32_t)(slot >> 1), (uint32_t)chain, 0u};
  philox4x32_10(ctr, k0, k1);
  const double r = sqrt(-2.0 * log(u53(ctr[0], ctr[1])));
  const double th = 2.0 * M_PI * u53(ctr[2], ctr[3]);
  return (slot & 1) ? r * sin(th) : r * cos(th);
}
__host__ __device__ inline double philox_uniform(uint32_t k0, uint32_t k1, int it, int chain) {
  uint32_t ctr[4] = {(uint32_t)it, 0xFFFFFFFFu, (uint32_t)chain, 1u};
  philox4x32_10(ctr, k0, k1);
  return u53(ctr[0], ctr[1]);
}

__device__ __forceinline__ double mh_normal(const MhDev& md, const double* __restrict__ z, int it, int c, int slot) {
  if (md.rng == GPB_RNG_HOST) return z[((size_t)it * md.chains + c) * md.zdim + slot];
  return philox_normal(md.k0, md.k1, it, md.chain0 + c, slot);
}
__device__ __forceinline__ double mh_uniform(const MhDev& md, const double* __restrict__ uu, int it, int c) {
  if (md.rng == GPB_RNG_HOST) return uu[(size_t)it * md.chains + c];
  return philox_uniform(md.k0, md.k1, it, md.chain0 + c);
}

__global__ void k_philox_fill(MhDev md, int iters, double* __restrict__ z, double* __restrict__ uu) {
  const size_t total = (size_t)iters * md.chains;
  for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
    const int it = (int)(e / md.chains), c = (int)(e % md.chains);
    for (int s = 0; s < md.zdim; ++s) z[e * md.zdim + s] = philox_normal(md.k0, md.k1, it, md.chain0 + c, s);
    uu[e] = philox_uniform(md.k0, md.k1, it, md.chain0 + c);
  }
}

// ---- the pieces of one Metropolis step, shared by the general and the one-launch path -------------
__device__ __forceinline__ double ig_logpdf(const MhDev& md, int p, double x) {
  // InverseGamma(shape, scale).logPdf (bayesian_dlms 0.4.0): a ln b - lnG(a) - (a+1) ln x - b/x
  return md.ig_const[p] - (md.ig_shape[p] + 1.0) * log(x) - md.ig_scale[p] / x;
}
// priorKernel: ps.map(term => ...).sum  (SimulatedGp.scala:143-151)
__device__ __forceinline__ double mh_prior(const MhDev& md, const double* __restrict__ par) {
  double pr = 0.0;
  int p = 0;
  for (int t = 0; t < md.nterms; ++t) {
    if (md.kind[t] == GPB_KIND_SE) {
      const double v = ig_logpdf(md, p, par[p]) + ig_logpdf(md, p + 1, par[p + 1]);
      pr = pr + v; p += 2;
    } else {
      pr = pr + ig_logpdf(md, p, par[p]); p += 1;
    }
  }
  return pr;
}
__device__ __forceinline__ double mh_ll_value(const MhDev& md, double Q, double U1, double S) {
  const double C = md.n * 0.9189385332046727;   // n/2 log(2 pi)
  return md.formula == GPB_FORMULA_GPML ? (-0.5 * Q - S - C) : (-0.5 * Q + 0.5 * (S + C) * U1);
}
// Breeze metropolis with JVM semantics: scala.math.min propagates NaN and `u < NaN` is false.
__device__ __forceinline__ bool mh_accepts(double new_ll, double old_ll, double u) {
  const double dl = new_ll - old_ll;
  return (dl == dl) && u < fmin(1.0, exp(dl));
}
__device__ __forceinline__ void fill_hyper_dev(const MhDev& md, const double* __restrict__ par,
                                               const double* __restrict__ beta, int nbeta, Hyper& h) {
  h.nterms = md.nterms; h.nbeta = nbeta;
  for (int i = 0; i < 3; ++i) h.beta[i] = (i < nbeta) ? beta[i] : 0.0;
  int p = 0;
  for (int t = 0; t < MAXT; ++t) {
    h.kind[t] = 0; h.p0[t] = 0.0; h.p1[t] = 0.0;
    if (t >= md.nterms) continue;
    h.kind[t] = md.kind[t];
    h.p0[t] = par[p++];
    if (md.kind[t] == GPB_KIND_SE) h.p1[t] = par[p++];
  }
}
__device__ __forceinline__ void write_row(const MhDev& md, double* __restrict__ row, const double* __restrict__ cur,
                                          const double* __restrict__ beta, double ll, double acc) {
  for (int p = 0; p < md.P; ++p) row[p] = cur[p];
  for (int a = 0; a < md.d; ++a) row[md.P + a] = beta[a];
  row[md.W] = ll; row[md.W + 1] = acc;
}
// A stopped chain's row: NaN state, NaN ll, accept slot = -info (decoded on the host).
__device__ __forceinline__ void write_dead_row(const MhDev& md, double* __restrict__ row, int info) {
  for (int p = 0; p < md.W + 1; ++p) row[p] = NAN;
  row[md.W + 1] = -(double)info;
}

// ---- general path: stream-ordered kernels around k_build / k_chol_tasks ---------------------------
// Gibbs systems: current kernel parameters, raw y (MeanParameters.scala:91: no mean subtraction).
__global__ void k_mh_gibbs_hyper(MhDev md, MhState s, Hyper* __restrict__ hyper) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= md.chains) return;
  Hyper h;
  fill_hyper_dev(md, s.cur + (size_t)c * md.P, nullptr, 0, h);
  hyper[c] = h;
}
// beta | rest: the d x d algebra of samplePlane per chain.
__global__ void k_mh_gibbs(MhDev md, MhState s, int it, const int* __restrict__ info) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= md.chains || s.dead[c]) return;
  const int d = md.d;
  int bad = info[c];
  double beta[4], z[4];
  if (bad == 0) {
    for (int a = 0; a < d; ++a) z[a] = mh_normal(md, s.z, it, c, md.P + a);
    const double* st = s.gstat + (size_t)c * (d * d + d);
    bad = plane_draw(d, st, st + d * d, md.prior_mean, md.prior_var, z, beta);
    if (bad < 0) bad = md.n + 1;   // singular d x d precision: report past the last pivot
  }
  if (bad != 0) {
    atomicCAS(s.first_bad, 0, bad);
    if (!md.on_notpd) { s.dead[c] = it + 1; s.dinfo[c] = bad; }   // the reference throws from cholesky(kxx)
    return;                                                     // reject mode: beta stays
  }
  for (int a = 0; a < d; ++a) s.beta[(size_t)c * 3 + a] = beta[a];
}
// it < 0: evaluate the initial state (proposal = current).  With the mean step the batch holds both
// sides of the comparison at the NEW beta: hyper[c] = current, hyper[chains + c] = proposal.
__global__ void k_mh_propose(MhDev md, MhState s, int it, Hyper* __restrict__ hyper) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= md.chains) return;
  const double* cur = s.cur + (size_t)c * md.P;
  double* prop = s.prop + (size_t)c * md.P;
  const double* beta = s.beta + (size_t)c * 3;
  // SimulatedGp.scala:118-127: new = old * exp(z), z ~ Gaussian(0, delta), left to right over the terms
  for (int p = 0; p < md.P; ++p)
    prop[p] = (it >= 0) ? cur[p] * exp(md.delta * mh_normal(md, s.z, it, c, p)) : cur[p];
  Hyper h;
  if (md.d > 0) {
    fill_hyper_dev(md, cur, beta, md.nbeta, h);
    hyper[c] = h;
    fill_hyper_dev(md, prop, beta, md.nbeta, h);
    hyper[md.chains + c] = h;
  } else {
    fill_hyper_dev(md, prop, beta, md.nbeta, h);
    hyper[c] = h;
  }
}
__global__ void k_mh_accept(MhDev md, MhState s, int it, const gpb_ll* __restrict__ ll, double* __restrict__ ring_rows) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= md.chains) return;
  double* row = (it >= 0) ? ring_rows + (size_t)c * md.RW : nullptr;
  if (s.dead[c]) {                         // stopped earlier (or by this iteration's Gibbs step)
    if (row) write_dead_row(md, row, s.dinfo[c]);
    return;
  }
  double* cur = s.cur + (size_t)c * md.P;
  const double* prop = s.prop + (size_t)c * md.P;
  const gpb_ll r = ll[(md.d > 0 ? md.chains : 0) + c];
  double old_ll = s.cur_ll[c];
  int bad = r.info;
  if (md.d > 0) {
    // ll(t) at the new beta, like Breeze's metropolis (its K is the one accepted earlier: info == 0
    // unless the state itself is degenerate)
    const gpb_ll ro = ll[c];
    if (ro.info != 0 && bad == 0) bad = ro.info;
    old_ll = (ro.info == 0) ? mh_prior(md, cur) + mh_ll_value(md, ro.Q, ro.U1, ro.S) : NAN;
  }
  if (bad != 0) {
    atomicCAS(s.first_bad, 0, bad);
    if (!md.on_notpd) {                    // mirror: this chain ends here (the reference throws)
      s.dead[c] = (it < 0 ? 0 : it) + 1; s.dinfo[c] = bad;
      if (row) write_dead_row(md, row, bad);
      return;
    }
  }
  const double new_ll = (bad == 0) ? mh_prior(md, prop) + mh_ll_value(md, r.Q, r.U1, r.S) : NAN;
  if (it < 0) { s.cur_ll[c] = new_ll; return; }   // initial state
  const bool acc = (bad == 0) && mh_accepts(new_ll, old_ll, mh_uniform(md, s.uu, it, c));
  if (acc) {
    for (int p = 0; p < md.P; ++p) cur[p] = prop[p];
    old_ll = new_ll;
  }
  s.cur_ll[c] = old_ll;
  write_row(md, row, cur, s.beta + (size_t)c * 3, old_ll, acc ? 1.0 : 0.0);
}

__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
// Fixed mean: everything between two factorisations in ONE kernel, one CTA per chain -- the likelihood
// reductions of the finished system (k_ll_epilogue's sums in the same order), the accept step, the NEXT
// proposal with its hyper-parameters, and the reset of k_chol_tasks' scheduler state for the next launch.
// An iteration is then three launches: k_build, k_chol_tasks, k_mh_step.
__global__ void __launch_bounds__(256)
k_mh_step(MhDev md, MhState s, Geom g, int it, int propose_next, const double* __restrict__ mats,
          int* __restrict__ info, int* __restrict__ done, SchedState* sched, Hyper* __restrict__ hyper,
          double* __restrict__ ring_rows) {
  const int c = blockIdx.x, tid = threadIdx.x;
  const double* M = mats + (size_t)c * g.mat_stride;
  double q = 0.0, u1 = 0.0, sl = 0.0;
#pragma unroll 4   // the loads of four rounds in flight (same sums in the same order)
  for (int col = tid; col < g.n; col += blockDim.x) {
    const double u = M[(size_t)col * g.ld + g.rhs_row];
    const double l = M[(size_t)col * g.ld + col];
    q = fma(u, u, q); u1 += u; sl += log(l);
  }
  __shared__ double red[3][8];
  __shared__ double sprop[MAXP];
  __shared__ int alive;
  q = warp_sum_d(q); u1 = warp_sum_d(u1); sl = warp_sum_d(sl);
  if ((tid & 31) == 0) { red[0][tid >> 5] = q; red[1][tid >> 5] = u1; red[2][tid >> 5] = sl; }
  __syncthreads();
  double* cur = s.cur + (size_t)c * md.P;
  double* prop = s.prop + (size_t)c * md.P;
  if (tid == 0) {
    double Q = 0, U1 = 0, S = 0;
    for (int i = 0; i < 8; ++i) { Q += red[0][i]; U1 += red[1][i]; S += red[2][i]; }
    double* row = (it >= 0) ? ring_rows + (size_t)c * md.RW : nullptr;
    int bad = info[c];
    if (s.dead[c]) {
      if (row) write_dead_row(md, row, s.dinfo[c]);
    } else {
      if (bad != 0) atomicCAS(s.first_bad, 0, bad);
      if (bad != 0 && !md.on_notpd) {          // mirror: this chain ends here (the reference throws)
        s.dead[c] = (it < 0 ? 0 : it) + 1; s.dinfo[c] = bad;
        if (row) write_dead_row(md, row, bad);
      } else {
        const double new_ll = (bad == 0) ? mh_prior(md, prop) + mh_ll_value(md, Q, U1, S) : NAN;
        if (it < 0) {
          s.cur_ll[c] = new_ll;                // initial state
        } else {
          double old_ll = s.cur_ll[c];
          const bool acc = (bad == 0) && mh_accepts(new_ll, old_ll, mh_uniform(md, s.uu, it, c));
          if (acc) {
            for (int p = 0; p < md.P; ++p) cur[p] = prop[p];
            old_ll = new_ll;
          }
          s.cur_ll[c] = old_ll;
          write_row(md, row, cur, s.beta + (size_t)c * 3, old_ll, acc ? 1.0 : 0.0);
        }
      }
    }
    alive = s.dead[c] ? 0 : 1;
    info[c] = 0;                               // scheduler state of the next launch: info, ...
  }
  __syncthreads();
  // ... progress counters and pre-accumulation flags of this chain, the ticket and the per-SM pause state
  for (int i = tid; i < g.rt; i += blockDim.x) done[(size_t)c * g.rt + i] = 0;
  for (int i = tid; i < g.nt; i += blockDim.x) {
    done[(size_t)g.batch * g.rt + (size_t)c * g.nt + i] = 0;
    done[(size_t)g.batch * (g.rt + g.nt) + (size_t)c * g.nt + i] = 0;
  }
  if (c == 0) {
    if (tid == 0) sched->next_task = 0;
    for (int e = tid; e < SCHED_MAXSM; e += blockDim.x) { sched->smslot[e] = 0; sched->pause[2 * e] = 0; sched->pause[2 * e + 1] = 0; }
  }
  if (!propose_next || !alive) return;
  // the next proposal: one lane per parameter (SimulatedGp.scala:118-127), then the hyper-parameters
  if (tid < md.P) {
    const double v = cur[tid] * exp(md.delta * mh_normal(md, s.z, it + 1, c, tid));
    prop[tid] = v; sprop[tid] = v;
  }
  __syncthreads();
  if (tid == 0) {
    Hyper h;
    fill_hyper_dev(md, sprop, s.beta + (size_t)c * 3, md.nbeta, h);
    hyper[c] = h;
  }
}

// ---- one-launch path for n <= 64 ------------------------------------------------------------------
// One CTA per chain walks `count` iterations: proposal, covariance (the reference's exact operation
// order), left-looking Cholesky with the rhs row bordered below the matrix (u = L^-1 r falls out of the
// same column steps), reductions, accept -- everything in shared memory, no launch per iteration.  The
// shipped example runs at n = 20 (SimulatedGp.scala:112), where a launch costs more than the algebra.
struct TinySmem {
  double A[TINY_N * TINY_LD];   // column-major [c][r], rows 0..n-1 the matrix, row n the rhs
  double x[2][TINY_N], y[TINY_N];
  double piv;
  double red[3][TINY_THREADS / 32];
  double cur[MAXP], prop[MAXP], pterm[MAXP], beta[3];
  double cur_ll, uni;
  double ninv2[MAXT];           // -1 / sigma^2 per squared-exponential term
  Hyper h;
  int info;
};
// Log-likelihood reductions of the system for sm.h (info in sm.info); every thread returns them.
// blockDim.x = the n + 1 stacked rows rounded up to whole warps: at n = 20 every barrier is one warp.
__device__ void tiny_eval(const MhDev& md, TinySmem& sm, double& Q, double& U1, double& S) {
  const int tid = threadIdx.x, n = md.n, nthr = blockDim.x;
  // covariance, lower triangle (KernelFunction.buildCov + nugget), entries dealt out linearly; the
  // hot-path form of k_build (h exp(d^2 (-1/s^2)): no sqrt / division per entry, <= 2 ulp(h) from the
  // reference's operation order, tests/test_gpu_parity.py::test_hot_path_build_entrywise)
  const int tri = n * (n + 1) / 2;
  for (int e = tid; e < tri; e += nthr) {
    int r = (int)((sqrtf(8.0f * (float)e + 1.0f) - 1.0f) * 0.5f);
    while (r * (r + 1) / 2 > e) --r;
    while ((r + 1) * (r + 2) / 2 <= e) ++r;
    const int c = e - r * (r + 1) / 2;
    const double dx = sm.x[0][r] - sm.x[0][c];
    double d2 = __dmul_rn(dx, dx);
    if (md.dim == 2) { const double dy = sm.x[1][r] - sm.x[1][c]; d2 = __dadd_rn(d2, __dmul_rn(dy, dy)); }
    double v = 0.0;
    for (int t = 0; t < sm.h.nterms; ++t) {
      const double term = (sm.h.kind[t] == GPB_KIND_SE) ? sm.h.p0[t] * exp(d2 * sm.ninv2[t]) : ((d2 == 0.0) ? sm.h.p0[t] : 0.0);
      v = __dadd_rn(v, term);
    }
    if (r == c) v = __dadd_rn(v, md.nugget);
    sm.A[c * TINY_LD + r] = v;
  }
  for (int c = tid; c < n; c += nthr) {   // rhs row: residual y - m(x)
    double mu = 0.0;
    if (sm.h.nbeta > 0) {
      mu = __dadd_rn(sm.h.beta[0], __dmul_rn(sm.x[0][c], sm.h.beta[1]));
      if (md.dim == 2) mu = __dadd_rn(mu, __dmul_rn(sm.x[1][c], sm.h.beta[2]));
    }
    sm.A[c * TINY_LD + n] = sm.y[c] - mu;
  }
  if (tid == 0) sm.info = 0;
  __syncthreads();
  // column j: s_i = a_ij - sum_{k<j} l_ik l_jk for the rows i >= j (row n: the rhs), then scale by
  // 1/sqrt(pivot) (hardware approximation + two Newton steps, as in the diagonal blocks of k_chol_tasks:
  // the IEEE sqrt and division subroutines are most of a column's dependency chain at this size)
  for (int j = 0; j < n; ++j) {
    double s = 0.0;
    if (tid >= j && tid <= n) {
      // four partial sums, eight loads in flight: the loop is bound by shared-memory latency otherwise
      double s0 = sm.A[j * TINY_LD + tid], s1 = 0.0, s2 = 0.0, s3 = 0.0;
      const double* ai = &sm.A[tid];
      const double* aj = &sm.A[j];
      int k = 0;
#pragma unroll 2
      for (; k + 3 < j; k += 4) {
        const double x0 = ai[k * TINY_LD], x1 = ai[(k + 1) * TINY_LD], x2 = ai[(k + 2) * TINY_LD], x3 = ai[(k + 3) * TINY_LD];
        const double y0 = aj[k * TINY_LD], y1 = aj[(k + 1) * TINY_LD], y2 = aj[(k + 2) * TINY_LD], y3 = aj[(k + 3) * TINY_LD];
        s0 = fma(-x0, y0, s0); s1 = fma(-x1, y1, s1); s2 = fma(-x2, y2, s2); s3 = fma(-x3, y3, s3);
      }
      for (; k < j; ++k) s0 = fma(-ai[k * TINY_LD], aj[k * TINY_LD], s0);
      s = (s0 + s1) + (s2 + s3);
      if (tid == j) sm.piv = s;
    }
    __syncthreads();
    const double a = sm.piv;
    if (!(a > 0.0)) { if (tid == 0) sm.info = j + 1; break; }   // uniform: every thread reads the same pivot
    double rs;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(rs) : "d"(a));
    double e1 = fma(-(a * rs), rs, 1.0);
    rs = fma(0.5 * rs, e1, rs);
    e1 = fma(-(a * rs), rs, 1.0);
    rs = fma(0.5 * rs, e1, rs);
    if (tid >= j && tid <= n) {
      double l = s * rs;
      if (tid == j) { const double sv = a * rs; l = fma(0.5 * rs, fma(-sv, sv, a), sv); }   // sqrt(a) to ~0.5 ulp
      sm.A[j * TINY_LD + tid] = l;
    }
    __syncthreads();
  }
  __syncthreads();
  double q = 0.0, u1 = 0.0, sl = 0.0;
  if (sm.info == 0 && tid < n) {
    const double u = sm.A[tid * TINY_LD + n];
    q = u * u; u1 = u; sl = log(sm.A[tid * TINY_LD + tid]);
  }
  q = warp_sum_d(q); u1 = warp_sum_d(u1); sl = warp_sum_d(sl);
  if ((tid & 31) == 0) { sm.red[0][tid >> 5] = q; sm.red[1][tid >> 5] = u1; sm.red[2][tid >> 5] = sl; }
  __syncthreads();
  Q = U1 = S = 0.0;
  for (int w = 0; w < (nthr >> 5); ++w) { Q += sm.red[0][w]; U1 += sm.red[1][w]; S += sm.red[2][w]; }
}

__global__ void __launch_bounds__(TINY_THREADS)
k_mh_tiny(MhDev md, MhState s, const double* __restrict__ x, const double* __restrict__ y, int it0, int count,
          int first, double* __restrict__ ring) {
  __shared__ TinySmem sm;
  const int c = blockIdx.x, tid = threadIdx.x, n = md.n;
  for (int i = tid; i < n; i += blockDim.x) {
    sm.x[0][i] = x[i];
    sm.x[1][i] = (md.dim == 2) ? x[n + i] : 0.0;
    sm.y[i] = y[i];
  }
  if (tid < md.P) sm.cur[tid] = s.cur[(size_t)c * md.P + tid];
  if (tid < 3) sm.beta[tid] = s.beta[(size_t)c * 3 + tid];
  if (tid == 0) sm.cur_ll = s.cur_ll[c];
  __syncthreads();
  int dead = s.dead[c], dinfo = s.dinfo[c];
  for (int it = first ? -1 : it0; it < it0 + count; it = (it < 0 ? it0 : it + 1)) {
    double* row = (it >= 0) ? ring + ((size_t)(it - it0) * md.chains + c) * md.RW : nullptr;
    if (dead) {
      if (tid == 0 && row) write_dead_row(md, row, dinfo);
      continue;
    }
    // one lane per parameter: its normal, its proposal, its prior term; one more lane draws the uniform
    if (tid < md.P) {
      const double v = (it >= 0) ? sm.cur[tid] * exp(md.delta * mh_normal(md, s.z, it, c, tid)) : sm.cur[tid];
      sm.prop[tid] = v;
      sm.pterm[tid] = ig_logpdf(md, tid, v);
    } else if (tid == md.P && it >= 0) {
      sm.uni = mh_uniform(md, s.uu, it, c);
    }
    __syncthreads();
    if (tid == 0) {
      fill_hyper_dev(md, sm.prop, sm.beta, md.nbeta, sm.h);
      for (int t = 0; t < md.nterms; ++t)
        sm.ninv2[t] = (md.kind[t] == GPB_KIND_SE) ? -1.0 / (sm.h.p1[t] * sm.h.p1[t]) : 0.0;
    }
    __syncthreads();
    double Q, U1, S;
    tiny_eval(md, sm, Q, U1, S);
    const int bad = sm.info;
    if (bad != 0) {
      if (tid == 0) atomicCAS(s.first_bad, 0, bad);
      if (!md.on_notpd) {
        dead = (it < 0 ? 0 : it) + 1; dinfo = bad;
        if (tid == 0 && row) write_dead_row(md, row, bad);
        __syncthreads();
        continue;
      }
    }
    if (tid == 0) {
      double pr = 0.0;   // priorKernel in the reference's order: ps.map(...).sum (SimulatedGp.scala:143-151)
      for (int t = 0, p = 0; t < md.nterms; ++t) {
        if (md.kind[t] == GPB_KIND_SE) { pr = pr + (sm.pterm[p] + sm.pterm[p + 1]); p += 2; }
        else { pr = pr + sm.pterm[p]; p += 1; }
      }
      const double new_ll = (bad == 0) ? pr + mh_ll_value(md, Q, U1, S) : NAN;
      if (it < 0) {
        sm.cur_ll = new_ll;
      } else {
        const bool acc = (bad == 0) && mh_accepts(new_ll, sm.cur_ll, sm.uni);
        if (acc) {
          for (int p = 0; p < md.P; ++p) sm.cur[p] = sm.prop[p];
          sm.cur_ll = new_ll;
        }
        write_row(md, row, sm.cur, sm.beta, sm.cur_ll, acc ? 1.0 : 0.0);
      }
    }
    __syncthreads();
  }
  if (tid < md.P) s.cur[(size_t)c * md.P + tid] = sm.cur[tid];
  if (tid == 0) { s.cur_ll[c] = sm.cur_ll; s.dead[c] = dead; s.dinfo[c] = dinfo; }
}
// Segment trailer: [first_bad] so that the flushed (and gathered) data carries the status with it.
__global__ void k_mh_trailer(const int* __restrict__ first_bad, double* __restrict__ trailer) {
  trailer[0] = (double)*first_bad;
  trailer[1] = 0.0;
}

struct Partition { int first, count; };
Partition shard_of(int total, int nranks, int rank) {
  Partition p;
  gpb_comm_shard(total, nranks, rank, &p.first, &p.count);
  return p;
}

}  // namespace

extern "C" {

int gpb_philox_streams(unsigned long long seed, int chain0, int chains, int iters, int P, double* z,
                       double* uu) {
  if (chains < 1 || iters < 0 || P < 0 || (P > 0 && !z) || !uu) return GPB_E_INVALID;
  const uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
  for (int it = 0; it < iters; ++it)
    for (int c = 0; c < chains; ++c) {
      double* zz = z + ((size_t)it * chains + c) * P;
      for (int s = 0; s < P; ++s) zz[s] = philox_normal(k0, k1, it, chain0 + c, s);
      uu[(size_t)it * chains + c] = philox_uniform(k0, k1, it, chain0 + c);
    }
  return GPB_OK;
}

int gpb_philox_streams_device(gpb_ctx* ctx, unsigned long long seed, int chain0, int chains, int iters,
                              int P, double* z, double* uu) {
  if (!ctx || chains < 1 || iters < 0 || P < 0 || (P > 0 && !z) || !uu) return GPB_E_INVALID;
  if (iters == 0) return GPB_OK;
  CK(cudaSetDevice(ctx->device));
  MhDev md;
  std::memset(&md, 0, sizeof(md));
  md.chains = chains; md.zdim = P; md.chain0 = chain0;
  md.k0 = (uint32_t)seed; md.k1 = (uint32_t)(seed >> 32);
  const size_t ic = (size_t)iters * chains;
  int rc = grow(ctx, &ctx->mh, &ctx->mh_doubles, ic * P + ic);
  if (rc) return rc;
  double* d_z = ctx->mh;
  double* d_u = ctx->mh + ic * P;
  const int blocks = (int)std::min<size_t>((ic + 127) / 128, 1184);
  k_philox_fill<<<blocks, 128, 0, ctx->stream>>>(md, iters, d_z, d_u);
  ctx->launches += 1;
  if (P) CK(cudaMemcpyAsync(z, d_z, sizeof(double) * ic * P, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaMemcpyAsync(uu, d_u, sizeof(double) * ic, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  CK(cudaGetLastError());
  return GPB_OK;
}

int gpb_mh_run_ex(gpb_ctx* ctx, const gpb_dataset* ds, const gpb_mh_config* cfg_in, int chains, int iters,
                  const gpb_term* init, int nterms, const double* beta0, int nbeta, const double* ig,
                  const double* z, const double* uu, const double* design, double* out_states,
                  double* out_ll, int* out_accept, int* out_stop, gpb_mh_result* res) {
  if (!ctx || !ds || ds->ctx != ctx || !cfg_in || chains < 1 || iters < 0 || !init || !ig || !out_states ||
      !out_ll || !out_accept)
    return GPB_E_INVALID;
  if (cfg_in->struct_size < (int)offsetof(gpb_mh_config, small_n) || cfg_in->struct_size > (int)sizeof(gpb_mh_config))
    return GPB_E_INVALID;
  gpb_mh_config cfg;
  std::memset(&cfg, 0, sizeof(cfg));
  std::memcpy(&cfg, cfg_in, (size_t)cfg_in->struct_size);
  if (nterms < 1 || nterms > GPB_MAX_TERMS) return GPB_E_INVALID;
  if (!(nbeta == 0 || nbeta == ds->dim + 1) || (nbeta > 0 && !beta0)) return GPB_E_INVALID;
  if (cfg.formula != GPB_FORMULA_AS_PARSED && cfg.formula != GPB_FORMULA_GPML) return GPB_E_INVALID;
  if (cfg.rng != GPB_RNG_HOST && cfg.rng != GPB_RNG_PHILOX) return GPB_E_INVALID;
  if (!(cfg.delta >= 0.0) || !(cfg.nugget >= 0.0)) return GPB_E_INVALID;
  if (cfg.rng == GPB_RNG_HOST && iters > 0 && (!z || !uu)) return GPB_E_INVALID;
  if (cfg.mean_step && (nbeta == 0 || !design || !(cfg.mean_prior_var > 0.0) || !(cfg.mean_nugget >= 0.0)))
    return GPB_E_INVALID;
  if (cfg.gather && (!ctx->comm || cfg.total_chains < chains)) return cfg.gather && !ctx->comm ? GPB_E_NOCOMM : GPB_E_INVALID;
  CK(cudaSetDevice(ctx->device));

  MhDev md;
  std::memset(&md, 0, sizeof(md));
  md.chains = chains; md.nterms = nterms; md.nbeta = nbeta; md.formula = cfg.formula;
  md.n = ds->n; md.dim = ds->dim; md.delta = cfg.delta; md.nugget = cfg.nugget;
  md.rng = cfg.rng; md.chain0 = cfg.chain0; md.on_notpd = cfg.on_notpd ? 1 : 0;
  md.k0 = (uint32_t)cfg.seed; md.k1 = (uint32_t)(cfg.seed >> 32);
  md.prior_mean = cfg.mean_prior_mean; md.prior_var = cfg.mean_prior_var;
  int P = 0;
  for (int t = 0; t < nterms; ++t) {
    if (init[t].kind == GPB_KIND_MATERN) return GPB_E_UNSUPPORTED;
    if (init[t].kind != GPB_KIND_SE && init[t].kind != GPB_KIND_WHITE) return GPB_E_INVALID;
    md.kind[t] = init[t].kind;
    P += init[t].kind == GPB_KIND_SE ? 2 : 1;
  }
  const int d = cfg.mean_step ? nbeta : 0;
  md.P = P; md.d = d; md.zdim = P + d; md.W = P + d; md.RW = md.W + 2;
  for (int p = 0; p < P; ++p) {
    md.ig_shape[p] = ig[2 * p]; md.ig_scale[p] = ig[2 * p + 1];
    md.ig_const[p] = ig[2 * p] * std::log(ig[2 * p + 1]) - std::lgamma(ig[2 * p]);
  }
  // all chains must share the kernel structure of chain 0
  std::vector<double> h_init((size_t)chains * (P + 3), 0.0);
  for (int c = 0; c < chains; ++c) {
    int p = 0;
    for (int t = 0; t < nterms; ++t) {
      const gpb_term& tm = init[(size_t)c * nterms + t];
      if (tm.kind != md.kind[t]) return GPB_E_INVALID;
      h_init[(size_t)c * P + p++] = tm.p0;
      if (tm.kind == GPB_KIND_SE) h_init[(size_t)c * P + p++] = tm.p1;
    }
    for (int a = 0; a < nbeta; ++a) h_init[(size_t)chains * P + (size_t)c * 3 + a] = beta0[(size_t)c * nbeta + a];
  }

  // communicator geometry: every rank's block, padded to the widest
  const int nranks = cfg.gather ? ctx->comm_nranks : 1, rank = cfg.gather ? ctx->comm_rank : 0;
  const int total = cfg.gather ? cfg.total_chains : chains;
  int cpad = chains;
  if (cfg.gather) {
    const Partition me = shard_of(total, nranks, rank);
    if (me.count != chains || me.first != cfg.chain0) return GPB_E_INVALID;
    cpad = shard_of(total, nranks, 0).count;   // rank 0 owns the widest block
  }
  const bool tiny = cfg.small_n == GPB_SMALL_AUTO && ds->n <= TINY_N && !cfg.mean_step;
  int T = cfg.flush_every > 0 ? cfg.flush_every : (tiny ? 1024 : 32);
  if (T > iters) T = iters > 0 ? iters : 1;
  const size_t seg = (size_t)T * cpad * md.RW + 2;                 // + trailer
  const int nseg = (iters + T - 1) / T;

  // general-path geometries: the likelihood systems (one per chain, or current + proposal per chain
  // with the mean step) and the Gibbs systems (design rows appended, nugget 1e-6)
  const Geom g_ll = make_geom(ds->n, ds->dim, 0, true, d > 0 ? 2 * chains : chains);
  const Geom g_gibbs = make_geom(ds->n, ds->dim, 0, true, chains, d);
  int rc = GPB_OK;
  if (!tiny) {
    if ((rc = ensure_workspace(ctx, g_ll))) return rc;
    if (d > 0 && (rc = ensure_workspace(ctx, g_gibbs))) return rc;
  }
  // device state
  const size_t cp = (size_t)chains * P, ic = (size_t)iters * chains;
  const size_t nz = cfg.rng == GPB_RNG_HOST ? ic * md.zdim : 0, nu = cfg.rng == GPB_RNG_HOST ? ic : 0;
  const size_t ngs = (size_t)chains * (d * d + d), nds = (size_t)d * ds->n;
  const size_t nd = 2 * cp + (size_t)chains * 3 + chains + ngs + nz + nu + nds;
  const size_t ni = 2 * (size_t)chains + 2;
  if ((rc = grow(ctx, &ctx->mh, &ctx->mh_doubles, nd + (ni + 1) / 2 + 1))) return rc;
  if ((rc = grow(ctx, &ctx->ring, &ctx->ring_doubles, 2 * seg * (size_t)(1 + (cfg.gather ? nranks : 0))))) return rc;
  if ((rc = grow_pinned(ctx, &ctx->h_stage, &ctx->h_stage_doubles, 2 * seg * (size_t)nranks))) return rc;
  MhState s;
  s.cur = ctx->mh; s.beta = s.cur + cp; s.prop = s.beta + (size_t)chains * 3; s.cur_ll = s.prop + cp;
  s.gstat = s.cur_ll + chains; s.z = s.gstat + ngs; s.uu = s.z + nz; s.design = s.uu + nu;
  s.dead = reinterpret_cast<int*>(s.design + nds); s.dinfo = s.dead + chains; s.first_bad = s.dinfo + chains;
  double* ring[2] = {ctx->ring, ctx->ring + seg};
  double* gath[2] = {ctx->ring + 2 * seg, ctx->ring + 2 * seg + seg * nranks};
  double* stage[2] = {ctx->h_stage, ctx->h_stage + seg * nranks};

  cudaStream_t st = ctx->stream, cs = ctx->copy_stream;
  CK(cudaMemcpyAsync(s.cur, h_init.data(), sizeof(double) * cp, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(s.beta, h_init.data() + cp, sizeof(double) * chains * 3, cudaMemcpyHostToDevice, st));
  if (nz) CK(cudaMemcpyAsync(s.z, z, sizeof(double) * nz, cudaMemcpyHostToDevice, st));
  if (nu) CK(cudaMemcpyAsync(s.uu, uu, sizeof(double) * nu, cudaMemcpyHostToDevice, st));
  if (nds) CK(cudaMemcpyAsync(s.design, design, sizeof(double) * nds, cudaMemcpyHostToDevice, st));
  CK(cudaMemsetAsync(s.dead, 0, sizeof(int) * ni, st));
  CK(cudaMemsetAsync(ctx->ring, 0, sizeof(double) * 2 * seg, st));   // padded chain slots read as zero
  const int tb = 128, gb = (chains + tb - 1) / tb;
  CK(cudaEventRecord(ctx->ev[0], st));
  CK(cudaEventRecord(ctx->ev[1], st));

  // outputs of a stopped chain start as "ran to the end"
  const int C = total;
  if (out_stop) for (int c = 0; c < C; ++c) out_stop[c] = -1;
  int first_info = 0, stopped = 0;
  std::vector<int> stop_seen((size_t)C, -1);
  auto unpack = [&](int sg) {   // segment sg is in stage[sg & 1]: scatter it into the caller's arrays
    const int it0 = sg * T, cnt = std::min(T, iters - it0);
    for (int r = 0; r < nranks; ++r) {
      const Partition pr = cfg.gather ? shard_of(total, nranks, r) : Partition{0, chains};
      const double* src = stage[sg & 1] + (size_t)r * seg;
      for (int t = 0; t < cnt; ++t)
        for (int c = 0; c < pr.count; ++c) {
          const double* row = src + ((size_t)t * cpad + c) * md.RW;
          const size_t o = (size_t)(it0 + t) * C + pr.first + c;
          std::memcpy(out_states + o * md.W, row, sizeof(double) * md.W);
          out_ll[o] = row[md.W];
          const double a = row[md.W + 1];
          out_accept[o] = a > 0.5 ? 1 : 0;
          if (a < 0.0 && stop_seen[pr.first + c] < 0) {
            stop_seen[pr.first + c] = it0 + t; ++stopped;
            if (out_stop) out_stop[pr.first + c] = it0 + t;
          }
        }
      const int fb = (int)src[seg - 2];
      if (fb != 0 && first_info == 0) first_info = fb;
    }
  };

  for (int sg = 0; sg < nseg; ++sg) {
    const int it0 = sg * T, cnt = std::min(T, iters - it0), b = sg & 1;
    if (sg >= 2) {
      // ring segment b is being reused: its flush must be complete (device) and unpacked (host)
      CK(cudaEventSynchronize(ctx->ev_flushed[b]));
      unpack(sg - 2);
      CK(cudaStreamWaitEvent(st, ctx->ev_flushed[b], 0));
    }
    if (tiny) {
      // the n + 1 stacked rows (and the P + 1 sampling lanes) rounded up to whole warps
      const int tthreads = 32 * ((std::max(ds->n + 1, P + 1) + 31) / 32);
      k_mh_tiny<<<chains, tthreads, 0, st>>>(md, s, ds->x, ds->y, it0, cnt, sg == 0 ? 1 : 0, ring[b]);
      ctx->launches += 1;
    } else {
      for (int it = (sg == 0 ? -1 : it0); it < it0 + cnt; it = (it < 0 ? it0 : it + 1)) {
        if (d > 0 && it >= 0) {
          // MeanParameters.sample first (Mcmc.scala:73): beta | kernel parameters, y
          k_mh_gibbs_hyper<<<gb, tb, 0, st>>>(md, s, ctx->hyper);
          launch_build(st, g_gibbs, ds->x, ds->y, nullptr, ctx->hyper, cfg.mean_nugget, ctx->mats, BUILD_FAST, 0, s.design);
          ctx->launches += 2;
          if ((rc = do_chol(ctx, g_gibbs, 0))) return rc;
          launch_plane_epilogue(st, g_gibbs, ctx->mats, s.design, s.gstat);
          k_mh_gibbs<<<gb, tb, 0, st>>>(md, s, it, ctx->info);
          ctx->launches += 2;
        }
        double* rows_it = it >= 0 ? ring[b] + (size_t)(it - it0) * cpad * md.RW : nullptr;
        if (d == 0) {
          // fixed mean: k_build -> k_chol_tasks -> k_mh_step (reductions + accept + next proposal + reset)
          if (it < 0) { k_mh_propose<<<gb, tb, 0, st>>>(md, s, it, ctx->hyper); ctx->launches += 1; }
          launch_build(st, g_ll, ds->x, ds->y, nullptr, ctx->hyper, cfg.nugget, ctx->mats, BUILD_FAST, 0);
          if ((rc = do_chol(ctx, g_ll, 0, it >= 0 ? 1 : 0))) return rc;
          k_mh_step<<<chains, 256, 0, st>>>(md, s, g_ll, it, it + 1 < iters ? 1 : 0, ctx->mats, ctx->info, ctx->done,
                                            ctx->sched, ctx->hyper, rows_it);
          ctx->launches += 2;
          continue;
        }
        k_mh_propose<<<gb, tb, 0, st>>>(md, s, it, ctx->hyper);
        launch_build(st, g_ll, ds->x, ds->y, nullptr, ctx->hyper, cfg.nugget, ctx->mats, BUILD_FAST, 0);
        ctx->launches += 2;
        if ((rc = do_chol(ctx, g_ll, 0))) return rc;
        launch_ll_epilogue(st, g_ll, ctx->mats, ctx->info, ctx->ll);
        k_mh_accept<<<gb, tb, 0, st>>>(md, s, it, ctx->ll, rows_it);
        ctx->launches += 2;
      }
    }
    k_mh_trailer<<<1, 1, 0, st>>>(s.first_bad, ring[b] + seg - 2);
    ctx->launches += 1;
    CK(cudaEventRecord(ctx->ev_ready[b], st));
    // flush on the copy stream, overlapped with the next segment's iterations: all-gather device to
    // device (NCCL over NVLink), then one copy to pinned host memory
    CK(cudaStreamWaitEvent(cs, ctx->ev_ready[b], 0));
    const double* src = ring[b];
    if (cfg.gather) {
      if ((rc = comm_allgather(ctx, ring[b], gath[b], seg, cs))) return rc;
      src = gath[b];
      ctx->launches += 1;
    }
    CK(cudaMemcpyAsync(stage[b], src, sizeof(double) * seg * nranks, cudaMemcpyDeviceToHost, cs));
    CK(cudaEventRecord(ctx->ev_flushed[b], cs));
  }
  CK(cudaEventRecord(ctx->ev[2], st));
  if (nseg == 0 && !tiny) {
    // no iterations: still validate the initial state like the loop would
  }
  for (int sg = std::max(0, nseg - 2); sg < nseg; ++sg) {
    CK(cudaEventSynchronize(ctx->ev_flushed[sg & 1]));
    unpack(sg);
  }
  rc = finish_timing(ctx);
  if (rc) return rc;
  if (res) {
    res->info = first_info; res->stopped_chains = stopped; res->path = tiny ? 1 : 0; res->flushes = nseg;
    res->device_ms = ctx->timing[1];
  }
  if (stopped > 0) return GPB_E_NOTPD;
  return GPB_OK;
}

int gpb_mh_run(gpb_ctx* ctx, const gpb_dataset* ds, int chains, int iters, const gpb_term* init,
               int nterms, const double* beta, int nbeta, const double* ig, const gpb_mh_opts* opts,
               const double* z, const double* uu, double* out_states, double* out_ll,
               int* out_accept, int* info_out) {
  if (!opts || chains < 1) return GPB_E_INVALID;
  if (nbeta < 0 || nbeta > 3 || (nbeta > 0 && !beta)) return GPB_E_INVALID;
  gpb_mh_config cfg;
  std::memset(&cfg, 0, sizeof(cfg));
  cfg.struct_size = (int)sizeof(cfg);
  cfg.formula = opts->formula; cfg.on_notpd = opts->on_notpd; cfg.delta = opts->delta; cfg.nugget = opts->nugget;
  cfg.rng = GPB_RNG_HOST;
  std::vector<double> b0((size_t)chains * (nbeta > 0 ? nbeta : 1));
  for (int c = 0; c < chains; ++c)
    for (int a = 0; a < nbeta; ++a) b0[(size_t)c * nbeta + a] = beta[a];
  gpb_mh_result res;
  std::memset(&res, 0, sizeof(res));
  const int rc = gpb_mh_run_ex(ctx, ds, &cfg, chains, iters, init, nterms, nbeta ? b0.data() : nullptr, nbeta, ig,
                               z, uu, nullptr, out_states, out_ll, out_accept, nullptr, &res);
  if (info_out && (rc == GPB_OK || rc == GPB_E_NOTPD)) *info_out = res.info;
  return rc;
}

int gpb_mh_run_host(gpb_ctx* ctx, const gpb_dataset* ds, int chains, int iters, const gpb_term* init,
                    int nterms, const double* beta, int nbeta, const gpb_mh_opts* opts,
                    gpb_propose_fn propose, gpb_log_prior_fn log_prior, void* user, const double* uu,
                    double* out_states, double* out_ll, int* out_accept, int* info_out) {
  if (!ctx || !ds || ds->ctx != ctx || chains < 1 || iters < 0 || !init || !opts || !propose || !log_prior ||
      !out_states || !out_ll || !out_accept || (iters > 0 && !uu))
    return GPB_E_INVALID;
  if (nterms < 1 || nterms > GPB_MAX_TERMS) return GPB_E_INVALID;
  if (opts->formula != GPB_FORMULA_AS_PARSED && opts->formula != GPB_FORMULA_GPML) return GPB_E_INVALID;
  int P = 0;
  for (int t = 0; t < nterms; ++t) P += init[t].kind == GPB_KIND_SE ? 2 : 1;
  const size_t cp = (size_t)chains * P;
  std::vector<double> cur(cp), next(cp), pri(chains), cur_ll(chains), bet((size_t)chains * (nbeta > 0 ? nbeta : 1));
  std::vector<gpb_term> terms((size_t)chains * nterms);
  std::vector<gpb_ll> lls(chains);
  std::vector<char> dead(chains, 0);
  for (int c = 0; c < chains; ++c) {
    int p = 0;
    for (int t = 0; t < nterms; ++t) {
      const gpb_term& tm = init[(size_t)c * nterms + t];
      if (tm.kind != init[t].kind) return GPB_E_INVALID;
      cur[(size_t)c * P + p++] = tm.p0;
      if (tm.kind == GPB_KIND_SE) cur[(size_t)c * P + p++] = tm.p1;
    }
    for (int a = 0; a < nbeta; ++a) bet[(size_t)c * nbeta + a] = beta[a];
  }
  auto evaluate = [&](const std::vector<double>& par, std::vector<double>& out) -> int {
    for (int c = 0; c < chains; ++c) {
      int p = 0;
      for (int t = 0; t < nterms; ++t) {
        gpb_term& tm = terms[(size_t)c * nterms + t];
        tm.kind = init[t].kind; tm.p1 = tm.p2 = 0.0;
        tm.p0 = par[(size_t)c * P + p++];
        if (tm.kind == GPB_KIND_SE) tm.p1 = par[(size_t)c * P + p++];
      }
    }
    const int rc = gpb_loglik_batch(ctx, ds, chains, terms.data(), nterms, nbeta ? bet.data() : nullptr, nbeta,
                                    opts->nugget, lls.data());
    if (rc) return rc;
    log_prior(user, chains, par.data(), pri.data());
    for (int c = 0; c < chains; ++c) out[c] = pri[c] + gpb_ll_value(&lls[c], opts->formula);   // NaN when info != 0
    return GPB_OK;
  };
  int first_info = 0, stopped = 0;
  auto note_bad = [&](int c, int it) {
    if (lls[c].info == 0) return;
    if (first_info == 0) first_info = lls[c].info;
    if (!opts->on_notpd && !dead[c]) { dead[c] = 1; ++stopped; (void)it; }
  };
  int rc = evaluate(cur, cur_ll);
  if (rc) return rc;
  for (int c = 0; c < chains; ++c) note_bad(c, 0);
  std::vector<double> new_ll(chains);
  for (int it = 0; it < iters; ++it) {
    propose(user, it, chains, cur.data(), next.data());
    if ((rc = evaluate(next, new_ll))) return rc;
    for (int c = 0; c < chains; ++c) {
      const size_t o = (size_t)it * chains + c;
      note_bad(c, it);
      if (dead[c]) {
        for (int p = 0; p < P; ++p) out_states[o * P + p] = NAN;
        out_ll[o] = NAN; out_accept[o] = 0;
        continue;
      }
      const double dl = new_ll[c] - cur_ll[c];
      const bool acc = (dl == dl) && uu[o] < std::fmin(1.0, std::exp(dl));
      if (acc) {
        std::memcpy(&cur[(size_t)c * P], &next[(size_t)c * P], sizeof(double) * P);
        cur_ll[c] = new_ll[c];
      }
      std::memcpy(out_states + o * P, &cur[(size_t)c * P], sizeof(double) * P);
      out_ll[o] = cur_ll[c]; out_accept[o] = acc ? 1 : 0;
    }
  }
  if (info_out) *info_out = first_info;
  return stopped > 0 ? GPB_E_NOTPD : GPB_OK;
}

}  // extern "C"
