// gpb_mh.cu -- device-resident Metropolis-Hastings over the kernel hyper-parameters for many
// chains, Philox streams, and the matrix-level triangular solves.
//
// Replaces (reference file:line under core/src/main/scala/gp/ unless noted):
//   KernelParameters.sample            KernelParameters.scala:231-246  (ll = prior + loglikelihood)
//   breeze MarkovChain.Kernels.metropolis (Breeze 0.13.2, unvendored): a = min(1, exp(new - old)),
//                                        u ~ U(0,1), u < a ? next : t
//   proposal / prior of the shipped example  examples/src/main/scala/SimulatedGp.scala:115-151
//   Predict.forwardSolve               Predict.scala:40-45   -> gpb_trsm_lower
//   Mcmc.backSolve (vector / matrix)   Mcmc.scala:35-41, 52-58 -> gpb_trsm_upper
//
// The MH loop never returns to the host inside an iteration: propose -> build -> factor/solve ->
// reduce -> accept are stream-ordered kernels; ll(t) is cached (the reference re-evaluates it).
#include "gpb_host.h"

using namespace gpb;
using namespace gpbh;

namespace {

struct MhDev {
  int chains, P, nterms, nbeta, formula, n;
  double delta;
  int kind[MAXT];
  double beta[3];
  double ig_shape[2 * MAXT], ig_scale[2 * MAXT], ig_const[2 * MAXT];  // const = a ln b - lgamma(a)
};

// it < 0: evaluate the initial state (proposal = current).
__global__ void k_mh_propose(MhDev md, int it, const double* __restrict__ cur,
                             const double* __restrict__ z, double* __restrict__ prop,
                             Hyper* __restrict__ hyper) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= md.chains) return;
  Hyper h;
  h.nterms = md.nterms; h.nbeta = md.nbeta;
  for (int i = 0; i < 3; ++i) h.beta[i] = md.beta[i];
  int p = 0;
  for (int t = 0; t < MAXT; ++t) {
    h.kind[t] = 0; h.p0[t] = 0.0; h.p1[t] = 0.0;
    if (t >= md.nterms) continue;
    h.kind[t] = md.kind[t];
    const int np = md.kind[t] == GPB_KIND_SE ? 2 : 1;
    for (int k = 0; k < np; ++k, ++p) {
      double v = cur[(size_t)c * md.P + p];
      // SimulatedGp.scala:118-127: new = old * exp(z), z ~ Gaussian(0, delta)
      if (it >= 0) v = v * exp(md.delta * z[((size_t)it * md.chains + c) * md.P + p]);
      prop[(size_t)c * md.P + p] = v;
      if (k == 0) h.p0[t] = v; else h.p1[t] = v;
    }
  }
  hyper[c] = h;
}

__device__ __forceinline__ double ig_logpdf(const MhDev& md, int p, double x) {
  // InverseGamma(shape, scale).logPdf (bayesian_dlms 0.4.0): a ln b - lnG(a) - (a+1) ln x - b/x
  return md.ig_const[p] - (md.ig_shape[p] + 1.0) * log(x) - md.ig_scale[p] / x;
}

__global__ void k_mh_accept(MhDev md, int it, const gpb_ll* __restrict__ ll,
                            const double* __restrict__ prop, const double* __restrict__ uu,
                            double* __restrict__ cur, double* __restrict__ cur_ll,
                            double* __restrict__ out_states, double* __restrict__ out_ll,
                            int* __restrict__ out_accept, int* __restrict__ first_bad) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= md.chains) return;
  const gpb_ll r = ll[c];
  double new_ll = 0.0;
  bool ok = r.info == 0;
  if (!ok) atomicCAS(first_bad, 0, r.info);
  if (ok) {
    const double C = md.n * 0.5 * log(2 * M_PI);
    const double llv = md.formula == GPB_FORMULA_GPML ? (-0.5 * r.Q - r.S - C)
                                                     : (-0.5 * r.Q + 0.5 * (r.S + C) * r.U1);
    // priorKernel: ps.map(term => ...).sum  (SimulatedGp.scala:143-151)
    double pr = 0.0;
    int p = 0;
    for (int t = 0; t < md.nterms; ++t) {
      if (md.kind[t] == GPB_KIND_SE) {
        const double v = ig_logpdf(md, p, prop[(size_t)c * md.P + p]) +
                         ig_logpdf(md, p + 1, prop[(size_t)c * md.P + p + 1]);
        pr = pr + v; p += 2;
      } else {
        pr = pr + ig_logpdf(md, p, prop[(size_t)c * md.P + p]); p += 1;
      }
    }
    new_ll = pr + llv;
  }
  if (it < 0) {  // initial state
    cur_ll[c] = ok ? new_ll : NAN;
    return;
  }
  bool acc = false;
  if (ok) {
    const double a = fmin(1.0, exp(new_ll - cur_ll[c]));
    acc = uu[(size_t)it * md.chains + c] < a;
  }
  if (acc) {
    for (int p = 0; p < md.P; ++p) cur[(size_t)c * md.P + p] = prop[(size_t)c * md.P + p];
    cur_ll[c] = new_ll;
  }
  const size_t o = (size_t)it * md.chains + c;
  for (int p = 0; p < md.P; ++p) out_states[o * md.P + p] = cur[(size_t)c * md.P + p];
  out_ll[o] = cur_ll[c];
  out_accept[o] = acc ? 1 : 0;
}

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

// ---- Philox4x32-10 (Salmon et al., SC'11) ----------------------------------------------------------
inline void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
  for (int r = 0; r < 10; ++r) {
    const uint64_t p0 = (uint64_t)0xD2511F53u * c[0], p1 = (uint64_t)0xCD9E8D57u * c[2];
    const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0, n1 = (uint32_t)p1;
    const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1, n3 = (uint32_t)p0;
    c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
}
inline double u53(uint32_t hi, uint32_t lo) {  // (0, 1)
  const uint64_t v = (((uint64_t)hi << 32) | lo) >> 11;
  return ((double)v + 0.5) * (1.0 / 9007199254740992.0);
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

int gpb_philox_streams(unsigned long long seed, int chain0, int chains, int iters, int P, double* z,
                       double* uu) {
  if (chains < 1 || iters < 0 || P < 0 || (P > 0 && !z) || !uu) return GPB_E_INVALID;
  const uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
  for (int it = 0; it < iters; ++it)
    for (int c = 0; c < chains; ++c) {
      const uint32_t chain = (uint32_t)(chain0 + c);
      for (int s = 0; 2 * s < P; ++s) {  // Box-Muller pairs
        uint32_t ctr[4] = {(uint32_t)it, (uint32_t)s, chain, 0u};
        philox4x32_10(ctr, k0, k1);
        const double r = std::sqrt(-2.0 * std::log(u53(ctr[0], ctr[1])));
        const double th = 2.0 * M_PI * u53(ctr[2], ctr[3]);
        double* zz = z + ((size_t)it * chains + c) * P;
        zz[2 * s] = r * std::cos(th);
        if (2 * s + 1 < P) zz[2 * s + 1] = r * std::sin(th);
      }
      uint32_t ctr[4] = {(uint32_t)it, 0xFFFFFFFFu, chain, 1u};
      philox4x32_10(ctr, k0, k1);
      uu[(size_t)it * chains + c] = u53(ctr[0], ctr[1]);
    }
  return GPB_OK;
}

int gpb_mh_run(gpb_ctx* ctx, const gpb_dataset* ds, int chains, int iters, const gpb_term* init,
               int nterms, const double* beta, int nbeta, const double* ig, const gpb_mh_opts* opts,
               const double* z, const double* uu, double* out_states, double* out_ll,
               int* out_accept, int* info_out) {
  if (!ctx || !ds || ds->ctx != ctx || chains < 1 || iters < 0 || !init || !ig || !opts || !uu ||
      !out_states || !out_ll || !out_accept)
    return GPB_E_INVALID;
  if (nterms < 1 || nterms > GPB_MAX_TERMS) return GPB_E_INVALID;
  if (!(nbeta == 0 || nbeta == ds->dim + 1) || (nbeta > 0 && !beta)) return GPB_E_INVALID;
  CK(cudaSetDevice(ctx->device));
  MhDev md;
  std::memset(&md, 0, sizeof(md));
  md.chains = chains; md.nterms = nterms; md.nbeta = nbeta; md.formula = opts->formula;
  md.n = ds->n; md.delta = opts->delta;
  for (int i = 0; i < nbeta; ++i) md.beta[i] = beta[i];
  int P = 0;
  for (int t = 0; t < nterms; ++t) {
    if (init[t].kind == GPB_KIND_MATERN) return GPB_E_UNSUPPORTED;
    if (init[t].kind != GPB_KIND_SE && init[t].kind != GPB_KIND_WHITE) return GPB_E_INVALID;
    md.kind[t] = init[t].kind;
    P += init[t].kind == GPB_KIND_SE ? 2 : 1;
  }
  md.P = P;
  if (!z) return GPB_E_INVALID;
  for (int p = 0; p < P; ++p) {
    md.ig_shape[p] = ig[2 * p]; md.ig_scale[p] = ig[2 * p + 1];
    md.ig_const[p] = ig[2 * p] * std::log(ig[2 * p + 1]) - std::lgamma(ig[2 * p]);
  }
  // all chains must share the kernel structure of chain 0
  std::vector<double> h_cur((size_t)chains * P);
  for (int c = 0; c < chains; ++c) {
    int p = 0;
    for (int t = 0; t < nterms; ++t) {
      const gpb_term& tm = init[(size_t)c * nterms + t];
      if (tm.kind != md.kind[t]) return GPB_E_INVALID;
      h_cur[(size_t)c * P + p++] = tm.p0;
      if (tm.kind == GPB_KIND_SE) h_cur[(size_t)c * P + p++] = tm.p1;
    }
  }
  const Geom g = make_geom(ds->n, ds->dim, 0, true, chains);
  int rc = ensure_workspace(ctx, g);
  if (rc) return rc;
  // aux layout: cur[chains*P] prop[chains*P] cur_ll[chains] z[iters*chains*P] uu[iters*chains]
  //             out_states[iters*chains*P] out_ll[iters*chains] | ints: out_accept[iters*chains] bad
  const size_t cp = (size_t)chains * P, ic = (size_t)iters * chains;
  const size_t nd = 2 * cp + chains + ic * P + ic + ic * P + ic;
  const size_t ni = ic + 1;
  if ((rc = grow(ctx, &ctx->aux, &ctx->aux_doubles, nd + (ni + 1) / 2 + 1))) return rc;
  double* d_cur = ctx->aux;
  double* d_prop = d_cur + cp;
  double* d_curll = d_prop + cp;
  double* d_z = d_curll + chains;
  double* d_uu = d_z + ic * P;
  double* d_states = d_uu + ic;
  double* d_outll = d_states + ic * P;
  int* d_acc = reinterpret_cast<int*>(d_outll + ic);
  int* d_bad = d_acc + ic;
  cudaStream_t st = ctx->stream;
  CK(cudaMemcpyAsync(d_cur, h_cur.data(), sizeof(double) * cp, cudaMemcpyHostToDevice, st));
  if (ic) {
    CK(cudaMemcpyAsync(d_z, z, sizeof(double) * ic * P, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(d_uu, uu, sizeof(double) * ic, cudaMemcpyHostToDevice, st));
  }
  CK(cudaMemsetAsync(d_bad, 0, sizeof(int), st));
  const int tb = 128, gb = (chains + tb - 1) / tb;
  CK(cudaEventRecord(ctx->ev[0], st));
  CK(cudaEventRecord(ctx->ev[1], st));
  for (int it = -1; it < iters; ++it) {
    k_mh_propose<<<gb, tb, 0, st>>>(md, it, d_cur, d_z, d_prop, ctx->hyper);
    launch_build(st, g, ds->x, ds->y, nullptr, ctx->hyper, opts->nugget, ctx->mats, BUILD_FAST, 0);
    ctx->launches += 2;
    if ((rc = do_chol(ctx, g, 0))) return rc;
    launch_ll_epilogue(st, g, ctx->mats, ctx->info, ctx->ll);
    k_mh_accept<<<gb, tb, 0, st>>>(md, it, ctx->ll, d_prop, d_uu, d_cur, d_curll, d_states, d_outll,
                                   d_acc, d_bad);
    ctx->launches += 2;
    if (it < 0 || opts->on_notpd == 0) {
      // mirror mode: a non-PD proposal stops the chain group (the reference throws); checking
      // costs one 4-byte read per iteration only in that mode after the initial state
      if (it < 0) {
        int h_bad = 0;
        CK(cudaMemcpyAsync(&h_bad, d_bad, sizeof(int), cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        if (h_bad != 0) { if (info_out) *info_out = h_bad; return GPB_E_NOTPD; }
      }
    }
  }
  CK(cudaEventRecord(ctx->ev[2], st));
  if (ic) {
    CK(cudaMemcpyAsync(out_states, d_states, sizeof(double) * ic * P, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(out_ll, d_outll, sizeof(double) * ic, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(out_accept, d_acc, sizeof(int) * ic, cudaMemcpyDeviceToHost, st));
  }
  int h_bad = 0;
  CK(cudaMemcpyAsync(&h_bad, d_bad, sizeof(int), cudaMemcpyDeviceToHost, st));
  rc = finish_timing(ctx);
  if (rc) return rc;
  if (info_out) *info_out = h_bad;
  if (h_bad != 0 && opts->on_notpd == 0) return GPB_E_NOTPD;
  return GPB_OK;
}

}  // extern "C"
