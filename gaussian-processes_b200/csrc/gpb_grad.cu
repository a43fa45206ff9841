// gpb_grad.cu -- gradient of the log marginal likelihood with respect to the kernel hyper-parameters
// (SURVEY.md 8f-2), on the same factorisation kernel.
//
// Replaces (reference file:line under core/src/main/scala/gp/):
//   GaussianProcess.mllGradient       GaussianProcess.scala:166-187
//       alpha = kxx \ ys ;  grad_p = 0.5 * sum(diag(alpha alpha^T G_p - kxx \ G_p))
//   KernelParameters.tangentMatrix    KernelParameters.scala:289-300   (one n x n matrix G_p per parameter)
//   KernelParameters.gradient         KernelParameters.scala:270-284   (the entries of G_p)
//
// The reference materialises every G_p and does one n x n LU solve per parameter.  Here
//   grad_p = 0.5 * sum_ij (alpha_i alpha_j - Kinv_ij) G_p(d_ij)
// with Kinv = L^-T L^-1 = V V^T, where V = L^-T comes out of the factorisation kernel itself: an
// identity block is appended below K, and the panel solves turn it into I L^-T (the zero tiles left of
// the block diagonal are skipped, so this costs n^3/3 flop on top of the n^3/3 of the Cholesky).
//   k_grad_alpha     alpha = V u (u = L^-1 (y - m(x)) is the solved rhs row)           HBM-bound, 4 n^2 B
//   k_grad_contract  per 64 x 64 tile pair: Kinv tile = sum_k V_Ik V_Jk^T on the DMMA pipe (n^3/3 flop,
//                    TMA ring as in k_chol_tasks), fused epilogue that never stores Kinv: distances and
//                    exp() recomputed from the coordinates, weighted sums per basis function
//   k_grad_finalize  fixed-order reduction over the tiles
// Every G_p the reference defines is a linear combination of four kinds of entry functions: E_t(d) =
// exp(-d^2/s_t^2), d^2 E_t(d), 1 and [d == 0]; the device returns the weighted sums of those, the host
// applies the coefficients of the requested mode (gpb_mll_gradient below).
#include <cuda.h>

#include "gpb_host.h"

namespace gpb {

bool make_tmap(CUtensorMap* tm, const Geom& g, double* mats);   // gpb_chol.cu

constexpr int GKC = 16;                 // V columns per pipeline stage (same box as k_chol_tasks)
constexpr int GSTAGES = 4;
constexpr int GCONS = 128;              // 4 consumer warps, 32 x 32 sub-tiles
constexpr int GTHREADS = GCONS + 32;    // + 1 producer warp
constexpr int GSUMS = 2 * MAXT + 2;     // [t] sum W E_t, [MAXT + t] sum W d^2 E_t, then sum W, sum W [d == 0]
constexpr uint32_t GSTAGE_TX = 2 * GKC * LDT * sizeof(double);

struct GradSpec {
  int nse;            // SE terms
  double ninv2[MAXT]; // -1 / s_t^2 of E_t(d) = exp(-d^2 / s_t^2)
};

struct __align__(128) GradSmem {
  double stage[GSTAGES][2 * GKC * LDT];   // [0..GKC) columns of V_J ("c"), [GKC..2GKC) of V_I ("r")
  uint64_t full[GSTAGES], empty[GSTAGES];
  double xi[2][TILE], xj[2][TILE], ai[TILE], aj[TILE];
  double red[GCONS / 32][GSUMS];
};

constexpr int ASLICE = 256;             // V columns per CTA of k_grad_alpha

// alpha_i = sum_{k >= i} V_ik u_k, as partial sums over k-slices of ASLICE columns:
// apart[b][slice][i].  One CTA per (64 rows, slice); thread = (row, one of 4 k phases), 4 independent
// chains each.  k_grad_alpha_sum adds the slices in fixed order.
__global__ void __launch_bounds__(256)
k_grad_alpha(Geom g, const double* __restrict__ mats, double* __restrict__ apart) {
  const int I = blockIdx.x, sl = blockIdx.y, b = blockIdx.z;
  const int r = threadIdx.x & (TILE - 1), ph = threadIdx.x >> 6;
  const int nsl = gridDim.y;
  double* out = apart + ((size_t)b * nsl + sl) * g.npad + I * TILE + r;
  const int kbeg = max(sl * ASLICE, I * TILE), kend = min((sl + 1) * ASLICE, g.npad);
  if (kbeg >= kend) { if (ph == 0) *out = 0.0; return; }   // V is upper triangular: nothing left of the diagonal
  const double* M = mats + (size_t)b * g.mat_stride;
  const double* Vrow = M + (size_t)g.ident0 * TILE + (size_t)I * TILE + r;
  double s[4] = {0.0, 0.0, 0.0, 0.0};
  // kbeg, kend are multiples of 64: 16-column steps, phase ph takes columns ph, ph+4, ph+8, ph+12
  for (int k = kbeg + ph; k < kend; k += 16)
#pragma unroll
    for (int q = 0; q < 4; ++q) s[q] = fma(Vrow[(size_t)(k + 4 * q) * g.ld], M[(size_t)(k + 4 * q) * g.ld + g.rhs_row], s[q]);
  __shared__ double part[4][TILE];
  part[ph][r] = (s[0] + s[1]) + (s[2] + s[3]);
  __syncthreads();
  if (ph == 0) *out = (part[0][r] + part[1][r]) + (part[2][r] + part[3][r]);
}
// alpha[b][i] = sum of the slices, fixed order.
__global__ void __launch_bounds__(256)
k_grad_alpha_sum(Geom g, int nsl, const double* __restrict__ apart, double* __restrict__ alpha) {
  const int i = blockIdx.x * 256 + threadIdx.x, b = blockIdx.y;
  if (i >= g.npad) return;
  double a = 0.0;
  for (int sl = i / ASLICE; sl < nsl; ++sl) a += apart[((size_t)b * nsl + sl) * g.npad + i];
  alpha[(size_t)b * g.npad + i] = a;
}

// One CTA per tile pair (I >= J) of one proposal; small I first: those have the longest contraction (k
// runs over tiles I .. nt-1), and consecutive CTAs work on the same proposal (its V stays in L2).
// A persistent variant with dynamic tickets was measured and was slower (18.8 vs 25.0 TFLOP/s at
// n = 4096, batch 8: interleaved proposals thrash L2), see profiles/r01_grad_experiments.txt.
__global__ void __launch_bounds__(GTHREADS, 2)
k_grad_contract(const __grid_constant__ CUtensorMap tmap, Geom g, const GradSpec* __restrict__ specs,
                const double* __restrict__ x, const double* __restrict__ alpha, double* __restrict__ part) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  GradSmem& sm = *reinterpret_cast<GradSmem*>(smem_raw);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int idx = blockIdx.x, b = blockIdx.y, tiles = gridDim.x;
  int I = (int)((sqrtf(8.0f * (float)idx + 1.0f) - 1.0f) * 0.5f);
  while (I * (I + 1) / 2 > idx) --I;
  while ((I + 1) * (I + 2) / 2 <= idx) ++I;
  const int J = idx - I * (I + 1) / 2;
  const int nstage = (g.nt - I) * (TILE / GKC);
  uint32_t stage = 0, phase = 0;

  if (warp == GCONS / 32) {
    // producer: one elected lane initialises the ring and streams the two row panels of V through it; the
    // first GSTAGES stages are on their way before the rest of the CTA has finished its prologue
    const int rowI = (g.ident0 + I) * TILE, rowJ = (g.ident0 + J) * TILE;
    auto issue = [&](int it) {
      const int col = I * TILE + it * GKC;
      mbar_arrive_expect_tx(&sm.full[stage], GSTAGE_TX);
      tma_load_3d(&sm.stage[stage][0], &tmap, rowJ, col, b, &sm.full[stage]);
      tma_load_3d(&sm.stage[stage][GKC * LDT], &tmap, rowI, col, b, &sm.full[stage]);
      if (++stage == GSTAGES) { stage = 0; phase ^= 1; }
    };
    int it = 0;
    if (lane == 0) {
      tma_prefetch_desc(&tmap);
      for (int s = 0; s < GSTAGES; ++s) { mbar_init(&sm.full[s], 1); mbar_init(&sm.empty[s], GCONS / 32); }
      mbar_fence_init();
      for (; it < nstage && it < GSTAGES; ++it) issue(it);
    }
    __syncwarp();
    __syncthreads();
    if (lane == 0)
      for (; it < nstage; ++it) {
        if (!mbar_wait(&sm.empty[stage], phase ^ 1)) break;
        issue(it);
      }
    return;
  }
  if (tid < TILE) {
    const int i = I * TILE + tid, j = J * TILE + tid;
    sm.xi[0][tid] = i < g.n ? x[i] : 0.0;
    sm.xi[1][tid] = (i < g.n && g.dim == 2) ? x[g.n + i] : 0.0;
    sm.xj[0][tid] = j < g.n ? x[j] : 0.0;
    sm.xj[1][tid] = (j < g.n && g.dim == 2) ? x[g.n + j] : 0.0;
    sm.ai[tid] = alpha[(size_t)b * g.npad + i];
    sm.aj[tid] = alpha[(size_t)b * g.npad + j];
  }
  __syncthreads();   // the ring is initialised; coordinates and alpha of both tiles are staged

  // consumers: acc[mt][nt] holds Kinv[c][r], c = wc*32 + mt*8 + gq (tile J), r = wr*32 + nt*8 + 2*tq + {0,1} (tile I)
  const int gq = lane >> 2, tq = lane & 3, wc = warp >> 1, wr = warp & 1;
  auto wait_full = [&](uint32_t st, uint32_t ph) {
    if (mbar_test_wait(&sm.full[st], ph)) return;   // plain probe: the stage is almost always there
    mbar_wait(&sm.full[st], ph);
  };
  auto load_frag = [&](const double* P, int ks, double (&a)[4], double (&bb)[4]) {
    const int k = ks * 4 + tq;
#pragma unroll
    for (int mt = 0; mt < 4; ++mt) a[mt] = P[k * LDT + wc * 32 + mt * 8 + gq];
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) bb[nt] = P[(GKC + k) * LDT + wr * 32 + nt * 8 + gq];
  };
  {
    double acc[4][4][2];
#pragma unroll
    for (int mt = 0; mt < 4; ++mt)
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) acc[mt][nt][0] = acc[mt][nt][1] = 0.0;
    int pend = -1;   // a stage is released one iteration late (see k_chol_tasks: LDS vs. mbarrier.arrive)
    {
      wait_full(stage, phase);
      double a[4], bb[4];
      load_frag(&sm.stage[stage][0], 0, a, bb);
#pragma unroll 1
      for (int it = 0; it < nstage; ++it) {
        const double* P = &sm.stage[stage][0];
        uint32_t nst = stage + 1, nph = phase;
        if (nst == GSTAGES) { nst = 0; nph ^= 1; }
        bool next_ready = true;
#pragma unroll
        for (int ks = 0; ks < GKC / 4; ++ks) {
          // the next k-step's fragments (or the next stage's first) are fetched before these 16 DMMAs issue
          double an[4], bn[4];
          // the barrier of the next stage is probed one k-step before its result is needed
          if (ks == GKC / 4 - 2) next_ready = (it + 1 >= nstage) || mbar_test_wait(&sm.full[nst], nph);
          if (ks + 1 < GKC / 4) {
            load_frag(P, ks + 1, an, bn);
          } else if (it + 1 < nstage) {
            if (!next_ready) wait_full(nst, nph);
            load_frag(&sm.stage[nst][0], 0, an, bn);
          } else {
#pragma unroll
            for (int i = 0; i < 4; ++i) an[i] = bn[i] = 0.0;
          }
#pragma unroll
          for (int mt = 0; mt < 4; ++mt)
#pragma unroll
            for (int nt = 0; nt < 4; ++nt) dmma884(acc[mt][nt][0], acc[mt][nt][1], a[mt], bb[nt]);
#pragma unroll
          for (int i = 0; i < 4; ++i) { a[i] = an[i]; bb[i] = bn[i]; }
        }
        __syncwarp();
        if (lane == 0 && pend >= 0) mbar_arrive(&sm.empty[pend]);
        pend = (int)stage;
        stage = nst; phase = nph;
      }
    }

    // fused epilogue: W_ij = alpha_i alpha_j - Kinv_ij against the basis functions of the tangent matrices.
    // Off-diagonal tile pairs stand for both (I, J) and (J, I): weight 2.
    const GradSpec& spec = specs[b];
    const int nse = spec.nse;
    const double wgt = (I == J) ? 1.0 : 2.0;
    double sums[GSUMS];
#pragma unroll
    for (int s = 0; s < GSUMS; ++s) sums[s] = 0.0;
#pragma unroll
    for (int mt = 0; mt < 4; ++mt)
#pragma unroll
      for (int nt = 0; nt < 4; ++nt)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int c = wc * 32 + mt * 8 + gq, r = wr * 32 + nt * 8 + 2 * tq + e;
          if (I * TILE + r >= g.n || J * TILE + c >= g.n) continue;
          const double W = wgt * (sm.ai[r] * sm.aj[c] - acc[mt][nt][e]);
          const double dx = sm.xi[0][r] - sm.xj[0][c];
          double d2 = __dmul_rn(dx, dx);
          if (g.dim == 2) { const double dy = sm.xi[1][r] - sm.xj[1][c]; d2 = __dadd_rn(d2, __dmul_rn(dy, dy)); }
          // the reference squares sqrt(d2) again (KernelParameters.scala:276); d2 itself is within an ulp of
          // that, and this epilogue is a latency chain: no square root, no division (-1/s^2 comes from the host)
          sums[2 * MAXT] += W;
          if (d2 == 0.0) sums[2 * MAXT + 1] += W;
          const double Wd = W * d2;
#pragma unroll
          for (int t = 0; t < MAXT; ++t)
            if (t < nse) {
              const double E = exp(d2 * __ldg(&spec.ninv2[t]));
              sums[t] = fma(W, E, sums[t]);
              sums[MAXT + t] = fma(Wd, E, sums[MAXT + t]);
            }
        }
    // acc has been consumed: every DMMA of the last stage completed, so did its LDS
    __syncwarp();
    if (lane == 0 && pend >= 0) mbar_arrive(&sm.empty[pend]);
#pragma unroll
    for (int s = 0; s < GSUMS; ++s) {
      double v = sums[s];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0) sm.red[warp][s] = v;
    }
    bar_named(1, GCONS);
    if (tid < GSUMS) {
      const double v = (sm.red[0][tid] + sm.red[1][tid]) + (sm.red[2][tid] + sm.red[3][tid]);
      part[((size_t)b * tiles + idx) * GSUMS + tid] = v;
    }
  }
}

// out[b][s] = sum over the tile pairs, fixed order (one CTA per proposal and sum).
__global__ void __launch_bounds__(256)
k_grad_finalize(int tiles, const double* __restrict__ part, const int* __restrict__ info, double* __restrict__ out) {
  const int b = blockIdx.x, s = blockIdx.y, tid = threadIdx.x;
  __shared__ double red[256];
  double v = 0.0;
  for (int t = tid; t < tiles; t += 256) v += part[((size_t)b * tiles + t) * GSUMS + s];
  red[tid] = v;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (tid < o) red[tid] += red[tid + o];
    __syncthreads();
  }
  if (tid == 0) out[(size_t)b * GSUMS + s] = info[b] != 0 ? nan("") : red[0];
}


static int alpha_slices(const Geom& g) { return (g.npad + ASLICE - 1) / ASLICE; }

static int launch_grad(cudaStream_t st, const Geom& g, const GradSpec* specs, const double* x, double* mats,
                       const int* info, double* apart, double* alpha, double* part, double* out) {
  CUtensorMap tmap;
  if (!make_tmap(&tmap, g, mats)) return -1;
  cudaFuncSetAttribute(k_grad_contract, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(GradSmem));
  const int tiles = g.nt * (g.nt + 1) / 2;
  const int nsl = alpha_slices(g);
  k_grad_alpha<<<dim3(g.nt, nsl, g.batch), 256, 0, st>>>(g, mats, apart);
  k_grad_alpha_sum<<<dim3((g.npad + 255) / 256, g.batch), 256, 0, st>>>(g, nsl, apart, alpha);
  k_grad_contract<<<dim3(tiles, g.batch), GTHREADS, sizeof(GradSmem), st>>>(tmap, g, specs, x, alpha, part);
  k_grad_finalize<<<dim3(g.batch, GSUMS), 256, 0, st>>>(tiles, part, info, out);
  return 4;
}

}  // namespace gpb

using namespace gpb;
using namespace gpbh;

extern "C" int gpb_mll_gradient(gpb_ctx* ctx, const gpb_dataset* ds, int batch, const gpb_term* terms, int nterms,
                                const double* beta, int nbeta, double nugget, int mode, double* grad, int* info) {
  if (!ctx || !ds || ds->ctx != ctx || !terms || !grad || batch < 1) return GPB_E_INVALID;
  if (mode != GPB_GRAD_AS_EXECUTED && mode != GPB_GRAD_ANALYTIC) return GPB_E_INVALID;
  if (nterms < 1 || nterms > GPB_MAX_TERMS) return GPB_E_INVALID;
  CK(cudaSetDevice(ctx->device));
  const Geom g = make_geom(ds->n, ds->dim, 0, true, batch, 0, true);
  int rc = ensure_workspace(ctx, g);
  if (rc) return rc;
  const int tiles = g.nt * (g.nt + 1) / 2;
  const size_t spec_doubles = (sizeof(GradSpec) * batch + 7) / 8;
  const size_t n_apart = (size_t)batch * g.npad * alpha_slices(g), n_alpha = (size_t)batch * g.npad, n_part = (size_t)batch * tiles * GSUMS, n_out = (size_t)batch * GSUMS;
  if ((rc = grow(ctx, &ctx->aux, &ctx->aux_doubles, n_apart + n_alpha + n_part + n_out + spec_doubles))) return rc;
  double* d_apart = ctx->aux;
  double* d_alpha = d_apart + n_apart;
  double* d_part = d_alpha + n_alpha;
  double* d_out = d_part + n_part;
  GradSpec* d_spec = reinterpret_cast<GradSpec*>(d_out + n_out);

  // K is built from the values exactly as passed (as executed, GaussianProcess.scala:171-176) or from
  // their exponentials (analytic: the parameters are on the log scale); the tangent entries always use
  // the exponentials (boundedBelow(0), KernelParameters.scala:272-274,280).
  std::vector<GradSpec> specs(batch);
  std::vector<gpb_term> kt(nterms);
  for (int b = 0; b < batch; ++b) {
    GradSpec& sp = specs[b];
    std::memset(&sp, 0, sizeof(sp));
    for (int t = 0; t < nterms; ++t) {
      const gpb_term& th = terms[(size_t)b * nterms + t];
      // grad is [batch][P] with P taken from element 0: every element must share its term structure
      if (th.kind != terms[t].kind) return GPB_E_INVALID;
      kt[t] = th;
      if (th.kind == GPB_KIND_SE) {
        const double s = std::exp(th.p1);
        sp.ninv2[sp.nse++] = -1.0 / (s * s);
        if (mode == GPB_GRAD_ANALYTIC) { kt[t].p0 = std::exp(th.p0); kt[t].p1 = std::exp(th.p1); }
      } else if (th.kind == GPB_KIND_WHITE) {
        if (mode == GPB_GRAD_ANALYTIC) kt[t].p0 = std::exp(th.p0);
      }
    }
    if ((rc = fill_hyper(&ctx->h_hyper[b], kt.data(), nterms, beta ? beta + (size_t)b * nbeta : nullptr, nbeta, ds->dim)))
      return rc;
  }
  CK(cudaMemcpyAsync(ctx->hyper, ctx->h_hyper, sizeof(Hyper) * batch, cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaMemcpyAsync(d_spec, specs.data(), sizeof(GradSpec) * batch, cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaEventRecord(ctx->ev[0], ctx->stream));
  launch_build(ctx->stream, g, ds->x, ds->y, nullptr, ctx->hyper, nugget, ctx->mats, BUILD_FAST, 0, nullptr);
  ctx->launches += 1;
  CK(cudaEventRecord(ctx->ev[1], ctx->stream));
  if ((rc = do_chol(ctx, g, 0))) return rc;
  CK(cudaEventRecord(ctx->ev[2], ctx->stream));
  const int nl = launch_grad(ctx->stream, g, d_spec, ds->x, ctx->mats, ctx->info, d_apart, d_alpha, d_part, d_out);
  if (nl < 0) { ctx->cuda_err = "cuTensorMapEncodeTiled failed (gradient contraction)"; return GPB_E_CUDA; }
  ctx->launches += nl;
  std::vector<double> sums(n_out);
  std::vector<int> h_info(batch);
  CK(cudaMemcpyAsync(sums.data(), d_out, sizeof(double) * n_out, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaMemcpyAsync(h_info.data(), ctx->info, sizeof(int) * batch, cudaMemcpyDeviceToHost, ctx->stream));
  if ((rc = finish_timing(ctx))) return rc;

  // coefficients of the basis sums per parameter, in KernelParameters.gradient's flatMap order
  int P = 0;
  for (int t = 0; t < nterms; ++t) P += terms[t].kind == GPB_KIND_SE ? 2 : 1;
  for (int b = 0; b < batch; ++b) {
    const double* S = &sums[(size_t)b * GSUMS];
    const double S1 = S[2 * MAXT], S0 = S[2 * MAXT + 1];
    double* out = grad + (size_t)b * P;
    int p = 0, si = 0;
    for (int t = 0; t < nterms; ++t) {
      const gpb_term& th = terms[(size_t)b * nterms + t];
      if (th.kind == GPB_KIND_SE) {
        const double h = std::exp(th.p0), s = std::exp(th.p1);
        const double SE = S[si], SD = S[MAXT + si];
        if (mode == GPB_GRAD_AS_EXECUTED) {
          out[p++] = 0.5 * (SE + S1);                                   // exp(-d^2/s^2) + h.derivative       (:276)
          out[p++] = 0.5 * (-(2 * h / (s * s * s)) * SE + S1);          // -(2h/s^3) exp(-d^2/s^2) + derivative (:277)
        } else {
          out[p++] = 0.5 * h * SE;                                      // dK/dlog h  = h E
          out[p++] = 0.5 * (2 * h / (s * s)) * SD;                      // dK/dlog s  = h E 2 d^2 / s^2
        }
        ++si;
      } else {
        if (mode == GPB_GRAD_AS_EXECUTED) out[p++] = 0.5 * (2.0 * S1);  // 1.0 + s.derivative for every d      (:281)
        else out[p++] = 0.5 * std::exp(th.p0) * S0;                     // dK/dlog s_w = s_w [d == 0]
      }
    }
    if (info) info[b] = h_info[b];
  }
  return GPB_OK;
}
