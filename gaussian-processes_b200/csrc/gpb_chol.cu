// gpb_chol.cu -- k_chol_tasks: blocked FP64 Cholesky + the forward panel solves of the appended
// rows (rhs and K*), as ONE persistent kernel that walks a left-looking tile task graph.
//
// Replaces (reference file:line under core/src/main/scala/gp/):
//   breeze.linalg.cholesky -> F2J dpotrf("L")   call sites GaussianProcess.scala:123, Predict.scala:69
//   Predict.forwardSolve   -> F2J dtrsv("L")    Predict.scala:40-45 (called at GaussianProcess.scala:124,
//                                               Predict.scala:89-90): the rhs / K* rows are appended
//                                               below K as extra tile rows, so  u^T = y^T L^-T  and
//                                               V = K* L^-T fall out of the same panel solves.
//
// Task (b, i, j) = 64x64 tile (tile row i, tile column j) of proposal b:
//     acc  = A_ij - sum_{k<j} L_ik L_jk^T          FP64 tensor cores (DMMA.8x8x4), operands streamed
//                                                   by tiled TMA loads through a 4-stage mbarrier
//                                                   ring filled by a dedicated producer warp
//     i==j : L_jj = chol(acc), and L_jj^-T by factoring the stacked [acc; I] (one row per thread in
//            registers, two columns per barrier)
//     i>j  : L_ij = acc * L_jj^-T                   DMMA again
// Task kinds (K_PLAIN / K_DIAG0 / K_PART / K_FUSED below): diagonal blocks are pre-accumulated off the
// critical path as a balanced SYRK (syrk_task) and factored by the CTA that solves the tile above
// them; the tile row that only holds the rhs row runs one 8-row fragment ("light" tasks).
// Tickets are handed out in topological order (by block column), decoded by the producer warp and
// passed to the consumers through shared memory; dependencies are per-tile-row progress counters
// polled with ld.acquire; every wait is bounded by a watchdog.  Two instantiations: <TRACE, 2> for
// throughput launches (two CTAs per SM, 168 registers), <TRACE, 1> for launches bound by the
// dependency chain (one CTA per SM, no register cap).
// Summation order is fixed by the task, not by the schedule: results are bit-reproducible.
//
// Roofline: trailing contraction n^3/3 flop on the FP64 tensor pipe (37.1 TFLOP/s measured DMMA
// issue peak on this B200, profiles/r01_dmma_shapes_microbench.log).
#include <cuda.h>
#include <cstdlib>

#include "gpb_internal.h"

namespace gpb {

#ifndef GPB_KC
#define GPB_KC 16
#endif
#ifndef GPB_STAGES
#define GPB_STAGES 4
#endif
#ifndef GPB_CTAS_PER_SM
#define GPB_CTAS_PER_SM 2
#endif
#ifndef GPB_MINBLOCKS
#define GPB_MINBLOCKS GPB_CTAS_PER_SM
#endif
constexpr int KC = GPB_KC;             // panel columns per pipeline stage
constexpr int STAGES = GPB_STAGES;
constexpr int CTAS_PER_SM = GPB_CTAS_PER_SM;  // co-resident CTAs: one factors/solves while the other contracts
#ifndef GPB_CONS_WARPS
#define GPB_CONS_WARPS 4
#endif
constexpr int CW = GPB_CONS_WARPS;     // consumer warps per CTA: 4 (32 x 32 sub-tiles) or 8 (32 c x 16 r)
static_assert(CW == 4 || CW == 8, "consumer warps");
constexpr int NCONS = 32 * CW;         // consumer threads
constexpr int RW = TILE / (CW / 2);    // tile rows r per consumer warp in the contraction
constexpr int NTW = RW / 8;            // 8-row fragments per warp
constexpr int NTHREADS = NCONS + 32;   // + 1 producer warp
constexpr int STAGE_DOUBLES = 2 * KC * LDT;
constexpr int SCOLS = 2 * KC;          // tile columns (of L_jj^-T, or of the diagonal scratch) one stage holds
constexpr int LT_STAGES = TILE / SCOLS;  // stages that carry one 64 x 64 L_jj^-T tile
constexpr int AT_STAGES = TILE / SCOLS;  // stages that carry the task's own tile (accumulator start)
static_assert(SCOLS == 32 && AT_STAGES == 2, "each consumer warp's 32 columns of the tile sit in one stage");
static_assert(STAGES >= AT_STAGES + 2, "ring too small");
constexpr uint32_t STAGE_TX_BYTES = 2 * KC * LDT * sizeof(double);  // two {LDT x KC} TMA boxes
constexpr int TRACE_SLOTS = 16;   // == CHOL_TRACE_SLOTS
constexpr long long WATCHDOG_CLOCKS = 4000000000LL;  // ~2 s at 1.9 GHz
#ifndef GPB_PAUSE_TASKS_PER_SM
#define GPB_PAUSE_TASKS_PER_SM 48
#endif
constexpr int PAUSE_TASKS_PER_SM = GPB_PAUSE_TASKS_PER_SM;
constexpr long long PAUSE_MAX_CLOCKS = 120000LL;      // ~60 us: a pause is advisory and bounded

struct __align__(128) CholSmem {
  double stage[STAGES][STAGE_DOUBLES];  // [0..KC) panel j ("c" rows), [KC..2KC) panel i ("r" rows)
  double Cs[TILE * LDT];                // C tile, column-major [c][r]
  double colbuf[2][2][2 * TILE];        // diagonal panel: two columns per barrier, double-buffered
  uint64_t full[STAGES], empty[STAGES];
  uint64_t tq_full[2], tq_empty[2];
  uint64_t rel_full[2], rel_empty[2];   // consumer -> producer: "tile stored, publish it"
  int* rel_ptr[2];
  int rel_val[2];
  int taskq[2];
  int4 taskd[2];                        // the decoded ticket (b, i, j, kind): consumers do not decode again
  int bad_col;
  // early hand-over of the first half of L_jj^-T (chain-bound instantiation): request set by the caller of
  // diag_factor, one-slot channel to the producer lane's release service
  int* hf_req;
  int* hf_ptr;
  int hf_count;
  uint64_t hf_full, hf_empty;
  int* pause_mine;                      // chain_pause: this CTA's "inside a critical section" flag ...
  int* pause_other;                     // ... and the flag of the CTA it shares the SM with
};

size_t chol_smem_bytes() { return sizeof(CholSmem); }
// [batch][nt] tiles L_jj^-T, followed by ONE identity tile (same pitch): the start value of the
// stacked inverse rows, TMA-loaded into the scratch stages of every diagonal factorisation
size_t chol_linvt_doubles(const Geom& g) { return ((size_t)g.batch * g.nt + 1) * TILE * LDT; }
long long chol_task_count(const Geom& g) {
  return (long long)g.batch * ((long long)g.nt * g.rt - (long long)g.nt * (g.nt - 1) / 2);
}

// Task kinds.  The diagonal block j+1 is not a task of its own: its contraction over the block columns
// k < j is pre-accumulated off the critical path (K_PART, stored back in place), and the CTA that
// solves tile (j+1, j) keeps X in shared memory, applies the last update X X^T and factors the block
// on the spot (K_FUSED) -- the dependency chain diag(j) -> (j+1, j) -> diag(j+1) loses one publication,
// one acquire and one TMA round trip per block column.  Summation order is unchanged.
enum : int { K_PLAIN = 0, K_DIAG0 = 1, K_PART = 2, K_FUSED = 3 };
struct TaskDesc { int b, i, j, kind; };

// Optional per-task trace (debug builds of the schedule analysis; trace == nullptr in production):
// [0] globaltimer at receipt, [1] clock at receipt, [2] first stage landed, [3] contraction done,
// [4] dependency on the diagonal block satisfied, [5] factor / solve done, [6] clock at release,
// [7] globaltimer at release, [8] main-loop stages that were not ready at first probe, [9] cycles the main loop spent blocked on stage data.
// Producer lane (same SM, same clock): [10] ticket posted, [11] the task's own tile issued, [12] last
// contraction stage issued, [13] L_jj^-T stages issued (or scratch granted).  [14] / [15] clock before / behind the
// factorisation of a diagonal block (diagonal and fused tasks).
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long v;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(v));
  return v;
}
__device__ __forceinline__ unsigned int smid() {
  unsigned int v;
  asm volatile("mov.u32 %0, %smid;" : "=r"(v));
  return v;
}

// Ticket order (topological; chosen so the next diagonal block is factored while the bulk of the
// current block column is still being solved):
//   group -1 : diag(0,0) of every proposal
//   group  j : B  (j+1, j+1)    partial accumulation over k < j, stored back (K_PART)   [if j+1 < nt]
//              A  (j+1, j)      solve, then the update and factorisation of diag(j+1)
//                               (K_FUSED; a plain solve when row j+1 is an appended row)  [if j+1 < rt]
//              C  (i, j), i >= j+2
//   (triangular-solve mode keeps the older meaning of the two slots: B inverts the given diagonal
//   block, A is void)
// tickets before group j (j <= nt-1):  batch * (1 + j*rt - j(j+1)/2 + j)
__device__ __host__ __forceinline__ long long group_start(int batch, int rt, int j) {
  return (long long)batch * (1 + (long long)j * rt - (long long)j * (j + 1) / 2 + j);
}
// Tickets are dealt out proposal-group by proposal-group (g.tgroup proposals per group, each group
// in the column order above): the panels a group streams stay L2-resident, and the tail of one
// group overlaps the head of the next.
__host__ __device__ __forceinline__ TaskDesc decode_task(const Geom& g, int t, int& jcache) {
  TaskDesc d;
  const int per_mat = g.nt * g.rt - g.nt * (g.nt - 1) / 2;
  const int per_group = g.tgroup * per_mat;
  const int gi = t / per_group;
  t -= gi * per_group;
  const int b0 = gi * g.tgroup;
  const int gb = min(g.tgroup, g.batch - b0);
  d.kind = K_PLAIN;
  if (t < gb) { d.b = b0 + t; d.i = 0; d.j = 0; d.kind = K_DIAG0; return d; }
  while (jcache + 1 < g.nt && (long long)t >= group_start(gb, g.rt, jcache + 1)) ++jcache;
  while (jcache > 0 && (long long)t < group_start(gb, g.rt, jcache)) --jcache;
  const int j = jcache;
  int local = t - (int)group_start(gb, g.rt, j);
  const int nA = (j + 1 < g.rt) ? gb : 0, nB = (j + 1 < g.nt) ? gb : 0;
  if (local < nB) { d.b = b0 + local; d.i = j + 1; d.j = j + 1; d.kind = K_PART; return d; }
  local -= nB;
  if (local < nA) { d.b = b0 + local; d.i = j + 1; d.j = j; d.kind = (j + 1 < g.nt) ? K_FUSED : K_PLAIN; return d; }
  local -= nA;
  const int per = g.rt - j - 2;
  d.b = b0 + local / per; d.i = j + 2 + local % per; d.j = j;
  return d;
}

// First block column of the contraction of task (i, j): 0 in general.  Tile row I = i - ident0 of an
// appended identity block is zero left of block column I, so its contraction starts at I and the
// tasks left of it are void (k_build wrote the zeros, launch_chol presets done[i] = I).  -1 = void.
__device__ __forceinline__ int task_k0(const Geom& g, const TaskDesc& d) {
  if (g.ident0 < 0 || d.i < g.ident0) return 0;
  const int I = d.i - g.ident0;
  if (I >= g.nt) return 0;   // the rhs rows behind the identity block are dense
  return d.j < I ? -1 : I;
}
// One launch instead of three memsets (+ a preset): zero the progress counters and flags, info and the
// ticket counter (`error` stays sticky until the host has read it); preset done[i] = I for the rows of
// an appended identity block.
__global__ void k_chol_reset(Geom g, int* __restrict__ done, int* __restrict__ info, SchedState* sched,
                             double* __restrict__ linvt) {
  const int per = g.rt + 2 * g.nt;           // done [batch][rt], then part [batch][nt], then half [batch][nt]
  const long long total = (long long)g.batch * per;
  // the identity tile behind the L_jj^-T tiles (the workspace may have moved since the last launch)
  double* ident = linvt + (size_t)g.batch * g.nt * TILE * LDT;
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < TILE * LDT; e += gridDim.x * blockDim.x)
    ident[e] = (e / LDT == e % LDT) ? 1.0 : 0.0;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    int v = 0;
    if (g.ident0 >= 0 && i < (long long)g.batch * g.rt) {
      const int row = (int)(i % g.rt);
      if (row >= g.ident0 && row - g.ident0 < g.nt) v = row - g.ident0;
    }
    done[i] = v;
    if (i < g.batch) info[i] = 0;
  }
  if (blockIdx.x == 0) {
    if (threadIdx.x == 0) sched->next_task = 0;
    for (int e = threadIdx.x; e < SCHED_MAXSM; e += blockDim.x) {
      sched->smslot[e] = 0; sched->pause[2 * e] = 0; sched->pause[2 * e + 1] = 0;
    }
  }
}

__device__ __forceinline__ bool wait_counter(const int* ctr, int need, int* error) {
  if (ld_acquire(ctr) >= need) return true;
  long long t0 = clock64();
  for (;;) {
    if (ld_acquire(ctr) >= need) return true;
    if (*(volatile int*)error) return false;
    if (clock64() - t0 > WATCHDOG_CLOCKS) { atomicExch(error, 1); return false; }
    __nanosleep(40);
  }
}


// 1/a and 1/sqrt(a) from the hardware approximations (2^-23) plus two Newton steps each: the IEEE
// division / rsqrt subroutines are several times as long.  (The division-free panel elimination of
// diag_factor needs no reciprocal; the range fallback keeps them on its dependency chain.)
__device__ __forceinline__ double fast_rcp(double a) {
  double x;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(x) : "d"(a));
  double e = fma(-a, x, 1.0);
  x = fma(x, e, x);
  e = fma(-a, x, 1.0);
  return fma(x, e, x);
}
__device__ __forceinline__ double fast_rsqrt(double a) {
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(a));
  double e = fma(-(a * y), y, 1.0);
  y = fma(0.5 * y, e, y);
  e = fma(-(a * y), y, 1.0);
  return fma(0.5 * y, e, y);
}

// Sign flip on the integer pipe (the FP64 pipe is the one the DMMAs need).
__device__ __forceinline__ double flip_sign(double x) {
  return __hiloint2double(__double2hiint(x) ^ (int)0x80000000, __double2loint(x));
}

// Factor the stacked 128x64 [C; I]: top rows (sm.Cs) end as L_jj, bottom rows W (starts as the
// identity) as L_jj^-T -- the inverse costs no extra dependent step.  W lives in LT_STAGES ring
// stages the producer grants to the diagonal task (stage s0 + c / SCOLS holds columns c of W, pitch
// LDT), so the kernel needs no second tile buffer.  Blocked by 16 columns:
//   panel   : one thread per stacked row keeps its 16 panel entries in registers; right-looking
//             inside the panel, two columns per named barrier, division-free (the entries still to be
//             eliminated carry a common factor; no reciprocal on the dependency chain, see below);
//   trailing: rank-16 update of the remaining columns on the FP64 tensor cores, in place in
//             shared memory.  Only 8 of the 16 row blocks are live at any step (top rows below
//             the panel, bottom rows the panel has reached), two per warp.
constexpr int PB = 16;
static_assert(SCOLS % 8 == 0, "a stage holds whole 8-column blocks of the scratch tile");
__device__ __forceinline__ double* wcol(CholSmem& sm, int s0, int c) {
  const unsigned uc = (unsigned)c;
  return &sm.stage[((unsigned)s0 + uc / SCOLS) % STAGES][(uc % SCOLS) * LDT];
}
// store_L = false leaves the L_jj tile in sm.Cs for write_L_tile: no task reads a diagonal tile of L
// (the solves use L_jj^-T), so the chain-bound instantiation publishes L_jj^-T first and writes L_jj
// behind the publication.
__device__ __forceinline__ void write_L_tile(const CholSmem& sm, double* __restrict__ Aij, int ld, int tid) {
  const int r = tid & (TILE - 1), half = tid >> 6;
#pragma unroll 4
  for (int c = half; c < TILE; c += NCONS / TILE)
    if (r >= c) Aij[(size_t)c * ld + r] = sm.Cs[c * LDT + r];
}
// SCALED: the block in sm.Cs has been equilibrated (diag_factor below); the global stores undo it with this
// thread's row constants (uL for a row of L_jj, uW for a row of L_jj^-T).
template <bool SCALED>
__device__ __noinline__ void diag_factor_t(CholSmem& sm, int s0, double* __restrict__ Aij,
                                           double* __restrict__ LT, int ld, int tid, bool store_L,
                                           double uL, double uW) {
  const int warp = tid >> 5, lane = tid & 31, gq = lane >> 2, tq = lane & 3;
  // W arrives as the identity (TMA, producer warp); every thread has waited on those stages
  const bool top = tid < TILE, act = tid < 2 * TILE;   // threads beyond the 128 stacked rows only keep the barriers
  const int rr = top ? tid : (tid - TILE) & (TILE - 1);  // entry (R, c): top Cs[c * LDT + rr], bottom wcol(c)[rr]
#pragma unroll 1
  for (int c0 = 0; c0 < TILE; c0 += PB) {
    double p[PB];
    if (top) {
#pragma unroll
      for (int q = 0; q < PB; ++q) p[q] = sm.Cs[(c0 + q) * LDT + rr];
    } else if (!act) {
#pragma unroll
      for (int q = 0; q < PB; ++q) p[q] = 0.0;
    } else {
#pragma unroll
      for (int q = 0; q < PB; ++q) p[q] = wcol(sm, s0, c0 + q)[rr];
    }
    // Division-free elimination: the entries of the panel that are still to be eliminated carry a common
    // factor `sig` (1 at the start of the panel).  Eliminating the column pair (c, c+1) with the 2 x 2 pivot
    // block [[a0, w], [w, b1]], det = a0 b1 - w^2, as
    //     S' = det S - (b1 v0 v0^T - w (v0 v1^T + v1 v0^T) + a0 v1 v1^T)      (= det x the Schur complement)
    // needs no reciprocal on the dependency chain: per entry  p' = det p - x0 g0 - x1 g1  with the
    // thread's own g0 = b1 y0 - w y1, g1 = a0 y1 - w y0.  det is normalised into [1, 2) by an exact power
    // of two, so sig grows by less than 2 per pair.  The finished columns leave the chain as true L
    // entries: L(r, c) = y0 / sqrt(a0 sig), L(r, c+1) = g1 / sqrt(det a0 sig)  (rsqrt off the chain).
    double sig = 1.0;
#pragma unroll
    for (int q = 0; q < PB; q += 2) {
      // two columns per barrier: every thread rebuilds column c+1 after the elimination of column c
      // from the two published (pre-elimination) columns, so the second publish + barrier is saved
      double* cb0 = sm.colbuf[(q >> 1) & 1][0];
      double* cb1 = sm.colbuf[(q >> 1) & 1][1];
      if (act) { cb0[tid] = p[q]; cb1[tid] = p[q + 1]; }
      bar_named(1, NCONS);
      const int c = c0 + q;
      const double a0 = cb0[c];                    // pivot of column c (x sig)
      const double w = cb0[c + 1];                 // entry (c+1, c)
      const double b1 = cb1[c + 1];
      if (tid == c && !(a0 > 0.0) && sm.bad_col < 0) sm.bad_col = c;   // first bad column only
      const double det = fma(a0, b1, -(w * w));    // a0 x the pivot of column c+1 after column c
      if (tid == c + 1 && (a0 > 0.0) && !(det > 0.0) && sm.bad_col < 0) sm.bad_col = c + 1;
      int ex = (__double2hiint(det) >> 20) & 0x7ff;
      ex = min(max(ex, 1), 2045);
      const double sc = __hiloint2double((2046 - ex) << 20, 0);   // 2^-k, exact
      const double dets = det * sc;
      const double y0 = p[q], y1 = p[q + 1];
      const double g0 = fma(b1, y0, -(w * y1)) * sc;
      const double g1r = fma(a0, y1, -(w * y0));   // == det for the thread of row c+1, bit for bit
      const double g1 = g1r * sc;
#pragma unroll
      for (int q2 = q + 2; q2 < PB; ++q2) {
        const double x0 = cb0[c0 + q2], x1 = cb1[c0 + q2];
        p[q2] = fma(-x1, g1, fma(-x0, g0, dets * p[q2]));
      }
      // off the chain: the two finished columns; the pivots get one more Newton step (L_cc ~0.5 ulp)
      const double as = a0 * sig;
      const double das = det * as;
      const double rs0 = fast_rsqrt(as), rs1 = fast_rsqrt(das);
      double l0 = y0 * rs0, l1 = g1r * rs1;
      if (tid == c) {                              // l0^2 sig = a0
        const double l2 = l0 * l0, lr = fma(l0, l0, -l2);
        l0 = fma(0.5 * rs0, fma(-lr, sig, fma(-l2, sig, a0)), l0);
      }
      if (tid == c + 1) {                          // l1^2 (a0 sig) = det
        const double l2 = l1 * l1, lr = fma(l1, l1, -l2);
        l1 = fma(0.5 * rs1, fma(-lr, as, fma(-l2, as, det)), l1);
      }
      p[q] = l0; p[q + 1] = l1;
      sig *= dets;
    }
    // the panel's columns are final: back to shared memory for the trailing update, and straight out to
    // global memory from the registers (L_jj columns from the top rows, L_jj^-T rows -- pitch LDT, the
    // layout the TMA boxes of the solves read -- from the bottom rows), so that the stores drain behind
    // the remaining panels instead of in front of the publication
    if (top) {
#pragma unroll
      for (int q = 0; q < PB; ++q) sm.Cs[(c0 + q) * LDT + rr] = p[q];
      if (store_L) {
#pragma unroll
        for (int q = 0; q < PB; ++q)
          if (rr >= c0 + q) Aij[(size_t)(c0 + q) * ld + rr] = SCALED ? p[q] * uL : p[q];
      }
    } else if (act) {
#pragma unroll
      for (int q = 0; q < PB; ++q) wcol(sm, s0, c0 + q)[rr] = p[q];
#pragma unroll
      for (int q = 0; q < PB; ++q) LT[(c0 + q) * LDT + rr] = (c0 + q >= rr) ? (SCALED ? p[q] * uW : p[q]) : 0.0;
    }
    const int cN = c0 + PB;                        // first trailing column
    // rows 0..31 of L_jj^-T (the first stage of the solves that wait for this block) are final after the
    // second panel: published on the way when the caller asked for it (sm.hf_req)
    const bool early = (cN == SCOLS) && (tid == 0) && (sm.hf_req != nullptr);
    int hfn = 0;
    if (early) {
      hfn = sm.hf_count;
      (void)mbar_wait(&sm.hf_empty, (hfn & 1) ^ 1);   // bounded; the producer lane serves the slot within a microsecond
      sm.hf_ptr = sm.hf_req;
    }
    bar_named(1, NCONS);                           // the panel is in shared memory; its global stores are ordered before the post
    if (early) { mbar_arrive(&sm.hf_full); sm.hf_count = hfn + 1; }
    if (cN < TILE) {
      // M(R, c2) -= sum_q P(R, q) P(c2, q)  for c2 >= cN;  D[m = c2][n = R], k = q
      // the warp's row blocks side by side: they share the panel-column operand, and their DMMA chains
      // (4 k-steps each) are independent
      constexpr int NH = 8 / CW;
      const int ntop = (TILE - cN) / 8;            // live top blocks: rows cN .. 63
      bool btop[NH]; int r0[NH]; double bf[NH][PB / 4];
#pragma unroll
      for (int h = 0; h < NH; ++h) {
        const int blk = NH * warp + h;             // live row block 0..7
        btop[h] = blk < ntop;
        r0[h] = btop[h] ? cN + 8 * blk : 8 * (blk - ntop);
#pragma unroll
        for (int ks = 0; ks < PB / 4; ++ks) {
          const int c = c0 + ks * 4 + tq;
          bf[h][ks] = btop[h] ? sm.Cs[c * LDT + r0[h] + gq] : wcol(sm, s0, c)[r0[h] + gq];
        }
      }
#pragma unroll 2
      for (int c2 = cN; c2 < TILE; c2 += 8) {
        double a[PB / 4];
#pragma unroll
        for (int ks = 0; ks < PB / 4; ++ks) a[ks] = -sm.Cs[(c0 + ks * 4 + tq) * LDT + c2 + gq];
        double2* dp[NH]; double2 dv[NH];
#pragma unroll
        for (int h = 0; h < NH; ++h) {
          double* col = btop[h] ? &sm.Cs[(c2 + gq) * LDT] : wcol(sm, s0, c2) + gq * LDT;
          dp[h] = reinterpret_cast<double2*>(&col[r0[h] + 2 * tq]);
          dv[h] = *dp[h];
        }
#pragma unroll
        for (int ks = 0; ks < PB / 4; ++ks)
#pragma unroll
          for (int h = 0; h < NH; ++h) dmma884(dv[h].x, dv[h].y, a[ks], bf[h][ks]);
#pragma unroll
        for (int h = 0; h < NH; ++h) *dp[h] = dv[h];
      }
      bar_named(1, NCONS);
    }
  }
}

// The division-free elimination forms products of two entries of the block (a0 b1, w^2, b1 y0, ...).  Every
// covariance matrix of the path is hundreds of orders of magnitude away from where those overflow, but a
// matrix-level gpb_potrf near the ends of the double range must not fail where dpotrf does not: when a
// diagonal entry lies outside 2^-300 .. 2^300 the block is first equilibrated, D A D with D = diag(2^-k_r),
// k_r = floor(exponent(a_rr) / 2) -- exact, and every diagonal entry lands in [1, 4); entries of an SPD
// block are bounded by its diagonal, Schur complements only shrink it -- and the stores undo it exactly:
// L = D^-1 L_s, L^-T = D L_s^-T, both row scalings (one constant per thread).  One reducing barrier per block
// on the ordinary path.
__device__ __forceinline__ void diag_factor(CholSmem& sm, int s0, double* __restrict__ Aij,
                                            double* __restrict__ LT, int ld, int tid, bool store_L) {
  int ex = 1023;
  if (tid < TILE) ex = (__double2hiint(sm.Cs[tid * LDT + tid]) >> 20) & 0x7ff;   // sign bit set: not PD anyway
  const bool in_range = ex >= 1023 - 300 && ex <= 1023 + 300;
  if (bar_named_and(1, NCONS, in_range)) {
    diag_factor_t<false>(sm, s0, Aij, LT, ld, tid, store_L, 1.0, 1.0);
    return;
  }
  double* ks = &sm.colbuf[0][0][0];          // k_r of the 64 rows (free until the first column pair is published)
  if (tid < TILE) {
    const int e = min(max(ex, 1), 2046) - 1023;
    ks[tid] = (double)((e - (e & 1)) / 2);   // floor(e / 2)
  }
  bar_named(1, NCONS);
  const int kr = (int)ks[tid & (TILE - 1)];
  if (tid < TILE)
    for (int c = 0; c < TILE; ++c)
      sm.Cs[c * LDT + tid] *= __hiloint2double((1023 - kr - (int)ks[c]) << 20, 0);
  const double uL = __hiloint2double((1023 + kr) << 20, 0), uW = __hiloint2double((1023 - kr) << 20, 0);
  bar_named(1, NCONS);
  diag_factor_t<true>(sm, s0, Aij, LT, ld, tid, store_L, uL, uW);
  if (tid < TILE)                            // sm.Cs keeps L_jj for write_L_tile
    for (int c = 0; c < TILE; ++c) sm.Cs[c * LDT + tid] *= uL;
}

// Triangular-solve mode: sm.Cs holds a given lower-triangular L_jj; thread 64+k computes row k of
// L_jj^-T by substitution on the stacked [L; I] (no factorisation, no barriers: L is read-only).
__device__ __noinline__ void diag_inverse(CholSmem& sm, double* __restrict__ LT, int tid) {
  if (tid < TILE || tid >= 2 * TILE) return;
  const int k = tid - TILE;
  double row[TILE];
#pragma unroll
  for (int c = 0; c < TILE; ++c) row[c] = (k == c) ? 1.0 : 0.0;
#pragma unroll
  for (int c = 0; c < TILE; ++c) {
    const double xv = row[c] / sm.Cs[c * LDT + c];
    row[c] = xv;
#pragma unroll
    for (int c2 = c + 1; c2 < TILE; ++c2) row[c2] = fma(-xv, sm.Cs[c * LDT + c2], row[c2]);
  }
#pragma unroll
  for (int c = 0; c < TILE; ++c) LT[c * LDT + k] = (c >= k) ? row[c] : 0.0;
}

// Pre-accumulation of a diagonal block (K_PART) as a balanced SYRK, a function of its own so that the
// ordinary contraction loop keeps its code and registers.  Only fragments on or below the diagonal
// of the block are ever read (36 of 64): warp 2, whose sub-tile lies strictly above the diagonal,
// takes tile columns 16..31 of warp 1's full sub-tile (8 fragments each); warps 0 and 3 run the 10
// lower fragments of their diagonal sub-tiles -- the task lasts 10/16 of the full tile.  The same
// operations per entry in the same order as the full-tile loop (negated accumulator, k ascending):
// bit-identical results; the strict upper triangle of the block keeps what it held.
// Consumes the two tile stages and `nstage` contraction stages of the ring and releases all of them.
template <int SH>   // 1: two column blocks x four row blocks;  2: lower triangle of a 4 x 4 block grid
__device__ __forceinline__ void syrk_contract(CholSmem& sm, double* __restrict__ Aij, int ld, uint32_t& stage,
                                              uint32_t& phase, int pre0, int pre1, int nstage, int lane,
                                              int cE, int rE, int* error) {
  constexpr int MT = (SH == 1) ? 2 : 4;
  const int gq = lane >> 2, tq = lane & 3;
  auto wait_full = [&](uint32_t st, uint32_t ph) {
    if (mbar_test_wait(&sm.full[st], ph)) return;
    if (!mbar_wait(&sm.full[st], ph)) atomicExch(error, 5);
  };
  double acc[MT][4][2];
  {
    const double* T = &sm.stage[cE < 32 ? pre0 : pre1][0];   // tile columns cE .. of this warp
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
#pragma unroll
      for (int nt = 0; nt < 4; ++nt)
        if (SH != 2 || nt >= mt) {
          const double2 v = *reinterpret_cast<const double2*>(&T[((cE & 31) + mt * 8 + gq) * LDT + rE + nt * 8 + 2 * tq]);
          acc[mt][nt][0] = flip_sign(v.x); acc[mt][nt][1] = flip_sign(v.y);
        }
  }
  auto load_frag = [&](const double* P, int ks, double (&a)[MT], double (&bb)[4]) {
    const int k = ks * 4 + tq;
#pragma unroll
    for (int mt = 0; mt < MT; ++mt) a[mt] = P[k * LDT + cE + mt * 8 + gq];
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) bb[nt] = P[(KC + k) * LDT + rE + nt * 8 + gq];
  };
  int pend = -1;
  wait_full(stage, phase);
  double a[MT], bb[4];
  load_frag(&sm.stage[stage][0], 0, a, bb);
#pragma unroll 1
  for (int it = 0; it < nstage; ++it) {
    const double* P = &sm.stage[stage][0];
    uint32_t nst = stage + 1, nph = phase;
    if (nst == STAGES) { nst = 0; nph ^= 1; }
    bool next_ready = true;
#pragma unroll
    for (int ks = 0; ks < KC / 4; ++ks) {
      double an[MT], bn[4];
      if (ks == KC / 4 - 2) next_ready = (it + 1 >= nstage) || mbar_test_wait(&sm.full[nst], nph);
      if (ks + 1 < KC / 4) {
        load_frag(P, ks + 1, an, bn);
      } else if (it + 1 < nstage) {
        if (!next_ready) wait_full(nst, nph);
        load_frag(&sm.stage[nst][0], 0, an, bn);
      } else {
#pragma unroll
        for (int i = 0; i < MT; ++i) an[i] = 0.0;
#pragma unroll
        for (int i = 0; i < 4; ++i) bn[i] = 0.0;
      }
#pragma unroll
      for (int mt = 0; mt < MT; ++mt)
#pragma unroll
        for (int nt = 0; nt < 4; ++nt)
          if (SH != 2 || nt >= mt) dmma884(acc[mt][nt][0], acc[mt][nt][1], a[mt], bb[nt]);
#pragma unroll
      for (int i = 0; i < MT; ++i) a[i] = an[i];
#pragma unroll
      for (int i = 0; i < 4; ++i) bb[i] = bn[i];
    }
    __syncwarp();
    if (lane == 0) {
      // the tile stages after the first stage (their LDS fed the flips above, long since issued);
      // every other stage one stage late (see the ordinary loop)
      if (it == 0) { mbar_arrive(&sm.empty[pre0]); mbar_arrive(&sm.empty[pre1]); }
      if (pend >= 0) mbar_arrive(&sm.empty[pend]);
    }
    pend = (int)stage;
    stage = nst; phase = nph;
  }
#pragma unroll
  for (int mt = 0; mt < MT; ++mt)
#pragma unroll
    for (int nt = 0; nt < 4; ++nt)
      if (SH != 2 || nt >= mt)
        *reinterpret_cast<double2*>(Aij + (size_t)(cE + mt * 8 + gq) * ld + rE + nt * 8 + 2 * tq) =
            make_double2(flip_sign(acc[mt][nt][0]), flip_sign(acc[mt][nt][1]));
  __syncwarp();                          // the stores consumed every accumulator: all LDS of the ring are done
  if (lane == 0 && pend >= 0) mbar_arrive(&sm.empty[pend]);
}
__device__ __noinline__ void syrk_task(CholSmem& sm, double* __restrict__ Aij, int ld, uint32_t* stage_io,
                                       uint32_t* phase_io, int nstage, int tid, int* error) {
  const int warp = tid >> 5, lane = tid & 31;
  uint32_t stage = *stage_io, phase = *phase_io;
  const int pre0 = (int)stage;
  if (!mbar_test_wait(&sm.full[stage], phase) && !mbar_wait(&sm.full[stage], phase)) atomicExch(error, 5);
  if (++stage == STAGES) { stage = 0; phase ^= 1; }
  const int pre1 = (int)stage;
  if (!mbar_test_wait(&sm.full[stage], phase) && !mbar_wait(&sm.full[stage], phase)) atomicExch(error, 5);
  if (++stage == STAGES) { stage = 0; phase ^= 1; }
  // warp: 0 -> block (c 0..31, r 0..31) lower; 3 -> (32.., 32..) lower; 1 -> (c 0..15, r 32..63); 2 -> (c 16..31, r 32..63)
  if (warp == 1 || warp == 2)
    syrk_contract<1>(sm, Aij, ld, stage, phase, pre0, pre1, nstage, lane, warp == 1 ? 0 : 16, 32, error);
  else
    syrk_contract<2>(sm, Aij, ld, stage, phase, pre0, pre1, nstage, lane, warp == 0 ? 0 : 32, warp == 0 ? 0 : 32, error);
  *stage_io = stage; *phase_io = phase;
}

// Deferred publication.  A consumer group that has stored its tile does not wait for the stores
// to become visible: it posts (counter, value) and moves on to the next task.  The producer lane,
// which spends most of its time waiting anyway, drains the post: gpu-scope fence (cumulative over
// the consumers' stores, ordered before it by the CTA barrier + mbarrier), proxy fence (the
// readers are TMA loads), release store.  Every producer spin loop calls this, so a pending
// publication can never be starved.
struct RelState { int r; int h; };
__device__ __forceinline__ void service_release(CholSmem& sm, RelState& rs) {
  if (mbar_test_wait(&sm.hf_full, rs.h & 1)) {     // early half of L_jj^-T (see diag_factor_t)
    int* p = sm.hf_ptr;
    __threadfence();
    fence_proxy_async();
    st_release(p, 1);
    mbar_arrive(&sm.hf_empty);
    ++rs.h;
  }
  for (;;) {
    const int slot = rs.r & 1, par = (rs.r >> 1) & 1;
    if (!mbar_test_wait(&sm.rel_full[slot], par)) return;
    int* p = sm.rel_ptr[slot];
    const int v = sm.rel_val[slot];
    __threadfence();
    fence_proxy_async();
    st_release(p, v);
    mbar_arrive(&sm.rel_empty[slot]);
    ++rs.r;
  }
}
__device__ __forceinline__ bool mbar_wait_serving(uint64_t* bar, uint32_t parity, CholSmem& sm, RelState& rs) {
  // sleep in hardware (<= ~0.3 us per try) instead of spinning: a busy producer lane steals issue
  // slots and SYNCS bandwidth from the consumer warp that shares its scheduler
  for (uint32_t it = 0; it < (1u << 23); ++it) {
    if (mbar_try_wait_ns(bar, parity, 300)) return true;
    service_release(sm, rs);
  }
  return false;
}
// Returns the observed counter value (>= need), or -1 on error / watchdog.
__device__ __forceinline__ int wait_counter_serving(const int* ctr, int need, int* error, CholSmem& sm, RelState& rs) {
  int v = ld_acquire(ctr);
  if (v >= need) return v;
  long long t0 = clock64();
  for (;;) {
    service_release(sm, rs);
    v = ld_acquire(ctr);
    if (v >= need) return v;
    if (*(volatile int*)error) return -1;
    if (clock64() - t0 > WATCHDOG_CLOCKS) { atomicExch(error, 1); return -1; }
    __nanosleep(20);
  }
}

// Wait for the first half of L_jj^-T (rows 0..31: the running factorisation raises `half` behind its second
// panel) or, with first == false, for the whole block (the row's progress counter).  Returns 1 when the
// block is complete, 0 when only its first half is, -1 on error / watchdog.
__device__ __forceinline__ int wait_lt_serving(const int* half, bool first, const int* donej, int need, int* error,
                                               CholSmem& sm, RelState& rs) {
  long long t0 = 0;
  for (int spin = 0;; ++spin) {
    if (ld_acquire(donej) >= need) return 1;
    if (first && ld_acquire(half) >= 1) return 0;
    if (spin == 0) t0 = clock64();
    service_release(sm, rs);
    if ((spin & 15) == 15) {
      if (*(volatile int*)error) return -1;
      if (clock64() - t0 > WATCHDOG_CLOCKS) { atomicExch(error, 1); return -1; }
    }
  }
}

// TRACE = true is the schedule-analysis build of the same kernel (gpb_set_trace); the production
// instantiation carries no timestamp code or registers.
// MINB = CTAs per SM the register budget is sized for: 2 (168 registers) for throughput launches, 1
// (no register cap, no spills in the factor / solve phases) for launches bound by the dependency
// chain, which run one CTA per SM anyway (n=4096, one proposal: 1.53 -> 1.38 ms).
template <bool TRACE, int MINB>
__global__ void __launch_bounds__(NTHREADS, MINB)
k_chol_tasks(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ CUtensorMap tmapL, Geom g, double* __restrict__ mats, double* __restrict__ linvt, int* __restrict__ done,
             int* __restrict__ info, SchedState* sched, int task_begin, int task_end,
             int static_sched, int solve_only, unsigned long long* __restrict__ trace, int chain_pause) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  CholSmem& sm = *reinterpret_cast<CholSmem*>(smem_raw);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  // The features that shorten the chain diag(j) -> (j+1, j) -> diag(j+1) -- direct publication of a factored
  // block by consumer thread 0, the first half of L_jj^-T handed over while the factorisation still runs, the
  // X X^T update of the fused task as a balanced SYRK -- were specific to the chain-bound instantiation until
  // the end of round 2.  Re-measured together on the final kernels they also pay in the throughput
  // instantiation: 1024 x 32 -3.3 %, 4096 x 2 -4.9 %, 2048 x 8 -4.4 %, 1024 x 256 -0.9 %, 4096 x 8 -0.35 %,
  // 8192 x 8 +0.3 % (profiles/r02_chain_pause_sweep.txt).  On in both.
  constexpr bool DIRECT_DIAG = true;
  constexpr bool EARLY_HALF = true;

  if (tid == 0) {
    tma_prefetch_desc(&tmap);
    tma_prefetch_desc(&tmapL);
    for (int s = 0; s < STAGES; ++s) { mbar_init(&sm.full[s], 1); mbar_init(&sm.empty[s], NCONS / 32); }
    for (int s = 0; s < 2; ++s) { mbar_init(&sm.tq_full[s], 1); mbar_init(&sm.tq_empty[s], NCONS / 32); }
    for (int s = 0; s < 2; ++s) { mbar_init(&sm.rel_full[s], 1); mbar_init(&sm.rel_empty[s], 1); }
    mbar_init(&sm.hf_full, 1); mbar_init(&sm.hf_empty, 1);
    sm.hf_req = nullptr; sm.hf_ptr = nullptr; sm.hf_count = 0;
    mbar_fence_init();
    sm.pause_mine = sm.pause_other = nullptr;
    if (MINB == 2 && chain_pause) {   // the one-CTA-per-SM instantiation has no neighbour to pause
      // which of the (up to two) CTAs of this SM am I?  Arrival order on a per-SM counter.
      const unsigned sid = smid() % SCHED_MAXSM;
      const int slot = atomicAdd(&sched->smslot[sid], 1) & 1;
      sm.pause_mine = &sched->pause[2 * sid + slot];
      sm.pause_other = &sched->pause[2 * sid + (slot ^ 1)];
    }
  }
  __syncthreads();

  uint32_t stage = 0, phase = 0;
  int jcache = 0;

  if (warp == NCONS / 32) {
    // =============================== producer warp ===============================
    int q = 0, static_next = task_begin + blockIdx.x;
    RelState rs; rs.r = 0; rs.h = 0;
    int posted_h = 0; // early halves the consumers will post
    int posted = 0;   // tasks handed to the consumers that will post a publication
    // The next ticket is drawn when every load of the current task has been issued (its last stages
    // are still in flight then, so the atomic's latency stays hidden).  The later, the better for
    // launches that are partly bound by the dependency chain: a busy CTA that draws early sits on a
    // ticket of the critical path.  Measured: at the start of the task n=1024 batch 32 0.64 -> 0.72 ms;
    // before the solve stages (the previous place) 0.65 ms; behind them 0.62 ms (n=300 batch 32
    // 0.124 -> 0.101 ms, n=1024 batch 16 0.435 -> 0.404 ms; throughput launches unchanged).  Waiting
    // further, until the consumers have posted the task's publication, is slower again
    // (n=1024 batch 16 0.443 ms): the next task's own tile then arrives late.
    auto fetch = [&]() {
      int tn;
      if (static_sched) { tn = static_next; static_next += gridDim.x; }
      else tn = task_begin + atomicAdd(&sched->next_task, 1);
      return tn;   // raw: clamped at the loop top
    };
    int t_next = (lane == 0) ? fetch() : 0;
    // chain_pause: do not feed this CTA's contraction while the CTA that shares the SM is inside a
    // critical section (advisory: bounded, and publications keep being served)
    int* const pause_mine = (MINB == 2) ? sm.pause_mine : nullptr;
    int* const pause_other = (MINB == 2) ? sm.pause_other : nullptr;
    auto hold_while_paused = [&]() {
      if (MINB != 2 || pause_other == nullptr || *(volatile int*)pause_other <= 0) return;
      const long long t0 = clock64();
      while (*(volatile int*)pause_other > 0) {
        service_release(sm, rs);
        if (clock64() - t0 > PAUSE_MAX_CLOCKS) break;
        __nanosleep(200);
      }
    };
    for (;;) {
      int t = __shfl_sync(0xffffffffu, t_next, 0);
      if (t >= task_end) t = -1;
      const int slot = q & 1, qpar = (q >> 1) & 1;
      TaskDesc d;
      d.b = d.i = d.j = 0; d.kind = K_PLAIN;
      if (t >= 0) d = decode_task(g, t, jcache);
      if (lane == 0) {
        if (!mbar_wait_serving(&sm.tq_empty[slot], qpar ^ 1, sm, rs)) atomicExch(&sched->error, 2);
        sm.taskq[slot] = t;
        sm.taskd[slot] = make_int4(d.b, d.i, d.j, d.kind);
        mbar_arrive(&sm.tq_full[slot]);
      }
      ++q;
      if (t < 0) break;
      const int k0 = task_k0(g, d);
      const int kind = solve_only ? K_PLAIN : d.kind;
      // L is given: no accumulation for top-square tiles; void task; nothing to pre-accumulate for block 1
      if ((solve_only && d.i < g.nt) || k0 < 0 || (kind == K_PART && d.j == 1)) {
        if (lane == 0) t_next = fetch();
        continue;
      }
      // (the chain-bound instantiation publishes diagonal blocks itself, see factor_block)
      if (!solve_only) posted += (DIRECT_DIAG) ? ((kind == K_FUSED || kind == K_PART || (kind == K_PLAIN && d.i != d.j)) ? 1 : 0)
                                             : ((kind == K_FUSED) ? 2 : 1);
      if (EARLY_HALF && !solve_only && (kind == K_FUSED || kind == K_DIAG0)) posted_h += 1;
      const int kend = (kind == K_PART) ? d.j - 1 : d.j;   // K_PART leaves the last block column to K_FUSED
      const int* done_b = done + (size_t)d.b * g.rt;
      const int* part_b = done + (size_t)g.batch * g.rt + (size_t)d.b * g.nt;
      unsigned long long* ptr_ = (TRACE && trace != nullptr && lane == 0) ? trace + (size_t)t * TRACE_SLOTS : nullptr;
      if (ptr_) ptr_[10] = clock64();
      // scratch stages of a diagonal factorisation: they start as the identity (the stacked [C; I]),
      // fetched by TMA from the identity tile behind linvt while the consumers are still updating C
      auto load_identity = [&]() {
        const int row0 = g.batch * g.nt * TILE;
#pragma unroll 1
        for (int s = 0; s < LT_STAGES; ++s) {
          if (!mbar_wait_serving(&sm.empty[stage], phase ^ 1, sm, rs)) atomicExch(&sched->error, 3);
          mbar_arrive_expect_tx(&sm.full[stage], STAGE_TX_BYTES);
          tma_load_2d(&sm.stage[stage][0], &tmapL, 0, row0 + s * SCOLS, &sm.full[stage]);
          tma_load_2d(&sm.stage[stage][KC * LDT], &tmapL, 0, row0 + s * SCOLS + KC, &sm.full[stage]);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      };
      if (lane == 0) {
        // the tile itself (accumulator start value) comes first, as AT_STAGES ordinary stages:
        // stage s holds tile columns s*SCOLS .. in [column][LDT] order
#pragma unroll 1
        for (int s = 0; s < AT_STAGES; ++s) {
          if (!mbar_wait_serving(&sm.empty[stage], phase ^ 1, sm, rs)) atomicExch(&sched->error, 3);
          mbar_arrive_expect_tx(&sm.full[stage], STAGE_TX_BYTES);
          const int col = d.j * TILE + s * SCOLS;
          tma_load_3d(&sm.stage[stage][0], &tmap, d.i * TILE, col, d.b, &sm.full[stage]);
          tma_load_3d(&sm.stage[stage][KC * LDT], &tmap, d.i * TILE, col + KC, d.b, &sm.full[stage]);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        if (ptr_) ptr_[11] = clock64();
        // one elected lane drives the ring: two TMA boxes per stage ({LDT rows x KC cols}: the
        // 4 extra rows give the padded, bank-conflict-free smem pitch for free)
        // progress counters only grow: poll again only when the last value seen is not enough
        int seen_i = 0, seen_j = 0;
        for (int kc = k0; kc < kend; ++kc) {
          bool polled = false;
          if (seen_i < kc + 1) { seen_i = wait_counter_serving(done_b + d.i, kc + 1, &sched->error, sm, rs); polled = true; }
          if (seen_j < kc + 1) { seen_j = wait_counter_serving(done_b + d.j, kc + 1, &sched->error, sm, rs); polled = true; }
          if (seen_i < 0 || seen_j < 0) { seen_i = seen_j = 1 << 30; }   // error: drain without waiting
          if (polled) fence_proxy_async();
#pragma unroll 1
          for (int s = 0; s < TILE / KC; ++s) {
            hold_while_paused();
            if (!mbar_wait_serving(&sm.empty[stage], phase ^ 1, sm, rs)) atomicExch(&sched->error, 3);
            const int col = kc * TILE + s * KC;
            mbar_arrive_expect_tx(&sm.full[stage], STAGE_TX_BYTES);
            tma_load_3d(&sm.stage[stage][0], &tmap, d.j * TILE, col, d.b, &sm.full[stage]);
            tma_load_3d(&sm.stage[stage][KC * LDT], &tmap, d.i * TILE, col, d.b, &sm.full[stage]);
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
          }
        }
        if (ptr_) ptr_[12] = clock64();
        if (kind == K_PART) {
          // nothing more to stream: the consumers store the partial sum back in place
        } else if (d.i != d.j) {
          // the solve X = C L_jj^-T streams L_jj^-T (pitch LDT in global memory) through the same
          // ring: LT_STAGES stages of SCOLS rows each, as soon as diag(j) is published
          const int row0 = (d.b * g.nt + d.j) * TILE;
          int whole = 1;
          if constexpr (EARLY_HALF) {
            whole = (seen_j >= d.j + 1) ? 1 : 0;
          } else {
            if (seen_j < d.j + 1) wait_counter_serving(done_b + d.j, d.j + 1, &sched->error, sm, rs);
            fence_proxy_async();
          }
#pragma unroll 1
          for (int s = 0; s < LT_STAGES; ++s) {
            if constexpr (EARLY_HALF) {
              // stage 0 (rows 0..31) goes out as soon as the running factorisation has released its first half
              if (whole == 0) {
                whole = wait_lt_serving(done + (size_t)g.batch * (g.rt + g.nt) + (size_t)d.b * g.nt + d.j, s == 0,
                                        done_b + d.j, d.j + 1, &sched->error, sm, rs);
                fence_proxy_async();
              } else if (s == 0) {
                fence_proxy_async();
              }
            }
            if (!mbar_wait_serving(&sm.empty[stage], phase ^ 1, sm, rs)) atomicExch(&sched->error, 3);
            mbar_arrive_expect_tx(&sm.full[stage], STAGE_TX_BYTES);
            tma_load_2d(&sm.stage[stage][0], &tmapL, 0, row0 + s * SCOLS, &sm.full[stage]);
            tma_load_2d(&sm.stage[stage][KC * LDT], &tmapL, 0, row0 + s * SCOLS + KC, &sm.full[stage]);
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
          }
          if (kind == K_FUSED) {
            // ... followed by the pre-accumulated diagonal block (d.i, d.i) and the scratch stages of
            // its factorisation
            if (d.j > 0) wait_counter_serving(part_b + d.i, 1, &sched->error, sm, rs);
            fence_proxy_async();
            // everything the critical section needs exists: the SM is this CTA's until the block is factored
            if (MINB == 2 && pause_mine != nullptr) atomicAdd(pause_mine, 1);
#pragma unroll 1
            for (int s = 0; s < AT_STAGES; ++s) {
              if (!mbar_wait_serving(&sm.empty[stage], phase ^ 1, sm, rs)) atomicExch(&sched->error, 3);
              mbar_arrive_expect_tx(&sm.full[stage], STAGE_TX_BYTES);
              const int col = d.i * TILE + s * SCOLS;
              tma_load_3d(&sm.stage[stage][0], &tmap, d.i * TILE, col, d.b, &sm.full[stage]);
              tma_load_3d(&sm.stage[stage][KC * LDT], &tmap, d.i * TILE, col + KC, d.b, &sm.full[stage]);
              if (++stage == STAGES) { stage = 0; phase ^= 1; }
            }
            load_identity();
          }
        } else if (!solve_only) {
          // diagonal task: LT_STAGES stages as scratch for the stacked inverse rows
          if (MINB == 2 && pause_mine != nullptr) atomicAdd(pause_mine, 1);
          load_identity();
        }
        // the next ticket is drawn only now, behind every load of this task: see `fetch` above
        t_next = fetch();
        if (ptr_) ptr_[13] = clock64();
      }
      __syncwarp();
    }
    if (lane == 0) {
      // drain: every task this CTA took must be published before the producer leaves
      const long long t0 = clock64();
      while (rs.r < posted || rs.h < posted_h) {
        service_release(sm, rs);
        if (clock64() - t0 > WATCHDOG_CLOCKS) { atomicExch(&sched->error, 6); break; }
      }
    }
    return;
  }

  // ================================= consumer warps =================================
  const int gq = lane >> 2, tq = lane & 3;          // DMMA fragment coordinates
  const int wc = warp / (CW / 2), wr = warp % (CW / 2);   // warp's sub-tile: 32 columns c x RW rows r
  int q = 0, relc = 0;
  // post (counter, value) for the producer lane to publish; all consumer stores are issued
  auto post_release = [&](int* ptr, int val) {
    const int rsl = relc & 1, rpar = (relc >> 1) & 1;
    if (tid == 0) {
      if (!mbar_wait(&sm.rel_empty[rsl], rpar ^ 1)) atomicExch(&sched->error, 7);
      sm.rel_ptr[rsl] = ptr; sm.rel_val[rsl] = val;
    }
    bar_named(1, NCONS);               // every consumer's stores are ordered before the post
    if (tid == 0) mbar_arrive(&sm.rel_full[rsl]);
    ++relc;
  };
  for (;;) {
    const int slot = q & 1, qpar = (q >> 1) & 1;
    if (!mbar_wait(&sm.tq_full[slot], qpar)) atomicExch(&sched->error, 4);
    const int t = sm.taskq[slot];
    const int4 td = sm.taskd[slot];
    // one arrival per warp, no CTA barrier: the end of the previous task (post_release / factor_block)
    // already held every warp until all of them were done with Cs
    __syncwarp();
    if (lane == 0) mbar_arrive(&sm.tq_empty[slot]);
    ++q;
    if (t < 0) break;
    unsigned long long* tr = (TRACE && trace != nullptr && tid == 0) ? trace + (size_t)t * TRACE_SLOTS : nullptr;
    if (tr) { tr[0] = globaltimer_ns(); tr[1] = clock64(); }
    TaskDesc d;
    d.b = td.x; d.i = td.y; d.j = td.z; d.kind = td.w;   // decoded by the producer warp
    const int k0 = task_k0(g, d);
    const int kind = solve_only ? K_PLAIN : d.kind;
    if (k0 < 0 || (kind == K_PART && d.j == 1)) continue;
    const int kend = (kind == K_PART) ? d.j - 1 : d.j;
    // Tile row with at most 8 rows in use (the rhs row of the likelihood): one 8-row DMMA fragment per
    // 8 tile columns instead of the whole 64-row tile -- the same operations on the live entries, in
    // the same order, so the results do not change; the dead rows are zero and stay untouched.
    const bool light = (d.i == g.light_row);
    double* M = mats + (size_t)d.b * g.mat_stride;
    int* done_b = done + (size_t)d.b * g.rt;
    int* part_b = done + (size_t)g.batch * g.rt + (size_t)d.b * g.nt;
    double* Aij = M + (size_t)(d.j * TILE) * g.ld + (size_t)d.i * TILE;  // (r, c) at Aij[c*ld + r]
    if (solve_only && d.i < g.nt && d.i != d.j) {
      // triangular-solve mode: the factor is an input, its off-diagonal tiles are already final
      if (tid == 0) atomicMax(done_b + d.i, d.j + 1);
      continue;
    }

    if (CW == 4 && kind == K_PART) {
      // diagonal block pre-accumulated as a balanced SYRK (syrk_task), back in place, flag posted
      if (tr) tr[2] = clock64();
      syrk_task(sm, Aij, g.ld, &stage, &phase, (kend - k0) * (TILE / KC), tid, &sched->error);
      if (tr) { tr[3] = tr[4] = tr[5] = clock64(); tr[8] = tr[9] = 0; }
      post_release(part_b + d.i, 1);
      if (tr) { tr[6] = clock64(); tr[7] = globaltimer_ns(); }
      continue;
    }

    int not_ready = 0;
    // fast path: the stage is almost always there already (test_wait is a plain probe; try_wait
    // costs ~100 cycles even on a completed phase)
    auto wait_full = [&](uint32_t st, uint32_t ph) {
      if (mbar_test_wait(&sm.full[st], ph)) return;
      ++not_ready;
      if (!mbar_wait(&sm.full[st], ph)) atomicExch(&sched->error, 5);
    };
    // acc[mt][nt] holds D[m = c][n = r]:  c = wc*32 + mt*8 + gq,  r = wr*RW + nt*8 + 2*tq + {0,1}
    double acc[4][NTW][2];
    int pre0 = -1, pre1 = -1;   // ring stages that carried the tile; released once acc has been consumed
    if (solve_only && d.i < g.nt) {
      // given factor (diagonal tile of the triangular-solve mode): straight from global memory
#pragma unroll
      for (int mt = 0; mt < 4; ++mt)
#pragma unroll
        for (int nt = 0; nt < NTW; ++nt) {
          const int c = wc * 32 + mt * 8 + gq, r = wr * RW + nt * 8 + 2 * tq;
          double2 v = ld_cg2(Aij + (size_t)c * g.ld + r);
          acc[mt][nt][0] = v.x; acc[mt][nt][1] = v.y;
        }
    } else {
      pre0 = (int)stage;
      wait_full(stage, phase);
      if (++stage == STAGES) { stage = 0; phase ^= 1; }
      pre1 = (int)stage;
      wait_full(stage, phase);
      if (++stage == STAGES) { stage = 0; phase ^= 1; }
      if (tr) tr[2] = clock64();
      const double* T = &sm.stage[wc == 0 ? pre0 : pre1][0];   // this warp's 32 tile columns
#pragma unroll
      for (int mt = 0; mt < 4; ++mt)
#pragma unroll
        for (int nt = 0; nt < NTW; ++nt) {
          const double2 v = *reinterpret_cast<const double2*>(&T[(mt * 8 + gq) * LDT + wr * RW + nt * 8 + 2 * tq]);
          acc[mt][nt][0] = v.x; acc[mt][nt][1] = v.y;
        }
    }
    // release of the tile stages: only after something that needed acc has issued (LDS done)
    auto release_pre = [&]() {
      if (pre0 >= 0) {
        if (lane == 0) { mbar_arrive(&sm.empty[pre0]); mbar_arrive(&sm.empty[pre1]); }
        pre0 = pre1 = -1;
      }
    };

    const int nstage = (solve_only && d.i < g.nt) ? 0 : (kend - k0) * (TILE / KC);
    // Stage release is DELAYED by one stage: ptxas schedules mbarrier.arrive right behind the last
    // LDS of a stage without waiting for those loads to return, and with two CTAs per SM the
    // refill (async proxy) was observed to overtake them.  Releasing stage s only while stage s+1
    // is being consumed guarantees every DMMA that read s has issued, i.e. its LDS completed.
    int pend = -1;
    // The contraction runs on the NEGATED accumulator, acc' = -A + sum L L^T, so that the fragments
    // enter the DMMAs as loaded: negating them at the stage boundary put four DADDs that wait on the
    // LDS in front of the next 16 DMMAs.  Rounding is sign-symmetric, so -acc' is bit for bit the
    // A - sum L L^T of the direct form (only the sign of an exact zero can differ).
    auto load_frag = [&](const double* P, int ks, double (&a)[4], double (&bb)[NTW]) {
      const int k = ks * 4 + tq;
#pragma unroll
      for (int mt = 0; mt < 4; ++mt) a[mt] = P[k * LDT + wc * 32 + mt * 8 + gq];
#pragma unroll
      for (int nt = 0; nt < NTW; ++nt) bb[nt] = P[(KC + k) * LDT + wr * RW + nt * 8 + gq];
    };
    if (light) {
      // the tile goes to Cs as it is (dead rows: zeros); the live fragment (rows 0..7, warps wr == 0)
      // stays in registers and is written over it after the contraction
      double la[4][2];
#pragma unroll
      for (int mt = 0; mt < 4; ++mt) {
        la[mt][0] = acc[mt][0][0]; la[mt][1] = acc[mt][0][1];
#pragma unroll
        for (int nt = 0; nt < NTW; ++nt) {
          const int c = wc * 32 + mt * 8 + gq, r = wr * RW + nt * 8 + 2 * tq;
          *reinterpret_cast<double2*>(&sm.Cs[c * LDT + r]) = make_double2(acc[mt][nt][0], acc[mt][nt][1]);
        }
      }
      __syncwarp();                      // the stores consumed the tile: its LDS are done
      release_pre();
      auto light_frag = [&](const double* P, double (&a)[KC / 4][4], double (&b)[KC / 4]) {
#pragma unroll
        for (int ks = 0; ks < KC / 4; ++ks) {
          const int k = ks * 4 + tq;
#pragma unroll
          for (int mt = 0; mt < 4; ++mt) a[ks][mt] = -P[k * LDT + wc * 32 + mt * 8 + gq];
          b[ks] = P[(KC + k) * LDT + gq];
        }
      };
      if (nstage > 0 && wr == 0) {
        wait_full(stage, phase);
        double a[KC / 4][4], b[KC / 4];
        light_frag(&sm.stage[stage][0], a, b);
#pragma unroll 1
        for (int it = 0; it < nstage; ++it) {
          uint32_t nst = stage + 1, nph = phase;
          if (nst == STAGES) { nst = 0; nph ^= 1; }
          double an[KC / 4][4], bn[KC / 4];
          if (it + 1 < nstage) {
            wait_full(nst, nph);
            light_frag(&sm.stage[nst][0], an, bn);
          } else {
#pragma unroll
            for (int ks = 0; ks < KC / 4; ++ks) {
              bn[ks] = 0.0;
#pragma unroll
              for (int mt = 0; mt < 4; ++mt) an[ks][mt] = 0.0;
            }
          }
#pragma unroll
          for (int ks = 0; ks < KC / 4; ++ks)
#pragma unroll
            for (int mt = 0; mt < 4; ++mt) dmma884(la[mt][0], la[mt][1], a[ks][mt], b[ks]);
#pragma unroll
          for (int ks = 0; ks < KC / 4; ++ks) {
            b[ks] = bn[ks];
#pragma unroll
            for (int mt = 0; mt < 4; ++mt) a[ks][mt] = an[ks][mt];
          }
          // this stage's fragments were fetched one iteration ago and its DMMAs have issued
          __syncwarp();
          if (lane == 0) mbar_arrive(&sm.empty[stage]);
          stage = nst; phase = nph;
        }
#pragma unroll
        for (int mt = 0; mt < 4; ++mt)
          *reinterpret_cast<double2*>(&sm.Cs[(wc * 32 + mt * 8 + gq) * LDT + 2 * tq]) = make_double2(la[mt][0], la[mt][1]);
      } else {
        // the warps of the dead rows only keep the ring turning
#pragma unroll 1
        for (int it = 0; it < nstage; ++it) {
          wait_full(stage, phase);
          if (lane == 0) mbar_arrive(&sm.empty[stage]);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    } else if (nstage > 0) {
#pragma unroll
      for (int mt = 0; mt < 4; ++mt)
#pragma unroll
        for (int nt = 0; nt < NTW; ++nt) { acc[mt][nt][0] = flip_sign(acc[mt][nt][0]); acc[mt][nt][1] = flip_sign(acc[mt][nt][1]); }
      wait_full(stage, phase);
      double a[4], bb[NTW];
      load_frag(&sm.stage[stage][0], 0, a, bb);
#pragma unroll 1
      for (int it = 0; it < nstage; ++it) {
        const double* P = &sm.stage[stage][0];
        uint32_t nst = stage + 1, nph = phase;
        if (nst == STAGES) { nst = 0; nph ^= 1; }
        bool next_ready = true;
#pragma unroll
        for (int ks = 0; ks < KC / 4; ++ks) {
          // fragments of the next k-step (or of the next stage's first k-step) are fetched before
          // this step's 16 DMMAs issue, so neither the LDS latency nor the barrier probe is exposed
          double an[4], bn[NTW];
          // the barrier of the next stage is probed one k-step before its result is needed
          if (ks == KC / 4 - 2) next_ready = (it + 1 >= nstage) || mbar_test_wait(&sm.full[nst], nph);
          if (ks + 1 < KC / 4) {
            load_frag(P, ks + 1, an, bn);
          } else if (it + 1 < nstage) {
            if (!next_ready) wait_full(nst, nph);
            load_frag(&sm.stage[nst][0], 0, an, bn);
          } else {
#pragma unroll
            for (int i = 0; i < 4; ++i) an[i] = 0.0;
#pragma unroll
            for (int i = 0; i < NTW; ++i) bn[i] = 0.0;
          }
#pragma unroll
          for (int mt = 0; mt < 4; ++mt)
#pragma unroll
            for (int nt = 0; nt < NTW; ++nt) dmma884(acc[mt][nt][0], acc[mt][nt][1], a[mt], bb[nt]);
#pragma unroll
          for (int i = 0; i < 4; ++i) a[i] = an[i];
#pragma unroll
          for (int i = 0; i < NTW; ++i) bb[i] = bn[i];
        }
        __syncwarp();
        release_pre();
        if (lane == 0 && pend >= 0) mbar_arrive(&sm.empty[pend]);
        pend = (int)stage;
        stage = nst; phase = nph;
      }
#pragma unroll
      for (int mt = 0; mt < 4; ++mt)
#pragma unroll
        for (int nt = 0; nt < NTW; ++nt) { acc[mt][nt][0] = flip_sign(acc[mt][nt][0]); acc[mt][nt][1] = flip_sign(acc[mt][nt][1]); }
    }

    if (tr) { tr[3] = clock64(); tr[8] = (unsigned long long)not_ready; tr[9] = 0; }
    if (kind == K_PART) {
      // ---------- pre-accumulated diagonal block: back to its place, K_FUSED takes it from there ----------
#pragma unroll
      for (int mt = 0; mt < 4; ++mt)
#pragma unroll
        for (int nt = 0; nt < NTW; ++nt) {
          const int c = wc * 32 + mt * 8 + gq, r = wr * RW + nt * 8 + 2 * tq;
          *reinterpret_cast<double2*>(Aij + (size_t)c * g.ld + r) = make_double2(acc[mt][nt][0], acc[mt][nt][1]);
        }
      __syncwarp();                      // the stores consumed every accumulator: all LDS of the ring are done
      release_pre();
      if (lane == 0 && pend >= 0) mbar_arrive(&sm.empty[pend]);
      if (tr) { tr[4] = tr[5] = clock64(); }
      post_release(part_b + d.i, 1);
      if (tr) { tr[6] = clock64(); tr[7] = globaltimer_ns(); }
      continue;
    }
    // acc -> Cs (column-major [c][r])
    if (!light) {
#pragma unroll
      for (int mt = 0; mt < 4; ++mt)
#pragma unroll
        for (int nt = 0; nt < NTW; ++nt) {
          const int c = wc * 32 + mt * 8 + gq, r = wr * RW + nt * 8 + 2 * tq;
          *reinterpret_cast<double2*>(&sm.Cs[c * LDT + r]) = make_double2(acc[mt][nt][0], acc[mt][nt][1]);
        }
    }
    if (tid == 0) sm.bad_col = -1;
    bar_named(1, NCONS);
    release_pre();
    if (lane == 0 && pend >= 0) mbar_arrive(&sm.empty[pend]);  // accumulators are stored: all reads done

    // Cs holds a diagonal block: factor it (diag_factor: thread = one of the 128 rows of the stacked
    // [Cs; I]) in LT_STAGES scratch stages the producer grants, report a bad pivot, publish the block.
    auto factor_block = [&](int dj, double* Ajj, double* LTj) {
      const int s0 = (int)stage;
#pragma unroll 1
      for (int s = 0; s < LT_STAGES; ++s) {
        wait_full(stage, phase);
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
      if constexpr (EARLY_HALF) {
        if (tid == 0) sm.hf_req = done + (size_t)g.batch * (g.rt + g.nt) + (size_t)d.b * g.nt + dj;   // read behind diag_factor's first barrier
      }
      if (tr) tr[14] = clock64();
      diag_factor(sm, s0, Ajj, LTj, g.ld, tid, !DIRECT_DIAG);
      if (tr) tr[15] = clock64();
      bar_named(1, NCONS);             // bad_col is final; every read of the scratch stages is done
      if (tid == 0 && sm.bad_col >= 0) atomicCAS(&info[d.b], 0, dj * TILE + sm.bad_col + 1);  // 1-based pivot
      if (DIRECT_DIAG) {
        // chain-bound launches: the next block column waits for exactly this store, so thread 0
        // publishes it right here instead of handing it to the (possibly sleeping) producer lane;
        // L_jj itself (an output only) goes out behind the publication
        if (tid == 0) { __threadfence(); fence_proxy_async(); st_release(done_b + dj, dj + 1); }
        write_L_tile(sm, Ajj, g.ld, tid);
        bar_named(1, NCONS);           // every warp is done with Cs before the next task may write it
      } else {
        post_release(done_b + dj, dj + 1);
      }
      if (MINB == 2 && tid == 0 && sm.pause_mine != nullptr) atomicSub(sm.pause_mine, 1);   // end of the critical section
      fence_proxy_async();             // generic writes to the scratch stages before TMA refills them
      __syncwarp();
      if (lane == 0)
        for (int s = 0; s < LT_STAGES; ++s) mbar_arrive(&sm.empty[(s0 + s) % STAGES]);
    };

    if (d.i == d.j) {
      // ---------- diagonal task (block 0, or the given factor of the triangular-solve mode) ----------
      double* LTj = linvt + ((size_t)d.b * g.nt + d.j) * (TILE * LDT);
      if (tr) tr[4] = clock64();
      if (solve_only) {
        diag_inverse(sm, LTj, tid);
        if (tr) tr[5] = clock64();
        __threadfence();
        fence_proxy_async();
        bar_named(1, NCONS);
        if (tid == 0) atomicMax(done_b + d.j, d.j + 1);
      } else {
        factor_block(d.j, Aij, LTj);
        if (tr) tr[5] = clock64();
      }
      if (tr) { tr[6] = clock64(); tr[7] = globaltimer_ns(); }
    } else {
      // ---------- off-diagonal task: X = C * L_jj^-T, L_jj^-T streamed through the ring ----------
      // stage s carries rows c = s*SCOLS .. of L_jj^-T; the warps solve those columns of X and store
      // them straight to global memory
      // warp w: rows r = (w % 4) * 16 .. +16, and the (w / 4)-th share of the stage's column tiles
      constexpr int CT = SCOLS / 8;          // 8-column tiles of X per stage
      constexpr int CTW = CT / (CW / 4);     // ... per warp
      const int rq = warp & 3, ch = warp >> 2;
      double xs[LT_STAGES][CTW][2][2];       // X stays in registers: K_FUSED needs it again (s loop unrolled)
      if (light) {
        // rows 0..7 only: warp w solves the w-th 8-column tile of each stage (same k order as below)
#pragma unroll 1
        for (int s = 0; s < LT_STAGES; ++s) {
          wait_full(stage, phase);
          if (tr && s == 0) tr[4] = clock64();
          if (warp < CT) {
            const double* LTs = &sm.stage[stage][warp * 8 * LDT];
            double x0 = 0.0, x1 = 0.0;
            const int kmax = s * SCOLS + (warp + 1) * 8;
#pragma unroll 4
            for (int kb = 0; kb < kmax; kb += 4) {
              const int k = kb + tq;
              dmma884(x0, x1, LTs[gq * LDT + k], sm.Cs[k * LDT + gq]);
            }
            const int c = s * SCOLS + warp * 8 + gq;
            *reinterpret_cast<double2*>(Aij + (size_t)c * g.ld + 2 * tq) = make_double2(x0, x1);
          }
          __syncwarp();
          if (lane == 0) mbar_arrive(&sm.empty[stage]);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      } else
#pragma unroll
      for (int s = 0; s < LT_STAGES; ++s) {
        wait_full(stage, phase);
        if (tr && s == 0) tr[4] = clock64();
        const double* LTs = &sm.stage[stage][ch * CTW * 8 * LDT];
        double (&x)[CTW][2][2] = xs[s];
#pragma unroll
        for (int ct = 0; ct < CTW; ++ct)
#pragma unroll
          for (int nt = 0; nt < 2; ++nt) x[ct][nt][0] = x[ct][nt][1] = 0.0;
        const int kmax = s * SCOLS + (ch + 1) * CTW * 8;   // L^-T(k, c) == 0 for k > c
        const int kfull = kmax - (CTW - 1) * 8;   // below this every column tile of the warp is live
#pragma unroll 4
        for (int kb = 0; kb < kfull; kb += 4) {
          const int k = kb + tq;
          double a[CTW], bb[2];
#pragma unroll
          for (int ct = 0; ct < CTW; ++ct) a[ct] = LTs[(ct * 8 + gq) * LDT + k];
#pragma unroll
          for (int nt = 0; nt < 2; ++nt) bb[nt] = sm.Cs[k * LDT + rq * 16 + nt * 8 + gq];
#pragma unroll
          for (int ct = 0; ct < CTW; ++ct)
#pragma unroll
            for (int nt = 0; nt < 2; ++nt) dmma884(x[ct][nt][0], x[ct][nt][1], a[ct], bb[nt]);
        }
        // triangular tail: column tile ct only has k < kfull + ct*8 (L^-T is lower triangular)
#pragma unroll
        for (int t8 = 1; t8 < CTW; ++t8) {
#pragma unroll
          for (int kk = 0; kk < 8; kk += 4) {
            const int k = kfull + (t8 - 1) * 8 + kk + tq;
            double bb[2];
#pragma unroll
            for (int nt = 0; nt < 2; ++nt) bb[nt] = sm.Cs[k * LDT + rq * 16 + nt * 8 + gq];
#pragma unroll
            for (int ct = t8; ct < CTW; ++ct) {
              const double a = LTs[(ct * 8 + gq) * LDT + k];
#pragma unroll
              for (int nt = 0; nt < 2; ++nt) dmma884(x[ct][nt][0], x[ct][nt][1], a, bb[nt]);
            }
          }
        }
#pragma unroll
        for (int ct = 0; ct < CTW; ++ct)
#pragma unroll
          for (int nt = 0; nt < 2; ++nt) {
            const int c = s * SCOLS + (ch * CTW + ct) * 8 + gq, r = rq * 16 + nt * 8 + 2 * tq;
            *reinterpret_cast<double2*>(Aij + (size_t)c * g.ld + r) = make_double2(x[ct][nt][0], x[ct][nt][1]);
          }
        // the stores above consumed every DMMA of this stage, so its LDS have completed
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.empty[stage]);
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
      if (tr) tr[5] = clock64();
      if (solve_only) {
        __threadfence();
        fence_proxy_async();
        bar_named(1, NCONS);
        if (tid == 0) st_release(done_b + d.i, d.j + 1);
      } else {
        // X is published; the barrier inside also says every warp is done with Cs (as C) and with L_jj^-T
        post_release(done_b + d.i, d.j + 1);
      }
      if (kind == K_FUSED) {
        // ---------- ... and straight on to diag(d.i): block -= X X^T from shared memory, factor ----------
        // X -> Cs in the panel layout of one 64-column chunk ([k = c][r], pitch LDT)
#pragma unroll
        for (int s = 0; s < LT_STAGES; ++s)
#pragma unroll
          for (int ct = 0; ct < CTW; ++ct)
#pragma unroll
            for (int nt = 0; nt < 2; ++nt) {
              const int c = s * SCOLS + (ch * CTW + ct) * 8 + gq, r = rq * 16 + nt * 8 + 2 * tq;
              *reinterpret_cast<double2*>(&sm.Cs[c * LDT + r]) = make_double2(xs[s][ct][nt][0], xs[s][ct][nt][1]);
            }
        // the pre-accumulated block arrives like every task's own tile
        pre0 = (int)stage;
        wait_full(stage, phase);
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
        pre1 = (int)stage;
        wait_full(stage, phase);
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
        if constexpr (CW == 4) {
          // This update sits on the dependency chain, and it is a SYRK --
          // only fragments on or below the diagonal of the block are ever read (diag_factor works on
          // rows >= columns; what Cs holds above the diagonal never reaches a result).  Warp 2, whose
          // sub-tile lies strictly above the diagonal, takes tile columns 16..31 of warp 1's full
          // sub-tile (8 fragments each); warps 0 and 3 run the 10 lower fragments of theirs: 10/16
          // of the time, the same operations per entry in the same order.
          const bool halfw = (warp == 1 || warp == 2);
          const int wcE = (warp == 2) ? 0 : wc, wrE = (warp == 2) ? 1 : wr, cofs = (warp == 2) ? 16 : 0;
          const double* T = &sm.stage[wcE == 0 ? pre0 : pre1][0];
#pragma unroll
          for (int mt = 0; mt < 4; ++mt)
#pragma unroll
            for (int nt = 0; nt < NTW; ++nt) {
              if (halfw ? mt >= 2 : nt < mt) continue;
              const double2 v = *reinterpret_cast<const double2*>(&T[(cofs + mt * 8 + gq) * LDT + wrE * RW + nt * 8 + 2 * tq]);
              acc[mt][nt][0] = v.x; acc[mt][nt][1] = v.y;
            }
          bar_named(1, NCONS);             // X is complete in Cs
          if (halfw) {
#pragma unroll 4
            for (int ks = 0; ks < TILE / 4; ++ks) {
              const int k = ks * 4 + tq;
              double a[2], bb[NTW];
#pragma unroll
              for (int mt = 0; mt < 2; ++mt) a[mt] = -sm.Cs[k * LDT + cofs + mt * 8 + gq];
#pragma unroll
              for (int nt = 0; nt < NTW; ++nt) bb[nt] = sm.Cs[k * LDT + RW + nt * 8 + gq];
#pragma unroll
              for (int mt = 0; mt < 2; ++mt)
#pragma unroll
                for (int nt = 0; nt < NTW; ++nt) dmma884(acc[mt][nt][0], acc[mt][nt][1], a[mt], bb[nt]);
            }
          } else {
#pragma unroll 4
            for (int ks = 0; ks < TILE / 4; ++ks) {
              const int k = ks * 4 + tq;
              double a[4], bb[NTW];
#pragma unroll
              for (int mt = 0; mt < 4; ++mt) a[mt] = -sm.Cs[k * LDT + wc * 32 + mt * 8 + gq];
#pragma unroll
              for (int nt = 0; nt < NTW; ++nt) bb[nt] = sm.Cs[k * LDT + wr * RW + nt * 8 + gq];
#pragma unroll
              for (int mt = 0; mt < 4; ++mt)
#pragma unroll
                for (int nt = mt; nt < NTW; ++nt) dmma884(acc[mt][nt][0], acc[mt][nt][1], a[mt], bb[nt]);
            }
          }
          __syncwarp();
          release_pre();
          bar_named(1, NCONS);             // every warp has read X: Cs may take the block
#pragma unroll
          for (int mt = 0; mt < 4; ++mt)
#pragma unroll
            for (int nt = 0; nt < NTW; ++nt) {
              if (halfw ? mt >= 2 : nt < mt) continue;
              const int c = wcE * 32 + cofs + mt * 8 + gq, r = wrE * RW + nt * 8 + 2 * tq;
              *reinterpret_cast<double2*>(&sm.Cs[c * LDT + r]) = make_double2(acc[mt][nt][0], acc[mt][nt][1]);
            }
        } else {
        {
            const double* T = &sm.stage[wc == 0 ? pre0 : pre1][0];
  #pragma unroll
            for (int mt = 0; mt < 4; ++mt)
  #pragma unroll
              for (int nt = 0; nt < NTW; ++nt) {
                const double2 v = *reinterpret_cast<const double2*>(&T[(mt * 8 + gq) * LDT + wr * RW + nt * 8 + 2 * tq]);
                acc[mt][nt][0] = v.x; acc[mt][nt][1] = v.y;
              }
          }
          bar_named(1, NCONS);             // X is complete in Cs
          // same k order as the streamed chunk it replaces (16 k-steps of 4): bitwise the same block
  #pragma unroll 4
          for (int ks = 0; ks < TILE / 4; ++ks) {
            const int k = ks * 4 + tq;
            double a[4], bb[NTW];
  #pragma unroll
            for (int mt = 0; mt < 4; ++mt) a[mt] = -sm.Cs[k * LDT + wc * 32 + mt * 8 + gq];
  #pragma unroll
            for (int nt = 0; nt < NTW; ++nt) bb[nt] = sm.Cs[k * LDT + wr * RW + nt * 8 + gq];
  #pragma unroll
            for (int mt = 0; mt < 4; ++mt)
  #pragma unroll
              for (int nt = 0; nt < NTW; ++nt) dmma884(acc[mt][nt][0], acc[mt][nt][1], a[mt], bb[nt]);
          }
          __syncwarp();
          release_pre();
          bar_named(1, NCONS);             // every warp has read X: Cs may take the block
  #pragma unroll
          for (int mt = 0; mt < 4; ++mt)
  #pragma unroll
            for (int nt = 0; nt < NTW; ++nt) {
              const int c = wc * 32 + mt * 8 + gq, r = wr * RW + nt * 8 + 2 * tq;
              *reinterpret_cast<double2*>(&sm.Cs[c * LDT + r]) = make_double2(acc[mt][nt][0], acc[mt][nt][1]);
            }
        }
        if (tid == 0) sm.bad_col = -1;
        bar_named(1, NCONS);
        factor_block(d.i, M + (size_t)(d.i * TILE) * g.ld + (size_t)d.i * TILE,
                     linvt + ((size_t)d.b * g.nt + d.i) * (TILE * LDT));
      }
      if (tr) { tr[6] = clock64(); tr[7] = globaltimer_ns(); }
    }
  }
}

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                             const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                             CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                             CUtensorMapFloatOOBfill);
static EncodeFn tmap_encoder() {
  static EncodeFn encode = nullptr;
  if (!encode) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess || !fn)
      return nullptr;
    encode = reinterpret_cast<EncodeFn>(fn);
  }
  return encode;
}

// L_jj^-T tiles: [batch * nt * TILE rows][LDT] doubles (rows c, pitch LDT), box {LDT, KC}.
static bool make_tmap_linvt(CUtensorMap* tm, const Geom& g, double* linvt) {
  EncodeFn encode = tmap_encoder();
  if (!encode) return false;
  const cuuint64_t dims[2] = {(cuuint64_t)LDT, ((cuuint64_t)g.batch * g.nt + 1) * TILE};   // + the identity tile
  const cuuint64_t strides[1] = {(cuuint64_t)LDT * sizeof(double)};
  const cuuint32_t box[2] = {(cuuint32_t)LDT, (cuuint32_t)KC};
  const cuuint32_t estr[2] = {1u, 1u};
  return encode(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, linvt, dims, strides, box, estr,
                CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

bool make_tmap(CUtensorMap* tm, const Geom& g, double* mats) {
  EncodeFn encode = tmap_encoder();
  if (!encode) return false;
  const cuuint64_t dims[3] = {(cuuint64_t)g.ld, (cuuint64_t)g.npad, (cuuint64_t)g.batch};
  const cuuint64_t strides[2] = {(cuuint64_t)g.ld * sizeof(double), (cuuint64_t)g.mat_stride * sizeof(double)};
  const cuuint32_t box[3] = {(cuuint32_t)LDT, (cuuint32_t)KC, 1u};
  const cuuint32_t estr[3] = {1u, 1u, 1u};
  return encode(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, mats, dims, strides, box, estr,
                CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// Host-side view of the ticket order (the same decode_task the kernel runs): schedule introspection
// for the tests and the trace tools.
long long chol_schedule_task(const Geom& g_in, long long ticket, int* out4) {
  Geom g = g_in;
  if (g.tgroup < 1 || g.tgroup > g.batch) g.tgroup = g.batch;
  const long long total = chol_task_count(g);
  if (ticket < 0 || ticket >= total || !out4) return total;
  int jcache = 0;
  const TaskDesc d = decode_task(g, (int)ticket, jcache);
  out4[0] = d.b; out4[1] = d.i; out4[2] = d.j; out4[3] = d.kind;
  return total;
}

int launch_chol(cudaStream_t st, const Geom& g_in, double* mats, double* linvt, int* done, int* info,
                SchedState* sched, int mode, int num_sms, int solve_only, unsigned long long* trace,
                int chain_pause, int skip_reset) {
  // per host thread (one gpb_ctx per thread): the attribute is set once per device, and the two TMA
  // descriptors are re-encoded only when the workspace or the geometry changed -- both are pure host
  // latency on every small call otherwise
  struct Entry { const double* mats = nullptr; const double* linvt = nullptr; int ld = 0, npad = 0, batch = 0;
                 size_t stride = 0; CUtensorMap tmap, tmapL; };
  struct Cache { int dev = -1; int next = 0; Entry e[4]; };   // the Mcmc.sample loop alternates two geometries
  thread_local Cache cc;
  int dev = 0;
  cudaGetDevice(&dev);
  if (cc.dev != dev) {
    cudaFuncSetAttribute(k_chol_tasks<false, GPB_MINBLOCKS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(CholSmem));
    cudaFuncSetAttribute(k_chol_tasks<false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(CholSmem));
    cudaFuncSetAttribute(k_chol_tasks<true, GPB_MINBLOCKS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(CholSmem));
    cudaFuncSetAttribute(k_chol_tasks<true, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(CholSmem));
    cc.dev = dev;
    for (Entry& e : cc.e) e.mats = nullptr;
  }
  Geom g = g_in;
  if (mode != 0 || g.tgroup < 1 || g.tgroup > g.batch) g.tgroup = g.batch;  // debug mode: one group
  Entry* hit = nullptr;
  for (Entry& e : cc.e)
    if (e.mats == mats && e.linvt == linvt && e.ld == g.ld && e.npad == g.npad && e.batch == g.batch && e.stride == g.mat_stride) hit = &e;
  if (!hit) {
    hit = &cc.e[cc.next];
    cc.next = (cc.next + 1) % 4;
    hit->mats = nullptr;
    if (!make_tmap(&hit->tmap, g, mats)) return -1;
    if (!make_tmap_linvt(&hit->tmapL, g, linvt)) return -1;
    hit->mats = mats; hit->linvt = linvt; hit->ld = g.ld; hit->npad = g.npad; hit->batch = g.batch; hit->stride = g.mat_stride;
  }
  const CUtensorMap& tmap = hit->tmap;
  const CUtensorMap& tmapL = hit->tmapL;
  const long long total = (long long)g.batch * ((long long)g.nt * g.rt - (long long)g.nt * (g.nt - 1) / 2);
  // tickets are 32-bit inside the kernel (one atomic counter, decode_task): refuse what would not fit
  // instead of truncating (64 chains x n=8192 are 5.4e5 tickets, n=32768 alone 1.3e5)
  if (total > 0x7fff0000LL) return -2;
  int launches = 0;
  if (!skip_reset || mode != 0) {
    const long long total = (long long)chol_done_ints(g);
    const int blocks = (int)((total + 255) / 256 < 296 ? (total + 255) / 256 : 296);
    k_chol_reset<<<blocks, 256, 0, st>>>(g, done, info, sched, linvt);
    ++launches;
  }
  if (mode == 0) {
    // Two co-resident CTAs per SM keep the DMMA pipe fed in throughput mode; launches with few tasks
    // per SM are bound by the dependency chain diag(j) -> (j+1, j) -> diag(j+1) instead, and a CTA
    // that has its SM (and an uncapped register budget) to itself factors and solves faster (n=4096,
    // one proposal: 1.20 ms vs 1.47 ms).  GPB_CTAS_PER_SM_RT overrides (tuning).
    static int env_ctas = -1;
    if (env_ctas < 0) { const char* e = getenv("GPB_CTAS_PER_SM_RT"); env_ctas = e ? atoi(e) : 0; }
    // crossover re-measured with the final kernels (profiles/r02_chain_pause_sweep.txt): ~18 tasks per SM from 16 block
    // columns on, earlier for shorter chains (n = 300: ~9, n = 640: ~13)
    const long long thr = g.nt >= 16 ? 18 : (6 + (3 * g.nt) / 4 < 8 ? 8 : 6 + (3 * g.nt) / 4);
    int ctas = (total >= thr * num_sms) ? CTAS_PER_SM : 1;
    if (env_ctas > 0 && env_ctas <= CTAS_PER_SM) ctas = env_ctas;
    // GPB_GRID_CAP=k (tuning): at most k CTAs per SM in the grid while keeping the instantiation `ctas` selects
    static int env_cap = -1;
    if (env_cap < 0) { const char* e = getenv("GPB_GRID_CAP"); env_cap = e ? atoi(e) : 0; }
    const long long slots = (long long)num_sms * ((env_cap > 0 && env_cap < ctas) ? env_cap : ctas);
    int grid = (int)(total < slots ? total : slots);
    // Two CTAs per SM, but still few enough tasks per SM that the chain diag(j) -> (j+1, j) -> diag(j+1)
    // bounds part of the launch: the CTA inside a block column's critical section gets the SM to itself
    // (see SchedState::pause; its X X^T + factorisation then takes 19.6 instead of 26.6 us).  Measured
    // (tools/pause_sweep.py, profiles/r02_chain_pause_sweep.txt): n=4096 x 2 proposals -10.7 %, x 3 -3.2 %,
    // n=2048 x 8 -4.6 %, n=1024 x 32 neutral; a loss once the launch is bound by CTA time (n=1024 x 64
    // +3.5 %, x 256 +4.3 %) or the chain is short (n=300 x 256 +10 %).  GPB_CHAIN_PAUSE=0|1 overrides.
    static int env_pause = -2;
    if (env_pause == -2) { const char* e = getenv("GPB_CHAIN_PAUSE"); env_pause = e ? atoi(e) : -1; }
    int pause = (ctas == 2 && !solve_only && g.nt >= 32 && total < (long long)PAUSE_TASKS_PER_SM * num_sms) ? 1 : 0;
    if (chain_pause >= 0) pause = chain_pause;
    if (env_pause >= 0) pause = env_pause;
    if (ctas != 2 || num_sms > SCHED_MAXSM) pause = 0;
    if (trace && ctas == 1)   // the schedule trace runs the instantiation the untraced launch would
      k_chol_tasks<true, 1><<<grid, NTHREADS, sizeof(CholSmem), st>>>(tmap, tmapL, g, mats, linvt, done, info, sched, 0,
                                                                      (int)total, 0, solve_only, trace, 0);
    else if (trace)
      k_chol_tasks<true, GPB_MINBLOCKS><<<grid, NTHREADS, sizeof(CholSmem), st>>>(tmap, tmapL, g, mats, linvt, done, info, sched, 0,
                                                                                  (int)total, 0, solve_only, trace, pause);
    else if (ctas == 1)
      k_chol_tasks<false, 1><<<grid, NTHREADS, sizeof(CholSmem), st>>>(tmap, tmapL, g, mats, linvt, done, info, sched, 0,
                                                                       (int)total, 0, solve_only, trace, 0);
    else
      k_chol_tasks<false, GPB_MINBLOCKS><<<grid, NTHREADS, sizeof(CholSmem), st>>>(tmap, tmapL, g, mats, linvt, done, info, sched, 0,
                                                                                   (int)total, 0, solve_only, trace, pause);
    ++launches;
  } else {
    // debug / fallback: one launch per ticket phase; every dependency is in an earlier launch
    auto run = [&](long long begin, long long count) {
      if (count <= 0) return;
      int grid = (int)(count < 2LL * num_sms ? count : 2LL * num_sms);
      k_chol_tasks<false, GPB_MINBLOCKS><<<grid, NTHREADS, sizeof(CholSmem), st>>>(tmap, tmapL, g, mats, linvt, done, info, sched,
                                                                                   (int)begin, (int)(begin + count), 1, solve_only, nullptr, 0);
      ++launches;
    };
    run(0, g.batch);
    for (int j = 0; j < g.nt; ++j) {
      long long begin = group_start(g.batch, g.rt, j);
      const long long nA = (j + 1 < g.rt) ? g.batch : 0, nB = (j + 1 < g.nt) ? g.batch : 0;
      const long long nC = (long long)g.batch * (g.rt - j - 2 > 0 ? g.rt - j - 2 : 0);
      run(begin, nB); run(begin + nB, nA); run(begin + nA + nB, nC);
    }
  }
  return launches;
}

}  // namespace gpb
