// gpb_internal.h -- host-visible launchers of the sm_100a kernels (one .cu per kernel family).
#pragma once
#include "gpb_common.cuh"

namespace gpb {

// Scheduler state of the factorisation kernel (device memory, zeroed before every launch).
constexpr int SCHED_MAXSM = 256;
struct SchedState {
  int next_task;  // atomic ticket
  int error;      // != 0: a watchdog fired (GPB_E_WATCHDOG); kernels drain without waiting
  // Chain-bound launches with two CTAs per SM (k_chol_tasks, `chain_pause`): a CTA that runs the
  // critical section of a block column (solve tile (j+1, j), update and factor diag(j+1)) asks the
  // producer warp of the CTA it shares the SM with to stop feeding its ring for that long.
  int smslot[SCHED_MAXSM];      // arrival counter per SM: the CTA's slot (0 / 1) on its SM
  int pause[2 * SCHED_MAXSM];   // pause[2 * smid + slot] != 0: that CTA is inside a critical section
};
constexpr size_t SCHED_HEAD_BYTES = 2 * sizeof(int);   // what the host reads back after every call

// k_build modes: the reference's exact operation order (matrix-level gpb_build_cov), the distance
// matrix, and the hot-path form used by every factorising entry point (gpb_build.cu).
constexpr int BUILD_EXACT = 0, BUILD_DISTANCE = 1, BUILD_FAST = 2;
void launch_build(cudaStream_t st, const Geom& g, const double* x, const double* y,
                  const double* xstar, const Hyper* hyper, double nugget, double* mats, int mode,
                  int full, const double* rows = nullptr);

// Tile Cholesky + panel solves of the tall system (gpb_chol.cu).
//   done   [batch][rt]  per tile-row progress counters (tiles (i, 0..done-1) are final), followed by
//          [batch][nt]  flags "diagonal block pre-accumulated over all but its last block column" and
//          [batch][nt]  flags "rows 0..31 of L_jj^-T published by the running factorisation"
//          (chol_done_ints(g) ints in all)
//   linvt  [batch][nt][TILE][LDT]  inverse-transposes of the diagonal blocks (side output, pitch LDT)
//   info   [batch]      first non-positive pivot (1-based, LAPACK) or 0
// mode 0: one persistent launch over all tasks; mode 1: two launches per block column.
// Returns the number of kernel launches issued.
int launch_chol(cudaStream_t st, const Geom& g, double* mats, double* linvt, int* done,
                int* info, SchedState* sched, int mode, int num_sms, int solve_only = 0,
                unsigned long long* trace = nullptr, int chain_pause = -1, int skip_reset = 0);
// skip_reset: the caller has already zeroed the scheduler state of this launch (the MH loop's step kernel
// does it for the next iteration: one launch less per iteration)
inline size_t chol_done_ints(const Geom& g) { return (size_t)g.batch * (g.rt + 2 * g.nt); }
constexpr int CHOL_TRACE_SLOTS = 16;
long long chol_task_count(const Geom& g);
// ticket -> {proposal, tile row, tile column, kind}; returns the number of tickets
long long chol_schedule_task(const Geom& g, long long ticket, int* out4);
// solve_only: the top square already holds a lower-triangular factor; only the appended rows are
// solved (X = B L^-T) -- the matrix-level forward / backward substitutions.
size_t chol_smem_bytes();
// doubles of the L_jj^-T side buffer (tiles of TILE rows x LDT pitch)
size_t chol_linvt_doubles(const Geom& g);

// Likelihood reductions Q, U1, S per proposal (gpb_epilogue.cu).
void launch_ll_epilogue(cudaStream_t st, const Geom& g, const double* mats, const int* info,
                        gpb_ll* out);
// Plane-mean Gibbs statistics: G = V V^T over the solved design rows, w = X^T u (gpb_epilogue.cu).
void launch_plane_epilogue(cudaStream_t st, const Geom& g, const double* mats, const double* rows,
                           double* out /* [batch][design*design + design] */);
// mu + L z for a factored matrix (GaussianProcess.draw), one CTA per 64 rows.
void launch_draw(cudaStream_t st, const Geom& g, const double* mats, const double* z, const double* mu,
                 double* out);
// Predictive mean / variance per test point.
// part: scratch of 2 * pred_slices(g) * m doubles (column-slice partial sums, reduced in fixed order).
int pred_slices(const Geom& g);
void launch_pred_epilogue(cudaStream_t st, const Geom& g, const double* mats, const Hyper* hyper,
                          double* part, double* mean, double* var);
// Identity on the padded diagonal (matrix-level gpb_potrf input).
void launch_pad_identity(cudaStream_t st, const Geom& g, double* mats);
// Zero the strict upper triangle of the top square (matrix-level gpb_potrf output).
void launch_zero_upper(cudaStream_t st, const Geom& g, double* mats);

}  // namespace gpb
