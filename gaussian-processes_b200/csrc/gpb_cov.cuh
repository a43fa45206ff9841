// gpb_cov.cuh -- the covariance function in the reference's exact operation order, shared by k_build
// (matrix-level gpb_build_cov) and the one-launch small-n Metropolis-Hastings kernel.
#pragma once
#include "gpb_common.cuh"

namespace gpb {

__device__ __forceinline__ double cov_of_dist(const Hyper& h, double d) {
  // KernelFunction.scala:47-55: ps.map(k_i(d)).sum == ((0.0 + k0) + k1) + ...
  double s = 0.0;
#pragma unroll 1
  for (int t = 0; t < h.nterms; ++t) {
    double v;
    if (h.kind[t] == GPB_KIND_SE) {
      double sg = h.p1[t];
      // KernelFunction.scala:14; explicit _rn ops: the JVM never contracts a*b+c into an FMA
      v = __dmul_rn(h.p0[t], exp(-__ddiv_rn(__dmul_rn(d, d), __dmul_rn(sg, sg))));
    } else {
      v = (d == 0.0) ? h.p0[t] : 0.0;  // KernelFunction.scala:34
    }
    s = __dadd_rn(s, v);
  }
  return s;
}


// Location.euclidean (Location.scala:27-33): sqrt(dx^2 [+ dy^2]) with no FMA contraction.
__device__ __forceinline__ double euclid(int dim, double ax, double ay, double bx, double by) {
  const double dx = ax - bx;
  double d2 = __dmul_rn(dx, dx);
  if (dim == 2) {
    const double dy = ay - by;
    d2 = __dadd_rn(d2, __dmul_rn(dy, dy));
  }
  return sqrt(d2);
}

}  // namespace gpb
