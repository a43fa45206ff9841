// gpb_plane.h -- the d x d algebra of the plane-mean Gibbs step (MeanParameters.samplePlane,
// MeanParameters.scala:86-101), d = dim + 1 <= 4, shared by the host entry point gpb_sample_plane and
// the device-resident Mcmc.sample loop (k_mh_gibbs).  Breeze `\` on a square matrix is an LU solve
// with partial pivoting, `cholesky` the lower factor.
#pragma once
#ifdef __CUDACC__
#define GPB_HD __host__ __device__
#else
#define GPB_HD
#endif

namespace gpb {

GPB_HD inline bool plane_lu_solve(int d, const double* A, const double* b, double* x) {
  double M[4][5];
  for (int i = 0; i < d; ++i) { for (int j = 0; j < d; ++j) M[i][j] = A[i * d + j]; M[i][d] = b[i]; }
  for (int c = 0; c < d; ++c) {
    int p = c;
    for (int i = c + 1; i < d; ++i) if (fabs(M[i][c]) > fabs(M[p][c])) p = i;
    if (M[p][c] == 0.0) return false;
    if (p != c) for (int j = 0; j <= d; ++j) { const double t = M[p][j]; M[p][j] = M[c][j]; M[c][j] = t; }
    for (int i = c + 1; i < d; ++i) {
      const double f = M[i][c] / M[c][c];
      for (int j = c; j <= d; ++j) M[i][j] -= f * M[c][j];
    }
  }
  for (int i = d - 1; i >= 0; --i) {
    double s = M[i][d];
    for (int j = i + 1; j < d; ++j) s -= M[i][j] * x[j];
    x[i] = s / M[i][i];
  }
  return true;
}

// G = X^T K^-1 X [d*d], w = X^T L^-1 y [d] (the reference multiplies X^T by u = L^-1 y, :94), z: d
// standard normals.  Returns 0, or -1 (singular precision) / j+1 (precision not positive definite).
GPB_HD inline int plane_draw(int d, const double* G, const double* w, double prior_mean, double prior_var,
                             const double* z, double* beta) {
  double prec[16], rhs[4], mean[4], root[16];
  for (int i = 0; i < 16; ++i) root[i] = 0.0;
  for (int i = 0; i < d; ++i) {
    for (int j = 0; j < d; ++j) prec[i * d + j] = G[i * d + j] + (i == j ? 1.0 / prior_var : 0.0);
    rhs[i] = (1.0 / prior_var) * prior_mean + w[i];   // priorPrec * fill(mean) + x.t * u  (:94-95)
  }
  if (!plane_lu_solve(d, prec, rhs, mean)) return -1;
  for (int j = 0; j < d; ++j) {                       // root = cholesky(prec)  (:97)
    double s = prec[j * d + j];
    for (int k = 0; k < j; ++k) s -= root[j * d + k] * root[j * d + k];
    if (!(s > 0.0)) return j + 1;
    root[j * d + j] = sqrt(s);
    for (int i = j + 1; i < d; ++i) {
      double t = prec[i * d + j];
      for (int k = 0; k < j; ++k) t -= root[i * d + k] * root[j * d + k];
      root[i * d + j] = t / root[j * d + j];
    }
  }
  for (int i = 0; i < d; ++i) {                       // beta = mean + root \ z  (:99)
    double s = z[i];
    for (int k = 0; k < i; ++k) s -= root[i * d + k] * (beta[k] - mean[k]);
    beta[i] = mean[i] + s / root[i * d + i];
  }
  return 0;
}

}  // namespace gpb
