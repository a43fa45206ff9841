// gpb200.hpp -- C++17 host-side mirror of the reference's Scala API for the GP hot path, written
// above the C ABI of libgpb200.so (include/gpb200.h).  Header only.
//
// The reference is JVM code and no JVM/scalac exists in this image, so this is the compiled-language
// stand-in for its operator interface: same object / method names, argument meaning and error
// behaviour as core/src/main/scala/gp/ (file:line cited per method).  The Scala (Panama) binding a
// maintainer would add is gaussian-processes_b200/scala/Gpb200.scala; the Python mirror the test
// suite drives is gaussian-processes_b200/gpb200/api.py.  No CPU fallback: every numerical call
// goes through the C ABI and throws when there is no sm_100 device.
#pragma once
#include <algorithm>
#include <cmath>
#include <functional>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "gpb200.h"

namespace gp {

// ---- Location.scala:5-33 ------------------------------------------------------------------------
struct Location {
  int dim;       // One -> 1, Two -> 2
  double x, y;   // y unused for One
  std::vector<double> toVector() const { return dim == 1 ? std::vector<double>{x} : std::vector<double>{x, y}; }
  // Location.euclidean (Location.scala:27-33); mixed dimensions throw like the reference (:32)
  static double euclidean(const Location& a, const Location& b) {
    if (a.dim != b.dim) throw std::runtime_error("Dimension mismatch");
    if (a.dim == 1) return std::sqrt(std::pow(a.x - b.x, 2));
    return std::sqrt(std::pow(a.x - b.x, 2) + std::pow(a.y - b.y, 2));
  }
};
inline Location One(double x) { return Location{1, x, 0.0}; }
inline Location Two(double x, double y) { return Location{2, x, y}; }
using DistFn = double (*)(const Location&, const Location&);

// ---- KernelParameters.scala:11-85 (ADT), MeanParameters.scala:8-46 ------------------------------
struct KernelParameters {
  enum Kind { SquaredExp = GPB_KIND_SE, White = GPB_KIND_WHITE, Matern = GPB_KIND_MATERN } kind;
  double p0, p1, p2;
  static KernelParameters se(double h, double sigma) { return {SquaredExp, h, sigma, 0.0}; }
  static KernelParameters white(double sigma) { return {White, sigma, 0.0, 0.0}; }
  static KernelParameters matern(double sigma, double nu, double l) { return {Matern, sigma, nu, l}; }
  std::vector<double> toList() const {
    if (kind == SquaredExp) return {p0, p1};
    if (kind == White) return {p0};
    return {p0, p1, p2};
  }
};
struct MeanParameters {
  std::vector<double> beta;   // empty == Zero, dim+1 entries == Plane(beta)
  static MeanParameters zero() { return {}; }
  static MeanParameters plane(std::vector<double> b) { return {std::move(b)}; }
  std::vector<double> toList() const { return beta; }
};
// GaussianProcess.Parameters (GaussianProcess.scala:15-29); toList: mean first, then kernel terms
struct Parameters {
  MeanParameters meanParameters;
  std::vector<KernelParameters> kernelParameters;
  std::vector<double> toList() const {
    std::vector<double> v = meanParameters.toList();
    for (const auto& k : kernelParameters) for (double d : k.toList()) v.push_back(d);
    return v;
  }
};
struct Data { Location x; double y; };          // GaussianProcess.Data (GaussianProcess.scala:36)
struct Gaussian {                               // breeze Gaussian(mu, sigma) as Predict.fit returns it
  double mean, sigma;
  double variance() const { return sigma * sigma; }
};
// what breeze.linalg.cholesky throws on a matrix that is not positive definite
struct NotConvergedException : std::runtime_error {
  int info;
  explicit NotConvergedException(int i) : std::runtime_error("matrix not positive definite (info=" + std::to_string(i) + ")"), info(i) {}
};

namespace detail {
inline void check(int status, gpb_ctx* ctx = nullptr) {
  if (status == GPB_OK) return;
  std::string msg = std::string("gpb200: ") + gpb_strerror(status);
  if (ctx) { const char* e = gpb_last_cuda_error(ctx); if (e && *e) msg += std::string(" [") + e + "]"; }
  if (status == GPB_E_UNSUPPORTED) throw std::invalid_argument(msg);
  throw std::runtime_error(msg);   // includes GPB_E_NODEVICE: there is no CPU fallback
}
struct Ctx {
  gpb_ctx* h = nullptr;
  explicit Ctx(int device) { check(gpb_ctx_create(device, &h)); }
  ~Ctx() { gpb_ctx_destroy(h); }
  Ctx(const Ctx&) = delete;
  Ctx& operator=(const Ctx&) = delete;
};
// one context per host thread (a gpb_ctx is not shared between threads)
inline gpb_ctx* context() {
  thread_local std::unique_ptr<Ctx> c;
  if (!c) c.reset(new Ctx(0));
  return c->h;
}
inline void requireEuclidean(DistFn dist) {
  // dist is an arbitrary closure in the reference; only Location.euclidean ships.  Anything else is
  // refused instead of being emulated on the CPU.
  if (dist != &Location::euclidean) throw std::invalid_argument("gpb200 supports Location::euclidean only");
}
inline std::vector<double> soa(const std::vector<Location>& xs, int& dim) {
  if (xs.empty()) throw std::invalid_argument("empty location vector");
  dim = xs[0].dim;
  std::vector<double> v((size_t)dim * xs.size());
  for (size_t i = 0; i < xs.size(); ++i) {
    if (xs[i].dim != dim) throw std::runtime_error("Dimension mismatch");
    v[i] = xs[i].x;
    if (dim == 2) v[xs.size() + i] = xs[i].y;
  }
  return v;
}
inline std::vector<gpb_term> terms(const std::vector<KernelParameters>& ps) {
  std::vector<gpb_term> t;
  for (const auto& p : ps) t.push_back(gpb_term{(int)p.kind, p.p0, p.p1, p.p2});
  return t;
}
struct Dataset {
  gpb_dataset* h = nullptr;
  int n = 0, dim = 0;
  Dataset(const std::vector<Location>& xs, const std::vector<double>& y) {
    std::vector<double> x = soa(xs, dim);
    n = (int)xs.size();
    if (!y.empty() && (int)y.size() != n) throw std::runtime_error("Dimension mismatch");
    check(gpb_dataset_create(context(), n, dim, x.data(), y.empty() ? nullptr : y.data(), &h), context());
  }
  ~Dataset() { gpb_dataset_destroy(h); }
  Dataset(const Dataset&) = delete;
  Dataset& operator=(const Dataset&) = delete;
};
}  // namespace detail

// A column-major dense matrix (Breeze DenseMatrix layout), just enough for the boundary.
struct DenseMatrix {
  int rows = 0, cols = 0;
  std::vector<double> data;
  DenseMatrix() = default;
  DenseMatrix(int r, int c) : rows(r), cols(c), data((size_t)r * c, 0.0) {}
  double& operator()(int i, int j) { return data[(size_t)j * rows + i]; }
  double operator()(int i, int j) const { return data[(size_t)j * rows + i]; }
};

// breeze.linalg.cholesky as called at GaussianProcess.scala:123 / Predict.scala:69: lower factor with a
// zeroed upper triangle; throws NotConvergedException when A is not positive definite.
inline DenseMatrix cholesky(const DenseMatrix& A) {
  if (A.rows != A.cols) throw std::runtime_error("Dimension mismatch");
  DenseMatrix L = A; int info = 0;
  detail::check(gpb_potrf(detail::context(), A.rows, L.data.data(), &info), detail::context());
  if (info != 0) throw NotConvergedException(info);
  return L;
}

// ---- MeanFunction.scala:7-24 ---------------------------------------------------------------------
struct MeanFunction {
  static std::function<double(const Location&)> apply(const MeanParameters& p) {
    if (p.beta.empty()) return [](const Location&) { return 0.0; };
    std::vector<double> b = p.beta;
    return [b](const Location& l) {
      if (l.dim == 1) return b.at(0) + l.x * b.at(1);
      return b.at(0) + l.x * b.at(1) + l.y * b.at(2);
    };
  }
};

// ---- KernelFunction.scala:84-89 ------------------------------------------------------------------
struct KernelFunction {
  // buildCov(xs, covFn, dist): covFn is KernelFunction.apply(ps) in every reference call site, so the
  // mirror takes ps; no nugget (the callers add it, GaussianProcess.scala:117-118).
  static DenseMatrix buildCov(const std::vector<Location>& xs, const std::vector<KernelParameters>& ps, DistFn dist) {
    detail::requireEuclidean(dist);
    int dim; std::vector<double> x = detail::soa(xs, dim);
    int n = (int)xs.size();
    DenseMatrix K(n, n);
    auto t = detail::terms(ps);
    detail::check(gpb_build_cov(detail::context(), n, dim, x.data(), t.data(), (int)t.size(), 0.0, K.data.data()), detail::context());
    return K;
  }
};

// ---- GaussianProcess.scala -----------------------------------------------------------------------
struct GaussianProcess {
  // which scalar loglikelihood returns: what GaussianProcess.scala:126 executes (default) or GPML's
  static int& formula() { static int f = GPB_FORMULA_AS_PARSED; return f; }

  // GaussianProcess.scala:50-61
  static DenseMatrix distanceMatrix(const std::vector<Location>& xs, DistFn dist) {
    detail::requireEuclidean(dist);
    int dim; std::vector<double> x = detail::soa(xs, dim);
    int n = (int)xs.size();
    DenseMatrix D(n, n);
    detail::check(gpb_distance_matrix(detail::context(), n, dim, x.data(), D.data.data()), detail::context());
    return D;
  }

  // The curried closure of GaussianProcess.scala:109-127: the outer application uploads the
  // observations once, operator() is one factorisation; batch() is many proposals in one pass.
  class LogLikelihood {
    std::shared_ptr<detail::Dataset> ds_;
   public:
    explicit LogLikelihood(std::shared_ptr<detail::Dataset> ds) : ds_(std::move(ds)) {}
    double operator()(const Parameters& p) const { return batch({p})[0]; }
    std::vector<double> batch(const std::vector<Parameters>& ps) const {
      if (ps.empty()) return {};
      const size_t nterms = ps[0].kernelParameters.size(), nbeta = ps[0].meanParameters.beta.size();
      if (nbeta != 0 && (int)nbeta != ds_->dim + 1) throw std::runtime_error("Dimension mismatch");
      std::vector<gpb_term> t; std::vector<double> beta;
      for (const auto& p : ps) {
        if (p.kernelParameters.size() != nterms || p.meanParameters.beta.size() != nbeta)
          throw std::invalid_argument("batch proposals must share one parameter structure");
        for (const auto& g : detail::terms(p.kernelParameters)) t.push_back(g);
        beta.insert(beta.end(), p.meanParameters.beta.begin(), p.meanParameters.beta.end());
      }
      std::vector<gpb_ll> r(ps.size());
      detail::check(gpb_loglik_batch(detail::context(), ds_->h, (int)ps.size(), t.data(), (int)nterms,
                                     nbeta ? beta.data() : nullptr, (int)nbeta, 1e-3, r.data()), detail::context());
      std::vector<double> out;
      for (const auto& x : r) {
        if (x.info != 0) throw NotConvergedException(x.info);   // breeze cholesky throws (:123)
        out.push_back(gpb_ll_value(&x, formula()));
      }
      return out;
    }
  };
  static LogLikelihood loglikelihood(const std::vector<Data>& observed, DistFn dist) {
    detail::requireEuclidean(dist);
    std::vector<Location> xs; std::vector<double> y;
    for (const auto& d : observed) { xs.push_back(d.x); y.push_back(d.y); }
    return LogLikelihood(std::make_shared<detail::Dataset>(xs, y));
  }

  // mllGradient(observed, dist)(p) (GaussianProcess.scala:166-187): p holds the UNCONSTRAINED kernel
  // parameters, as sampleHmc passes them (KernelParameters.scala:141-144); one entry per kernel
  // parameter in toList order.  gradientMode(): as executed by the Scala (default) or analytic.
  static int& gradientMode() { static int m = GPB_GRAD_AS_EXECUTED; return m; }
  class MllGradient {
    std::shared_ptr<detail::Dataset> ds_;
   public:
    explicit MllGradient(std::shared_ptr<detail::Dataset> ds) : ds_(std::move(ds)) {}
    std::vector<double> operator()(const Parameters& p) const {
      const size_t nbeta = p.meanParameters.beta.size();
      if (nbeta != 0 && (int)nbeta != ds_->dim + 1) throw std::runtime_error("Dimension mismatch");
      auto t = detail::terms(p.kernelParameters);
      size_t P = 0; for (const auto& k : p.kernelParameters) P += k.toList().size();
      std::vector<double> g(P); int info = 0;
      detail::check(gpb_mll_gradient(detail::context(), ds_->h, 1, t.data(), (int)t.size(), nbeta ? p.meanParameters.beta.data() : nullptr,
                                     (int)nbeta, 1e-3, gradientMode(), g.data(), &info), detail::context());
      if (info != 0) throw NotConvergedException(info);   // the reference's LU would carry on (gpb200.h)
      return g;
    }
  };
  static MllGradient mllGradient(const std::vector<Data>& observed, DistFn dist) {
    detail::requireEuclidean(dist);
    std::vector<Location> xs; std::vector<double> y;
    for (const auto& d : observed) { xs.push_back(d.x); y.push_back(d.y); }
    return MllGradient(std::make_shared<detail::Dataset>(xs, y));
  }

  // GaussianProcess.scala:67-83: mu + root z, nugget 1e-3; z = the |xs| standard normals the
  // reference draws from breeze (passed in so the stream is explicit).  Cholesky root (8f-3).
  static std::vector<Data> draw(const std::vector<Location>& xs, DistFn dist, const Parameters& p, const std::vector<double>& z) {
    detail::requireEuclidean(dist);
    if (z.size() != xs.size()) throw std::runtime_error("Dimension mismatch");
    int dim; std::vector<double> x = detail::soa(xs, dim);
    auto mf = MeanFunction::apply(p.meanParameters);
    std::vector<double> mu; for (const auto& l : xs) mu.push_back(mf(l));
    auto t = detail::terms(p.kernelParameters);
    std::vector<double> out(xs.size()); int info = 0;
    detail::check(gpb_draw(detail::context(), (int)xs.size(), dim, x.data(), t.data(), (int)t.size(), 1e-3,
                           mu.data(), z.data(), out.data(), &info), detail::context());
    if (info != 0) throw NotConvergedException(info);
    std::vector<Data> r; for (size_t i = 0; i < xs.size(); ++i) r.push_back(Data{xs[i], out[i]});
    return r;
  }
};

// ---- Predict.scala -------------------------------------------------------------------------------
struct Predict {
  // Predict.scala:40-45 (dtrsv "L","N","N")
  static std::vector<double> forwardSolve(const DenseMatrix& A, const std::vector<double>& y) {
    if (A.rows != A.cols || (int)y.size() != A.rows) throw std::runtime_error("Dimension mismatch");
    std::vector<double> b = y;
    detail::check(gpb_trsm_lower(detail::context(), A.rows, 1, A.data.data(), b.data()), detail::context());
    return b;
  }
  // Predict.scala:57-94: raw y, nugget 1e-6, Gaussian(mean, sqrt(var))
  static std::vector<std::pair<Location, Gaussian>> fit(const std::vector<Location>& newXs, const std::vector<Data>& observed,
                                                        DistFn dist, const Parameters& p) {
    detail::requireEuclidean(dist);
    std::vector<Location> xs; std::vector<double> y;
    for (const auto& d : observed) { xs.push_back(d.x); y.push_back(d.y); }
    detail::Dataset ds(xs, y);
    int sdim; std::vector<double> xstar = detail::soa(newXs, sdim);
    if (sdim != ds.dim) throw std::runtime_error("Dimension mismatch");
    auto t = detail::terms(p.kernelParameters);
    const int m = (int)newXs.size();
    std::vector<double> mean(m), var(m); int info = 0;
    detail::check(gpb_predict(detail::context(), ds.h, t.data(), (int)t.size(), 1e-6, m, xstar.data(), mean.data(), var.data(), &info),
                  detail::context());
    if (info != 0) throw NotConvergedException(info);
    std::vector<std::pair<Location, Gaussian>> out;
    for (int i = 0; i < m; ++i) out.emplace_back(newXs[i], Gaussian{mean[i], var[i] >= 0 ? std::sqrt(var[i]) : std::nan("")});
    return out;
  }
  static std::vector<Data> buildData(const std::vector<Location>& xs, const std::vector<double>& ys) {
    std::vector<Data> d; for (size_t i = 0; i < xs.size() && i < ys.size(); ++i) d.push_back(Data{xs[i], ys[i]});
    return d;
  }
};

// ---- MeanParameters.scala:55-121 -----------------------------------------------------------------
struct MeanSampler {
  // makeDesignMatrix (MeanParameters.scala:55-65) INCLUDING its layout: the row-major flattened
  // coordinates are read back column-major, so for Two the x / y columns come out interleaved.
  static DenseMatrix makeDesignMatrix(const std::vector<Location>& xs) {
    if (xs.empty()) throw std::invalid_argument("empty location vector");
    const int n = (int)xs.size(), m = xs[0].dim;
    std::vector<double> flat;
    for (const auto& l : xs) { if (l.dim != m) throw std::runtime_error("Dimension mismatch"); for (double v : l.toVector()) flat.push_back(v); }
    DenseMatrix X(n, m + 1);
    for (int i = 0; i < n; ++i) { X(i, 0) = 1.0; for (int j = 0; j < m; ++j) X(i, j + 1) = flat[(size_t)j * n + i]; }
    return X;
  }
  // samplePlane(prior, obs, xs, dist, p) (MeanParameters.scala:74-103); prior = Gaussian(mean, sigma);
  // z = the dim+1 standard normals the reference draws.  nugget 1e-6 (:84).
  static MeanParameters samplePlane(const Gaussian& prior, const std::vector<Data>& obs, const std::vector<Location>& xs,
                                    DistFn dist, const Parameters& p, const std::vector<double>& z) {
    detail::requireEuclidean(dist);
    if (xs.size() != obs.size()) throw std::runtime_error("Dimension mismatch");
    DenseMatrix X = makeDesignMatrix(xs);
    if ((int)z.size() != X.cols) throw std::runtime_error("Dimension mismatch");
    std::vector<double> y; for (const auto& d : obs) y.push_back(d.y);
    detail::Dataset ds(xs, y);
    auto t = detail::terms(p.kernelParameters);
    std::vector<double> beta(X.cols); int info = 0;
    detail::check(gpb_sample_plane(detail::context(), ds.h, t.data(), (int)t.size(), 1e-6, X.cols, X.data.data(), prior.mean,
                                   prior.variance(), z.data(), beta.data(), &info), detail::context());
    if (info != 0) throw NotConvergedException(info);
    return MeanParameters::plane(beta);
  }
};

// ---- Mcmc.scala:35-58, KernelParameters.scala:231-246 --------------------------------------------
struct Mcmc {
  static std::vector<double> backSolve(const DenseMatrix& A, const std::vector<double>& y) {
    if (A.rows != A.cols || (int)y.size() != A.rows) throw std::runtime_error("Dimension mismatch");
    std::vector<double> b = y;
    detail::check(gpb_trsm_upper(detail::context(), A.rows, 1, A.data.data(), b.data()), detail::context());
    return b;
  }
  static DenseMatrix backSolve(const DenseMatrix& A, const DenseMatrix& Y) {
    if (A.rows != A.cols || Y.rows != A.rows) throw std::runtime_error("Dimension mismatch");
    DenseMatrix B = Y;
    detail::check(gpb_trsm_upper(detail::context(), A.rows, Y.cols, A.data.data(), B.data.data()), detail::context());
    return B;
  }
};

// KernelParameters.sample(obs, dist, prior, prop) with Breeze's metropolis kernel: step(p, z, u) -> p'.
// prior / prop stay host closures like in the reference; prop receives the standard normals z of this
// step, u is the acceptance uniform.  ll(next) and ll(current) go to the device as ONE batch of two.
struct KernelSampler {
  using Prior = std::function<double(const std::vector<KernelParameters>&)>;
  using Proposal = std::function<std::vector<KernelParameters>(const std::vector<KernelParameters>&, const std::vector<double>&)>;
  static std::function<Parameters(const Parameters&, const std::vector<double>&, double)>
  sample(const std::vector<Data>& obs, DistFn dist, Prior prior, Proposal prop) {
    auto ll = GaussianProcess::loglikelihood(obs, dist);
    return [ll, prior, prop](const Parameters& p, const std::vector<double>& z, double u) {
      Parameters nxt{p.meanParameters, prop(p.kernelParameters, z)};
      std::vector<double> v = ll.batch({nxt, p});
      // JVM semantics of Breeze's metropolis: a NaN difference rejects (scala.math.min propagates NaN,
      // `u < NaN` is false); std::min(1.0, NaN) would return 1.0 and always accept
      const double dl = prior(nxt.kernelParameters) + v[0] - prior(p.kernelParameters) - v[1];
      const bool accept = (dl == dl) && u < std::min(1.0, std::exp(std::min(dl, 700.0)));
      return accept ? nxt : p;
    };
  }
};

}  // namespace gp
