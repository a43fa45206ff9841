// Drives the C++ mirror of the reference API (include/gpb200.hpp) on the GPU and prints inputs and
// results as one JSON object; tests/test_cpp_mirror.py compares them with the CPU oracle.
// Build: g++ -std=c++17 -Iinclude tests/cpp/test_mirror.cpp -L<libdir> -lgpb200 -Wl,-rpath,<libdir>
#include <cstdint>
#include <cstdio>
#include <string>

#include "gpb200.hpp"

using namespace gp;

static uint64_t s_state = 0x9E3779B97F4A7C15ull;
static double unif() {   // splitmix64 -> (0,1); the inputs are printed, so the generator need not be mirrored
  uint64_t z = (s_state += 0x9E3779B97F4A7C15ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z ^= z >> 31;
  return ((z >> 11) + 0.5) / 9007199254740992.0;
}
static double normal() { return std::sqrt(-2.0 * std::log(unif())) * std::cos(6.283185307179586 * unif()); }

static void vec(const char* name, const std::vector<double>& v, bool comma = true) {
  std::printf("\"%s\": [", name);
  for (size_t i = 0; i < v.size(); ++i) std::printf("%s%.17g", i ? ", " : "", v[i]);
  std::printf("]%s\n", comma ? "," : "");
}
template <class F> static std::string thrown(F f) {
  try { f(); } catch (const NotConvergedException&) { return "NotConvergedException"; }
  catch (const std::invalid_argument&) { return "invalid_argument"; }
  catch (const std::runtime_error& e) { return e.what(); }
  return "nothing";
}

int main() {
  const std::vector<KernelParameters> ks = {KernelParameters::se(3.0, 5.5), KernelParameters::white(0.5)};
  std::printf("{\n");

  // ---- 1-D, Zero mean: loglikelihood, batch with a plane mean, Predict.fit ----------------------
  {
    const int n = 150, m = 40;
    std::vector<Location> xs; std::vector<double> x, y, xstar;
    for (int i = 0; i < n; ++i) { x.push_back(-10.0 + 20.0 * unif()); xs.push_back(One(x.back())); }
    xs[5] = xs[2]; x[5] = x[2];                       // duplicate location: white noise off the diagonal
    for (int i = 0; i < n; ++i) y.push_back(1.5 * std::sin(0.4 * x[i]) + 0.3 * normal());
    std::vector<Data> obs = Predict::buildData(xs, y);
    auto ll = GaussianProcess::loglikelihood(obs, &Location::euclidean);
    Parameters p{MeanParameters::zero(), ks};
    vec("x1", x); vec("y1", y);
    std::printf("\"ll_as_parsed\": %.17g,\n", ll(p));
    GaussianProcess::formula() = GPB_FORMULA_GPML;
    std::printf("\"ll_gpml\": %.17g,\n", ll(p));
    std::vector<Parameters> ps;
    for (int b = 0; b < 3; ++b)
      ps.push_back(Parameters{MeanParameters::plane({0.2 * b, -0.1 + 0.05 * b}),
                              {KernelParameters::se(2.0 + b, 4.0 + 0.5 * b), KernelParameters::white(0.4 + 0.1 * b)}});
    vec("ll_batch_gpml", ll.batch(ps));
    GaussianProcess::formula() = GPB_FORMULA_AS_PARSED;
    std::vector<Location> nx; for (int i = 0; i < m; ++i) { xstar.push_back(-12.0 + 24.0 * i / (m - 1)); nx.push_back(One(xstar.back())); }
    auto fit = Predict::fit(nx, obs, &Location::euclidean, p);
    std::vector<double> mu, sd; for (auto& f : fit) { mu.push_back(f.second.mean); sd.push_back(f.second.sigma); }
    vec("xstar1", xstar); vec("fit_mean", mu); vec("fit_sigma", sd);

    // KernelParameters.sample: 25 Metropolis steps, explicit streams, log-normal walk, IG priors
    const double delta = 0.05, ig[3][2] = {{3.0, 6.0}, {3.0, 10.0}, {3.0, 1.0}};
    auto lig = [](double a, double b, double v) { return a * std::log(b) - std::lgamma(a) - (a + 1.0) * std::log(v) - b / v; };
    auto prior = [&](const std::vector<KernelParameters>& k) {
      return (lig(ig[0][0], ig[0][1], k[0].p0) + lig(ig[1][0], ig[1][1], k[0].p1)) + lig(ig[2][0], ig[2][1], k[1].p0);
    };
    auto prop = [&](const std::vector<KernelParameters>& k, const std::vector<double>& z) {
      return std::vector<KernelParameters>{KernelParameters::se(k[0].p0 * std::exp(delta * z[0]), k[0].p1 * std::exp(delta * z[1])),
                                           KernelParameters::white(k[1].p0 * std::exp(delta * z[2]))};
    };
    auto step = KernelSampler::sample(obs, &Location::euclidean, prior, prop);
    std::vector<double> zs, us, states; Parameters cur = p;
    for (int it = 0; it < 25; ++it) {
      std::vector<double> z = {normal(), normal(), normal()}; double u = unif();
      cur = step(cur, z, u);
      zs.insert(zs.end(), z.begin(), z.end()); us.push_back(u);
      for (double v : cur.toList()) states.push_back(v);
    }
    vec("mh_z", zs); vec("mh_u", us); vec("mh_states", states);
  }

  // ---- 2-D: distanceMatrix, buildCov, cholesky, forwardSolve / backSolve, samplePlane, draw -----
  {
    const int n = 90;
    std::vector<Location> xs; std::vector<double> xa, xb, y, z;
    for (int i = 0; i < n; ++i) { xa.push_back(-5.0 + 10.0 * unif()); xb.push_back(-5.0 + 10.0 * unif()); xs.push_back(Two(xa[i], xb[i])); }
    for (int i = 0; i < n; ++i) { y.push_back(0.7 + 0.2 * xa[i] - 0.1 * xb[i] + normal()); z.push_back(normal()); }
    vec("xa", xa); vec("xb", xb); vec("y2", y); vec("z2", z);
    DenseMatrix D = GaussianProcess::distanceMatrix(xs, &Location::euclidean);
    DenseMatrix K = KernelFunction::buildCov(xs, ks, &Location::euclidean);
    vec("D", D.data); vec("K", K.data);
    for (int i = 0; i < n; ++i) K(i, i) += 1e-3;
    DenseMatrix L = cholesky(K);
    vec("L", L.data);
    std::vector<double> u = Predict::forwardSolve(L, y);
    vec("fwd", u);
    DenseMatrix U(n, n); for (int i = 0; i < n; ++i) for (int j = 0; j < n; ++j) U(i, j) = L(j, i);
    vec("back", Mcmc::backSolve(U, u));
    DenseMatrix Y(n, 2); for (int i = 0; i < n; ++i) { Y(i, 0) = y[i]; Y(i, 1) = z[i]; }
    vec("back2", Mcmc::backSolve(U, Y).data);
    vec("design", MeanSampler::makeDesignMatrix(xs).data);
    std::vector<double> zb = {normal(), normal(), normal()};
    vec("zb", zb);
    Parameters p{MeanParameters::plane({0.5, 0.1, -0.2}), ks};
    vec("plane_beta", MeanSampler::samplePlane(Gaussian{1.0, 2.0}, Predict::buildData(xs, y), xs, &Location::euclidean, p, zb).beta);
    // mllGradient: unconstrained parameters; as executed (positive values keep kxx PD) and analytic
    auto grad = GaussianProcess::mllGradient(Predict::buildData(xs, y), &Location::euclidean);
    vec("grad_as_executed", grad(Parameters{MeanParameters::zero(), {KernelParameters::se(1.1, 0.9), KernelParameters::white(0.3)}}));
    GaussianProcess::gradientMode() = GPB_GRAD_ANALYTIC;
    vec("grad_analytic", grad(Parameters{MeanParameters::plane({0.5, 0.1, -0.2}),
                                         {KernelParameters::se(std::log(3.0), std::log(5.5)), KernelParameters::white(std::log(0.5))}}));
    GaussianProcess::gradientMode() = GPB_GRAD_AS_EXECUTED;
    std::vector<double> dr; for (auto& d : GaussianProcess::draw(xs, &Location::euclidean, p, z)) dr.push_back(d.y);
    vec("draw", dr);
  }

  // ---- error behaviour ---------------------------------------------------------------------------
  {
    std::vector<Data> mixed = {Data{One(0.0), 1.0}, Data{Two(1.0, 2.0), 0.5}};
    std::string e1 = thrown([&] { GaussianProcess::loglikelihood(mixed, &Location::euclidean); });
    DenseMatrix A(2, 2); A(0, 0) = 1; A(0, 1) = 2; A(1, 0) = 2; A(1, 1) = 1;
    std::string e2 = thrown([&] { cholesky(A); });
    std::vector<Data> two = {Data{One(0.0), 1.0}, Data{One(1.0), 0.5}};
    std::string e3 = thrown([&] { GaussianProcess::loglikelihood(two, &Location::euclidean)(Parameters{MeanParameters::zero(), {KernelParameters::matern(1, 1.5, 1)}}); });
    DistFn manhattan = [](const Location& a, const Location& b) { return std::fabs(a.x - b.x) + std::fabs(a.y - b.y); };
    std::string e4 = thrown([&] { GaussianProcess::distanceMatrix({One(0.0), One(1.0)}, manhattan); });
    std::string e5 = thrown([&] { GaussianProcess::loglikelihood(two, &Location::euclidean)(Parameters{MeanParameters::zero(), {KernelParameters::se(1.0, 50.0), KernelParameters::white(-2.0)}}); });
    std::printf("\"err_mixed_dims\": \"%s\", \"err_not_pd\": \"%s\", \"err_matern\": \"%s\", \"err_dist\": \"%s\", \"err_bad_proposal\": \"%s\",\n",
                e1.c_str(), e2.c_str(), e3.c_str(), e4.c_str(), e5.c_str());
  }
  std::printf("\"launches\": %lld\n}\n", gpb_launch_count(detail::context()));
  return 0;
}
