// ref_cc_mvn.cc — TEST INFRASTRUCTURE.  The statement templates the reference's C++ backend emits for a GP marginal
// likelihood with per-evaluation hyper-parameters, on the reference's own vendored Eigen:
//   covariance  kernelSE' lsv lls2 (app/benchmark/Main.hs:65-74):  exp(lsv - distEucSq a b / (2 exp lls2)) + diag 0.01,
//               distEucSq a b = asColumn (square a) + asRow (square b) - 2 *> outer a b
//   multi_normal_lpdf (ccNode, CC.hs:221-225):  -0.5 * (C.llt().matrixL().solve(z - m).squaredNorm()
//                                                        + log(2*M_PI) * z.size() + LOG_DET(C))
//   LOG_DET (prelude macro, CC.hs:420-421)
//   priors lsv ~ normal 0 1, lls2 ~ normal (log 100) 2 (gpclas, Main.hs:84-91) through the generic rule CC.hs:226-228.
// Built by oracle/build_ref.sh into oracle/_ref/cc_mvn (headers under /root/reference/stochaskell/{eigen,boost}).
// usage: cc_mvn            stdin: n, x[n], y[n], then any number of "lsv lls2" points -> one %.17g lp per point
//        cc_mvn bench T    the same, but evaluates the FIRST point T times and prints "evals seconds" (CPU baseline)
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <stdexcept>
#include <boost/math/distributions.hpp>
#define eigen_assert(X) do { if(!(X)) throw std::runtime_error(#X); } while(false);
#include <Eigen/Dense>
using namespace std;
using namespace Eigen;
#define LOG_DET(m) ((m).isLowerTriangular() ? (m).diagonal().array().log().sum() : (m).llt().matrixL().toDenseMatrix().diagonal().array().log().sum() * 2)

static double lp_at(const VectorXd& x, const VectorXd& y, double lsv, double lls2) {
  const int n = x.size();
  MatrixXd d2 = x.array().square().matrix().replicate(1, n) + x.array().square().matrix().transpose().replicate(n, 1) - 2 * (x * x.transpose());
  MatrixXd C = (lsv - d2.array() / (2 * exp(lls2))).exp().matrix();
  C.diagonal().array() += 0.01;
  VectorXd m = VectorXd::Zero(n);
  double v = -0.5 * (C.llt().matrixL().solve((y - m)).squaredNorm() + log(2 * M_PI) * y.size() + LOG_DET(C));
  v += log(boost::math::pdf(boost::math::normal_distribution<>(0, 1), lsv));
  v += log(boost::math::pdf(boost::math::normal_distribution<>(log(100.), 2), lls2));
  return v;
}

int main(int argc, char** argv) {
  int n;
  if (!(cin >> n)) return 1;
  VectorXd x(n), y(n);
  for (int i = 0; i < n; ++i) cin >> x[i];
  for (int i = 0; i < n; ++i) cin >> y[i];
  double lsv, lls2;
  if (argc >= 3 && string(argv[1]) == "bench") {
    const int T = atoi(argv[2]);
    cin >> lsv >> lls2;
    auto t0 = chrono::steady_clock::now();
    double s = 0;
    for (int t = 0; t < T; ++t) s += lp_at(x, y, lsv + 1e-3 * t, lls2);
    const double dt = chrono::duration<double>(chrono::steady_clock::now() - t0).count();
    printf("%d %.6f %.17g\n", T, dt, s);
    return 0;
  }
  while (cin >> lsv >> lls2) printf("%.17g\n", lp_at(x, y, lsv, lls2));
  return 0;
}
