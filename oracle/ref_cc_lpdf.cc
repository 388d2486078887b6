// ref_cc_lpdf.cc — TEST INFRASTRUCTURE.  Evaluates the log-density statement templates of the reference's
// C++ backend (ccNode, CC.hs:198-228) with the reference's own vendored Boost.Math, so that the oracle's
// log-space formulas are pinned against "the CC backend's arithmetic".  Built by oracle/build_ref.sh into
// oracle/_ref/cc_lpdf from this file + the headers under /root/reference/stochaskell/boost.
// stdin: lines "<name> <x> <args...>";  stdout: one %.17g value per line.
#include <cmath>
#include <cstdio>
#include <iostream>
#include <sstream>
#include <string>
#include <boost/math/distributions.hpp>
#include <boost/math/special_functions/factorials.hpp>
using namespace std;
int main() {
  string line;
  while (getline(cin, line)) {
    istringstream ls(line);
    string f; double x, a = 0, b = 0;
    if (!(ls >> f >> x)) continue;
    ls >> a >> b;
    double v = NAN;
    try {
      if (f == "gamma_lpdf") v = log(boost::math::pdf(boost::math::gamma_distribution<>(a, 1. / b), x));                       // CC.hs:206-209
      else if (f == "neg_binomial_lpdf") v = log(boost::math::pdf(boost::math::negative_binomial_distribution<>(a, b / (b + 1)), x));  // :210-213
      else if (f == "normal_lpdf") v = log(boost::math::pdf(boost::math::normal_distribution<>(a, b), x));                     // :226-228 (generic rule)
      else if (f == "poisson_lpdf") v = log(boost::math::pdf(boost::math::poisson_distribution<>(a), x));
      else if (f == "cauchy_lpdf") v = log(boost::math::pdf(boost::math::cauchy_distribution<>(a, b), x));
      else if (f == "uniform_lpdf") v = log(boost::math::pdf(boost::math::uniform_distribution<>(a, b), x));
      else if (f == "logistic_lpdf") v = log(boost::math::pdf(boost::math::logistic_distribution<>(a, b), x));
      else if (f == "beta_lpdf") v = log(boost::math::pdf(boost::math::beta_distribution<>(a, b), x));
      else if (f == "bernoulliLogit_lpdf") v = x ? log(1. / (1 + exp(-a))) : log(1 - 1. / (1 + exp(-a)));                       // :198-200  (checked form)
      else if (f == "logFactorial") v = lgamma(1 + x);                                                                         // :167-168
    } catch (const std::exception& e) { v = NAN; }
    printf("%.17g\n", v);
  }
}
