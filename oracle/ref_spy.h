// ref_spy.h — TEST INFRASTRUCTURE.  Force-included (g++ -include) when the reference's own generated
// C++ samplers (the output of Language.CC, CC.hs:398-432) are compiled under oracle/_ref/.
// The generated program seeds std::mt19937 from std::random_device (CC.hs:424-425) and prints its
// state with 6 significant digits (CC.hs:379-384), so its arithmetic cannot be pinned as is.  This
// shim leaves every generated statement untouched and only
//   * replaces `random_device` by a source that returns $STK_REF_SEED (reproducible runs), and
//   * wraps the std distributions the generator emits (CC.hs:276-296) so that every draw — its
//     parameters and its value — is logged to stderr with 17 significant digits.
// The bernoulli parameter of an rjmc kernel IS the acceptance probability min(1, ratio)
// (Program.hs:1278-1283), so the log gives exact known-answer vectors
//   (input state, proposal draws) -> acceptance probability
// computed by the reference's own code generator + Boost/Eigen.
#pragma once
#include <cstdio>
#include <cstdlib>
#include <random>
namespace std {
struct stk_spy_random_device {
  unsigned operator()() const { const char* s = getenv("STK_REF_SEED"); return s ? (unsigned)strtoul(s, 0, 10) : 5489u; }
};
template <class T = double> struct stk_spy_normal_distribution {
  T m, s; stk_spy_normal_distribution(T m_ = 0, T s_ = 1) : m(m_), s(s_) {}
  template <class G> T operator()(G& g) { T v = normal_distribution<T>(m, s)(g); fprintf(stderr, "draw normal %.17g %.17g = %.17g\n", (double)m, (double)s, (double)v); return v; }
};
template <class T = double> struct stk_spy_uniform_real_distribution {
  T a, b; stk_spy_uniform_real_distribution(T a_ = 0, T b_ = 1) : a(a_), b(b_) {}
  template <class G> T operator()(G& g) { T v = uniform_real_distribution<T>(a, b)(g); fprintf(stderr, "draw uniform %.17g %.17g = %.17g\n", (double)a, (double)b, (double)v); return v; }
};
template <class T = double> struct stk_spy_gamma_distribution {
  T a, b; stk_spy_gamma_distribution(T a_ = 1, T b_ = 1) : a(a_), b(b_) {}
  template <class G> T operator()(G& g) { T v = gamma_distribution<T>(a, b)(g); fprintf(stderr, "draw gamma %.17g %.17g = %.17g\n", (double)a, (double)b, (double)v); return v; }
};
struct stk_spy_bernoulli_distribution {
  double p; stk_spy_bernoulli_distribution(double p_ = 0.5) : p(p_) {}
  template <class G> bool operator()(G& g) { bool v = bernoulli_distribution(p)(g); fprintf(stderr, "draw bernoulli %.17g = %d\n", p, (int)v); return v; }
};
template <class T = int> struct stk_spy_poisson_distribution {
  double l; stk_spy_poisson_distribution(double l_ = 1) : l(l_) {}
  template <class G> T operator()(G& g) { T v = poisson_distribution<T>(l)(g); fprintf(stderr, "draw poisson %.17g = %d\n", l, (int)v); return v; }
};
template <class T = int> struct stk_spy_uniform_int_distribution {
  T a, b; stk_spy_uniform_int_distribution(T a_ = 0, T b_ = 1) : a(a_), b(b_) {}
  template <class G> T operator()(G& g) { T v = uniform_int_distribution<T>(a, b)(g); fprintf(stderr, "draw uniform_int %d %d = %d\n", (int)a, (int)b, (int)v); return v; }
};
}  // namespace std
// Boost.Math headers are pulled in first so that its own classes of the same names keep their names.
#include <boost/math/distributions.hpp>
#include <boost/math/special_functions/factorials.hpp>
namespace boost { namespace math {  // ... and stay reachable under the macro-renamed spellings
template <class T = double, class P = policies::policy<> > using stk_spy_normal_distribution = normal_distribution<T, P>;
template <class T = double, class P = policies::policy<> > using stk_spy_gamma_distribution = gamma_distribution<T, P>;
template <class T = double, class P = policies::policy<> > using stk_spy_poisson_distribution = poisson_distribution<T, P>;
template <class T = double, class P = policies::policy<> > using stk_spy_bernoulli_distribution = bernoulli_distribution<T, P>;
} }
#define random_device stk_spy_random_device
#define normal_distribution stk_spy_normal_distribution
#define uniform_real_distribution stk_spy_uniform_real_distribution
#define gamma_distribution stk_spy_gamma_distribution
#define bernoulli_distribution stk_spy_bernoulli_distribution
#define poisson_distribution stk_spy_poisson_distribution
#define uniform_int_distribution stk_spy_uniform_int_distribution
