// host build of csrc/stk_fastmath.cuh for tests/test_fastmath_cpu.py (accuracy against libm)
#include "../../stochaskell_b200/csrc/stk_fastmath.cuh"
#include <vector>
static std::vector<double>& tab() {
  static std::vector<double> t;
  if (t.empty()) { t.resize(stk::kFmTableDoubles); stk::fm_build_tables(t.data(), 0, 1); }
  return t;
}
extern "C" double fm_log(double x) { return stk::fast_log(x, tab().data()); }
extern "C" double fm_exp(double x) { return stk::fast_exp(x, tab().data()); }
