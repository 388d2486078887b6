// ref_soccer_loop.cc — TEST / BENCH INFRASTRUCTURE.  In-process loop around the reference's own generated C++ sampler
// (the program Language.CC emitted for `soccer n `rjmcC` soccerJump`, rj.ipynb cell 10, extracted to oracle/_ref/rj_soccer.cc
// by oracle/build_ref.sh; never copied into the repository).  compileCC's closure spawns that program once per transition and
// talks to it through text (CC.hs:494-498); this driver keeps the generated main() exactly as it is (renamed by the
// preprocessor), feeds its stdin/stdout through string streams and calls it T times in one process, so that the
// arithmetic + text-format cost of a transition can be timed without the process spawn.
//   usage: rj_soccer_loop T < initial-state-line      prints: <transitions> <seconds> then the final state line
#include <chrono>
#include <iostream>
#include <sstream>
#include <string>
// `int main() {` becomes `int rj_unused_; void rj_soccer_generated_main() {`: the generated body ends without a return statement,
// which is only defined behaviour for main itself (as an int function g++ -O2 would fall off its end)
#define main() rj_unused_; void rj_soccer_generated_main()
#include RJ_SOCCER_CC
#undef main

int main(int argc, char** argv) {
  long T = argc > 1 ? atol(argv[1]) : 1000;
  std::string state;
  std::getline(std::cin, state);
  std::streambuf* in0 = std::cin.rdbuf(); std::streambuf* out0 = std::cout.rdbuf();
  auto t0 = std::chrono::steady_clock::now();
  for (long t = 0; t < T; ++t) {
    std::istringstream is(state + "\n"); std::ostringstream os;
    std::cin.rdbuf(is.rdbuf()); std::cout.rdbuf(os.rdbuf());
    std::cin.clear();
    rj_soccer_generated_main();
    state = os.str();
    while (!state.empty() && (state.back() == '\n' || state.back() == ' ')) state.pop_back();
  }
  double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  std::cin.rdbuf(in0); std::cout.rdbuf(out0);
  std::cout << T << " " << dt << "\n" << state << std::endl;
  return 0;
}
