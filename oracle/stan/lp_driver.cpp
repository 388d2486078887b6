// lp_driver.cpp — TEST INFRASTRUCTURE.  Full-precision log-density + gradient from the reference's own HMC
// backend: links the C++ that the vendored stanc (CmdStan 2.20.0) generated for a model (-DMODEL_HPP=...)
// and calls stan::model::log_prob_grad — the function base_static_hmc::transition evaluates through
// `hamiltonian_.update_potential_gradient` — printing %.17g instead of `diagnose`'s 6 significant digits.
//   usage: lp_driver <data.R> <propto 0|1> < points      (one unconstrained point per line)
//   output per point: lp g_1 ... g_P
// Built only by oracle/stan/make_golden.py in the container that has /root/reference.
#include MODEL_HPP
#include <stan/io/dump.hpp>
#include <stan/model/log_prob_grad.hpp>
#include <cstdio>
#include <fstream>
#include <iostream>
#include <sstream>

int main(int argc, char** argv) {
  if (argc < 3) return 2;
  std::ifstream f(argv[1]);
  stan::io::dump ctx(f);
  stan_model model(ctx, 0, &std::cerr);
  const bool propto = atoi(argv[2]) != 0;
  std::string line;
  while (std::getline(std::cin, line)) {
    std::istringstream ls(line);
    std::vector<double> q; double v;
    while (ls >> v) q.push_back(v);
    if (q.empty()) continue;
    if (q.size() != model.num_params_r()) { fprintf(stderr, "expected %zu values\n", (size_t)model.num_params_r()); return 3; }
    std::vector<int> qi; std::vector<double> g;
    double lp = propto ? stan::model::log_prob_grad<true, true>(model, q, qi, g, &std::cerr)
                       : stan::model::log_prob_grad<false, true>(model, q, qi, g, &std::cerr);
    printf("%.17g", lp);
    for (double x : g) printf(" %.17g", x);
    printf("\n");
  }
  return 0;
}
