// C1 linreg, d-dimensional extension of linreg.ipynb cell 1/7, written in the
// shape Language.Stan.stanProgram emits (data = given PNodes, parameters =
// latent PNodes with <lower=0> from SubrangeT, one `~` per PNode).
// TEST INFRASTRUCTURE (oracle): compiled with the vendored CmdStan 2.20 by
// oracle/stan/build_ref.sh; never part of the product path.
data {
  int<lower=1> n;
  int<lower=1> d;
  matrix[n,d] x_stan_0_0;
  vector[n] x_stan_0_4;
}
parameters {
  real x_stan_0_1;
  vector[d] x_stan_0_2;
  real<lower=0> x_stan_0_3;
}
model {
  vector[n] v_0_3;
  vector[n] v_0_4;
  vector[n] v_0_6;
  v_0_3 = x_stan_0_0 * x_stan_0_2;
  v_0_4 = x_stan_0_1 + v_0_3;
  for (i_1_1 in 1:n) {
    v_0_6[i_1_1] = x_stan_0_3;
  }
  x_stan_0_1 ~ normal(0, 10);
  x_stan_0_2 ~ normal(0, 10);
  x_stan_0_3 ~ cauchy(0, 5) T[0,];
  x_stan_0_4 ~ normal(v_0_4, v_0_6);
}
