"""C5 DP mixture oracle (PARITY UNPINNED: the reference only forward-simulates dpmm through its Church backend and
has no inference for it).  Checks that exist: the forward model against the model's own definition
(dp.ipynb cells 1-3), and that the blocked Gibbs sampler recovers a planted mixture."""
import numpy as np


def planted(N, seed=3):
    rng = np.random.default_rng(seed)
    centers = np.array([12.0, 31.0, 48.0, 66.0, 87.0])
    comp = rng.integers(0, 5, N)
    return centers[comp] + rng.standard_normal(N), centers[comp]


def test_forward_simulation_follows_the_model(orc):
    N, K = 4000, 32
    v, mu, z, ys = orc.dp_simulate(N, K, seed=5, chain=2)
    assert v[-1] == 1.0 and np.all((v[:-1] > 0) & (v[:-1] < 1)) and np.all((mu > 0) & (mu < 100))
    pr = v * np.concatenate([[1.0], np.cumprod(1 - v[:-1])])
    assert abs(pr.sum() - 1) < 1e-12
    assert z.min() >= 1 and z.max() <= K
    freq = np.bincount(z, minlength=K + 1)[1:] / N
    assert np.max(np.abs(freq - pr)) < 4 * np.sqrt(0.25 / N)           # stick_j ~ pmf probs
    resid = ys - mu[z - 1]
    assert abs(resid.mean()) < 0.06 and abs(resid.std() - 1) < 0.05     # values_j ~ normal atoms[stick_j] 1
    # sticks ~ beta 1 alpha over many chains: mean 1 / (1 + alpha)
    v0 = np.array([orc.dp_simulate(2, K, seed=5, chain=c)[0][0] for c in range(2000)])
    assert abs(v0.mean() - 1 / 6) < 0.01


def test_gibbs_recovers_planted_mixture(orc):
    y, centers = planted(3000)
    K = 32
    v, mu, _, _ = orc.dp_simulate(2, K, seed=9, chain=0)
    for t in range(60):
        z = orc.dp_sweep(y, v, mu, seed=9, chain=0, sweep=t)
    occ = np.bincount(z, minlength=K + 1)[1:]
    # a planted component may be shared by several atoms (all sitting on its centre): judge per observation
    assert np.mean(np.abs(mu[z - 1] - centers) < 1.0) > 0.97
    assert len(np.unique(np.round(mu[z - 1] / 10))) >= 5
    # counts are exact integers and the fixed-point sums make the atom posterior means order-independent
    assert occ.sum() == len(y)


def test_sweep_is_a_pure_function_of_seed_chain_sweep(orc):
    y, _ = planted(501)
    v0, mu0, _, _ = orc.dp_simulate(2, 16, seed=1, chain=7)
    a = (v0.copy(), mu0.copy()); b = (v0.copy(), mu0.copy())
    za = orc.dp_sweep(y, a[0], a[1], 1, 7, 3); zb = orc.dp_sweep(y, b[0], b[1], 1, 7, 3)
    assert np.array_equal(za, zb) and np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    c = (v0.copy(), mu0.copy())
    zc = orc.dp_sweep(y, c[0], c[1], 1, 8, 3)
    assert not np.array_equal(za, zc)
