"""Tier 4: posterior means / sds agree within 4 MCSE and r_hat < 1.01.

(a) against the numbers the reference prints in its README for non-centred eight schools
    (README.md:117-143: mean, std, n_eff per site from the reference's own run);
(b) against the CPU oracle run with the same keys and configuration (eight schools and a
    small logistic regression).
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from oracle import hmc as ohmc  # noqa: E402
from oracle import potentials as opot  # noqa: E402
from oracle import threefry as T  # noqa: E402
from oracle.hmc_util import Config  # noqa: E402

Y8 = np.array([28.0, 8.0, -3.0, 7.0, -1.0, 1.0, 18.0, 12.0])
S8 = np.array([15.0, 10.0, 16.0, 11.0, 9.0, 11.0, 10.0, 18.0])

# README.md:121-139 (mean, std, n_eff)
README = {
    "mu": (4.08, 3.51, 720.43), "tau": (3.96, 3.31, 488.63),
    "theta": [(6.48, 5.72, 801.59), (4.95, 5.10, 1183.06), (3.65, 5.58, 581.31),
              (4.56, 5.04, 1282.60), (3.41, 4.79, 801.25), (3.58, 4.80, 1101.33),
              (6.31, 5.17, 1081.11), (4.81, 5.38, 954.14)],
    "theta_base": [(0.41, 0.95, 851.45), (0.15, 0.95, 1568.11), (-0.08, 0.98, 1037.16),
                   (0.06, 0.89, 1745.02), (-0.14, 0.94, 719.85), (-0.10, 0.96, 1128.45),
                   (0.38, 0.95, 1026.50), (0.10, 0.97, 1190.98)],
}


@pytest.fixture(scope="module")
def nb():
    import __graft_entry__ as g

    return g.build()


def _mcse(x, nb):
    """x: [chain, draw]; Monte Carlo standard error of the mean from the Geyer ESS."""
    ess = nb.diagnostics.effective_sample_size(x)
    return x.std(ddof=1) / np.sqrt(max(float(ess), 1.0)), float(ess)


def test_eight_schools_matches_reference_readme(nb):
    mc = nb.infer.MCMC(nb.infer.NUTS(nb.models.eight_schools_noncentered), num_warmup=500,
                       num_samples=500, num_chains=64, chain_method="vectorized")
    mc.run(nb.random.PRNGKey(0), 8, S8, y=Y8, extra_fields=("potential_energy",))
    s = mc.get_samples(group_by_chain=True)
    assert mc.get_extra_fields()["diverging"].mean() < 0.01
    checks = [("mu", s["mu"], README["mu"]), ("tau", s["tau"], README["tau"])]
    for j in range(8):
        checks.append((f"theta[{j}]", s["theta"][..., j], README["theta"][j]))
        checks.append((f"theta_base[{j}]", s["theta_base"][..., j], README["theta_base"][j]))
    for name, x, (mean, std, n_eff) in checks:
        x = np.asarray(x, np.float64)
        assert nb.diagnostics.split_gelman_rubin(x) < 1.01, name
        mcse_ours, _ = _mcse(x, nb)
        mcse_ref = std / np.sqrt(n_eff)
        tol = 4.0 * np.hypot(mcse_ours, mcse_ref) + 0.005  # README prints 2 decimals
        assert abs(x.mean() - mean) < tol, (name, x.mean(), mean, tol)
        assert abs(x.std() - std) < 0.15 * std + 0.01, (name, x.std(), std)
    # README.md:143: expected log joint density -46.09 (from a 1000-draw run)
    pe = mc.get_extra_fields()["potential_energy"]
    assert abs(np.mean(-pe) - (-46.09)) < 0.5


def _compare_with_oracle(nb, model, args, kwargs, fam, sites, chains, nw, ns, ns_dev=1000, **kw):
    """Tier 4 (north_star): means / sds of the CUDA chains within 4 MCSE of the oracle's,
    split r_hat < 1.01 on the CUDA chains.  The device draws ``ns_dev`` samples per chain (r_hat
    of 4 x 300 draws is itself noisy at the 1 % level); the oracle's run length is what its
    Python loops afford -- both enter the tolerance through their own MCSE."""
    key = T.PRNGKey(3)
    mc = nb.infer.MCMC(nb.infer.NUTS(model, **kw), num_warmup=nw, num_samples=ns_dev,
                       num_chains=chains, chain_method="vectorized")
    mc.run(key, *args, **kwargs)
    z_dev = mc.get_samples(group_by_chain=True)
    res, _, _ = ohmc.run_mcmc(fam, key, nw, ns, num_chains=chains, cfg=Config(np.float32, True),
                              **kw)
    z_ora = res["z"]  # [chain, draw, D] unconstrained, flat sorted-site order
    off = 0
    for name, size, unconstrain in sites:
        a = np.asarray(z_dev[name], np.float64).reshape(chains, ns_dev, size)
        if unconstrain:
            a = np.log(a)
        for j in range(size):
            x, y = a[..., j], np.asarray(z_ora[..., off + j], np.float64)
            assert nb.diagnostics.split_gelman_rubin(x) < 1.01, (name, j)
            mx, _ = _mcse(x, nb)
            my, _ = _mcse(y, nb)
            tol = 4.0 * np.hypot(mx, my)
            assert abs(x.mean() - y.mean()) < tol, (name, j, x.mean(), y.mean(), tol)
            assert abs(x.std() - y.std()) < 0.25 * y.std(), (name, j, x.std(), y.std())
        off += size


def test_eight_schools_matches_oracle_posterior(nb):
    _compare_with_oracle(nb, nb.models.eight_schools_noncentered, (8, S8), {"y": Y8},
                         opot.EightSchools(S8, Y8),
                         [("mu", 1, False), ("tau", 1, True), ("theta_base", 8, False)],
                         chains=4, nw=300, ns=300)


def test_logistic_matches_oracle_posterior(nb):
    rng = np.random.default_rng(12)
    N, D = 2000, 5
    X = rng.normal(size=(N, D)).astype(np.float32)
    X[:, -1] = 1.0
    y = rng.random(N) < 1 / (1 + np.exp(-X @ (rng.normal(size=D) * 0.7)))
    _compare_with_oracle(nb, nb.models.logistic_regression, (X, y), {}, opot.Logistic(X, y),
                         [("coefs", D, False)], chains=4, nw=300, ns=300)
