"""Write ``refmodels.npz``: the four registered models evaluated by the REFERENCE's own code.

    python tests/golden/make_refmodels.py       (needs /root/reference; not for the GPU box)

``refshim_full.install_full()`` imports the whole unmodified ``numpyro`` package over the NumPy
stand-in for JAX.  The model functions are the reference's own: ``examples/covtype.py`` ``model``,
``examples/horseshoe_regression.py`` ``model_normal_likelihood`` / ``model_bernoulli_likelihood``,
``examples/stochastic_volatility.py`` ``model`` (imported from those files), and the
non-centred eight schools of ``README.md:101-110`` (a README snippet, restated below).  Recorded:

* ``potential_energy(model, args, kwargs, z)`` (``infer/util.py:302-327``: handlers, ``biject_to``
  transforms + log|J| factors, reparam, the distributions' ``log_prob``) at fixed unconstrained
  points -> pins the potentials' VALUES (gradients: finite differences of these, in the test);
* ``initialize_model(key, model, init_strategy=init_to_uniform)`` (``infer/util.py:632-800``,
  ``find_valid_initial_params`` ``:335-477``): the initial point drawn for a key -> pins the
  site order and key consumption of the init search (validity is judged on the potential
  value only: there is no autodiff in the stand-in, the gradient it sees is zeros);
* ``postprocess_fn`` (``constrain_fn`` with deterministic sites) at those points.
"""
import importlib.util
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, HERE)

import refpin_cases as rc  # noqa: E402
import refshim  # noqa: E402
import refshim_full  # noqa: E402

W = refshim._wrap


def _example(name):
    for m in ("matplotlib", "matplotlib.dates", "matplotlib.pyplot"):  # plotting is not installed
        sys.modules.setdefault(m, refshim_full._Permissive(m))
    sys.modules["matplotlib"].use = lambda *a, **k: None
    path = os.path.join(refshim.REFERENCE, "examples", name + ".py")
    spec = importlib.util.spec_from_file_location("refexample_" + name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def models(numpyro, data):
    import numpyro.distributions as dist
    from numpyro.distributions.transforms import AffineTransform
    from numpyro.infer.reparam import TransformReparam

    def eight_schools_noncentered(J, sigma, y=None):  # README.md:101-110
        mu = numpyro.sample("mu", dist.Normal(0, 5))
        tau = numpyro.sample("tau", dist.HalfCauchy(5))
        with numpyro.plate("J", J):
            with numpyro.handlers.reparam(config={"theta": TransformReparam()}):
                theta = numpyro.sample(
                    "theta", dist.TransformedDistribution(dist.Normal(0.0, 1.0), AffineTransform(mu, tau)))
            numpyro.sample("obs", dist.Normal(theta, sigma), obs=y)

    hs = _example("horseshoe_regression")
    sv = _example("stochastic_volatility")
    cov = _example("covtype")
    f32 = lambda a: W(np.asarray(a, np.float32))  # noqa: E731
    return {
        "es": (eight_schools_noncentered, (8, f32(rc.S8)), {"y": f32(rc.Y8)}),
        "logit": (cov.model, (f32(data["logistic_X"]), W(data["logistic_y"].astype(np.int32))), {}),
        "hsb": (hs.model_bernoulli_likelihood, (f32(data["hs_X"]), f32(data["hs_Yb"])), {}),
        "hsn": (hs.model_normal_likelihood, (f32(data["hs_X"]), f32(data["hs_Yn"])), {}),
        "sv": (sv.model, (f32(data["sv_returns"]),), {}),
    }


# name: (model key, NUTS kwargs, MCMC kwargs, seed)
MCMC_CASES = {
    "es_2chains_thin": ("es", dict(), dict(num_warmup=50, num_samples=20, num_chains=2, thinning=2), 4),
    "es_fixed_2chains": ("es", dict(step_size=0.3, adapt_step_size=False, adapt_mass_matrix=False),
                         dict(num_warmup=10, num_samples=30, num_chains=2, thinning=3), 5),
    "logit_dense": ("logit", dict(dense_mass=True), dict(num_warmup=60, num_samples=10), 6),
    "hsn_depth6": ("hsn", dict(max_tree_depth=6), dict(num_warmup=40, num_samples=10), 7),
    "sv_depth6": ("sv", dict(max_tree_depth=6), dict(num_warmup=30, num_samples=8), 8),
}


def mcmc_fixture(numpyro, data, fams):
    """The reference's whole ``MCMC.run`` stack (``mcmc.py``: chain keys, ``_single_chain_mcmc``,
    ``fori_collect``, ``postprocess_fn``; ``NUTS(model)`` with its real ``initialize_model``) on the
    reference's own model functions.  The one substitution: ``value_and_grad(potential_fn)`` ->
    the oracle family's closed form (value and gradient; no autodiff in the stand-in, and the
    oracle's value is held to the reference's ``potential_energy`` by the entries above)."""
    from numpyro.infer import MCMC, NUTS
    from oracle.threefry import PRNGKey

    cur = {}

    def vag(f, **kw):
        def run(x):
            flat, unravel = refshim.ravel_pytree(x)
            U, g = cur["fam"].potential_and_grad(np.asarray(flat, np.float32))
            return W(np.float32(U)), unravel(np.asarray(g, np.float32))
        return run

    sys.modules["jax"].value_and_grad = vag
    for m in ("numpyro.infer.util", "numpyro.infer.hmc_util"):
        sys.modules[m].value_and_grad = vag
    mods = models(numpyro, data)
    out = {}
    refshim.PARTITIONABLE = True
    for name, (key, kkw, mkw, seed) in MCMC_CASES.items():
        model, args, kwargs = mods[key]
        cur["fam"] = fams[key]
        with numpyro.util.control_flow_prims_disabled(), np.errstate(all="ignore"):
            mcmc = MCMC(NUTS(model, **kkw), chain_method="sequential", progress_bar=False, **mkw)
            mcmc.run(W(PRNGKey(seed)), *args, extra_fields=("num_steps", "potential_energy", "diverging"),
                     **kwargs)
        for k, v in mcmc.get_samples(group_by_chain=True).items():
            out[f"mcmc/{name}/samples/{k}"] = np.asarray(v)
        for k, v in mcmc.get_extra_fields(group_by_chain=True).items():
            out[f"mcmc/{name}/extra/{k}"] = np.asarray(v)
        print("mcmc", name, {k: np.asarray(v).shape for k, v in mcmc.get_samples(group_by_chain=True).items()})
    return out


def errors_fixture(numpyro, data):
    """Argument checking of the reference's ``HMC`` / ``NUTS`` / ``MCMC`` (``hmc.py:643-658``,
    ``mcmc.py:342-354,672-678``): exception type and message, or the warning text."""
    import warnings

    from numpyro.infer import HMC, MCMC, NUTS

    model, args, kwargs = models(numpyro, data)["es"]
    key = W(np.array([0, 1], np.uint32))
    bad_init = {"mu": W(np.zeros(3, np.float32)), "tau": W(np.zeros(3, np.float32)),
                "theta_base": W(np.zeros((3, 8), np.float32))}
    calls = {
        "nuts_no_model": lambda: NUTS(),
        "nuts_model_and_potential": lambda: NUTS(model, potential_fn=lambda z: 0.0),
        "hmc_no_length": lambda: HMC(model, trajectory_length=None),
        "hmc_both_lengths_warns": lambda: HMC(model, num_steps=3),
        "mcmc_thinning": lambda: MCMC(NUTS(model), num_warmup=1, num_samples=1, thinning=0),
        "mcmc_chain_method": lambda: MCMC(NUTS(model), num_warmup=1, num_samples=1, chain_method="foo"),
        "mcmc_init_params_leading_dim": lambda: MCMC(
            NUTS(model), num_warmup=1, num_samples=1, num_chains=2, chain_method="sequential",
            progress_bar=False).run(key, *args, init_params=bad_init, **kwargs),
    }
    out = {}
    for name, f in calls.items():
        try:
            with warnings.catch_warnings(record=True) as w:
                warnings.simplefilter("always")
                f()
            rec = ["warning", str(w[0].message)] if w else ["ok", ""]
        except Exception as e:  # noqa: BLE001
            rec = [type(e).__name__, str(e)]
        out["errors/" + name] = np.array(rec)
        print("errors", name, rec)
    k = NUTS(model, step_size=1, max_tree_depth=(5, 10))
    out["attrs/nuts"] = np.array([repr(k._step_size), k.sample_field, repr(k.default_fields), k._algo,
                                  repr(k._max_tree_depth)])
    return out


def main():
    numpyro = refshim_full.install_full(True)
    from numpyro.infer.util import initialize_model, potential_energy
    from oracle.threefry import PRNGKey

    jax = sys.modules["jax"]
    zeros_grad = lambda f, **kw: (lambda x: (f(x), refshim.tree_map(lambda v: v * 0, x)))  # noqa: E731
    jax.value_and_grad = zeros_grad
    sys.modules["numpyro.infer.util"].value_and_grad = zeros_grad

    data = rc.make_data()
    fams = rc.families(data)
    out = {}
    for key, (model, args, kwargs) in models(numpyro, data).items():
        fam = fams[key]
        shapes = None
        with numpyro.util.control_flow_prims_disabled():
            for part in (True, False):
                refshim.PARTITIONABLE = part
                for seed in (0, 1, 2):
                    info = initialize_model(W(PRNGKey(seed)), model, model_args=args, model_kwargs=kwargs)
                    z0 = info.param_info.z
                    if shapes is None:
                        shapes = {k: np.shape(v) for k, v in z0.items()}
                    tag = "part" if part else "legacy"
                    out[f"{key}/init_{tag}_{seed}"] = np.asarray(refshim.ravel_pytree(z0)[0])
            refshim.PARTITIONABLE = True
            post = info.postprocess_fn
            zs, Us = [], []
            for seed in (0, 1, 2, 3):
                z = rc.init_point(fam, seed)
                z = (z + np.random.default_rng(seed).normal(size=z.shape) * 0.3).astype(np.float32)
                zd = {k: W(v.reshape(shapes[k])) for k, v in fam.unravel(z).items()}
                U = potential_energy(model, args, kwargs, zd)
                zs.append(z)
                Us.append(np.asarray(U))
                if seed == 0:
                    for k, v in post(zd).items():
                        out[f"{key}/constrained/{k}"] = np.asarray(v)
        out[f"{key}/z"] = np.stack(zs)
        out[f"{key}/U"] = np.stack(Us)
        out[f"{key}/sites"] = np.array(sorted(shapes))
        print(key, "U", [float(u) for u in Us], "sites", {k: shapes[k] for k in sorted(shapes)})
    out.update(mcmc_fixture(numpyro, data, fams))
    out.update(errors_fixture(numpyro, data))
    if "--check" in sys.argv:  # re-derive from the reference source and compare with the committed file
        old = np.load(os.path.join(HERE, "refmodels.npz"))
        bad = [k for k in out if k not in old.files or not np.array_equal(np.asarray(out[k]), old[k])]
        bad += [k for k in old.files if k not in out]
        print("check:", "identical" if not bad else f"DIFFERENT {bad[:8]}")
        sys.exit(1 if bad else 0)
    np.savez_compressed(os.path.join(HERE, "refmodels.npz"), **out)
    print("wrote refmodels.npz", os.path.getsize(os.path.join(HERE, "refmodels.npz")), "bytes")


if __name__ == "__main__":
    main()
