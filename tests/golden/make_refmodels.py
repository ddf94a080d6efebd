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
    np.savez_compressed(os.path.join(HERE, "refmodels.npz"), **out)
    print("wrote refmodels.npz", os.path.getsize(os.path.join(HERE, "refmodels.npz")), "bytes")


if __name__ == "__main__":
    main()
