"""Write ``refpredict.npz``: ``log_likelihood`` and ``Predictive`` of the REFERENCE
(``numpyro/infer/util.py:803-1148``) on its own model functions, executed from /root/reference
over the NumPy stand-in (``refshim_full``; ``soft_vmap``'s batching is emulated as a loop).

    python tests/golden/make_refpredict.py      (needs /root/reference; not for the GPU box)

Per model: constrained posterior draws (fixed, S = 6), ``log_likelihood(model, draws, *args)``
and, where every observation site can be sampled with the emulated ``jax.random``
(Bernoulli, Normal), ``Predictive(model, draws)(PRNGKey(3), *args_without_observations)`` in
both Threefry layouts.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, HERE)

import refpin_cases as rc  # noqa: E402
import refshim  # noqa: E402
import refshim_full  # noqa: E402

W = refshim._wrap
S = 6


def draws(data):
    rng = np.random.default_rng(11)
    f = lambda *s: rng.normal(size=s).astype(np.float32)  # noqa: E731
    Dl, Dh, T = data["logistic_X"].shape[1], data["hs_X"].shape[1], data["sv_returns"].shape[0]
    return {
        "es": {"mu": f(S) * 3 + 4, "tau": np.exp(f(S)), "theta_base": f(S, 8)},
        "logit": {"coefs": f(S, Dl) * 0.7},
        "hsb": {"lambdas": np.exp(f(S, Dh)), "tau": np.exp(f(S, 1) - 1), "unscaled_betas": f(S, Dh)},
        "hsn": {"lambdas": np.exp(f(S, Dh)), "tau": np.exp(f(S, 1) - 1), "unscaled_betas": f(S, Dh),
                "prec_obs": np.exp(f(S) + 1)},
        "sv": {"sigma": np.exp(f(S) - 3), "nu": np.exp(f(S) + 2), "s": f(S, T) * 0.3 - 4.5},
    }


def main():
    numpyro = refshim_full.install_full()
    from numpyro.infer.util import Predictive, log_likelihood
    import make_refmodels as mr

    data = rc.make_data()
    mods = mr.models(numpyro, data)
    post = draws(data)
    out = {}
    for key, (model, args, kwargs) in mods.items():
        p = {k: W(v) for k, v in post[key].items()}
        for k, v in post[key].items():
            out[f"{key}/draws/{k}"] = v
        with np.errstate(all="ignore"):
            ll = log_likelihood(model, p, *args, **kwargs)
        for k, v in ll.items():
            out[f"{key}/loglik/{k}"] = np.asarray(v)
        if key == "sv":
            continue  # StudentT sampling goes through jax.random.gamma (rejection sampler): not emulated
        # the observations are the last positional argument / the `y` keyword: dropped for Predictive
        pargs = args if key == "es" else args[:-1] + (None,)
        pmodel = model
        if key == "logit":
            # examples/covtype.py's model indexes `labels[idx]` and so cannot run unobserved; the
            # predictive fixture uses the same two sites without the subsample plate
            import numpyro.distributions as dist
            import jax.numpy as jnp

            def pmodel(data, labels=None):
                coefs = numpyro.sample("coefs", dist.Normal(jnp.zeros(data.shape[1]), jnp.ones(data.shape[1])))
                return numpyro.sample("obs", dist.Bernoulli(logits=jnp.dot(data, coefs)), obs=labels)
        for part in (True, False):
            refshim.PARTITIONABLE = part
            pred = Predictive(pmodel, p)(W(np.array([0, 3], np.uint32)), *pargs)
            for k, v in pred.items():
                out[f"{key}/predictive/{'part' if part else 'legacy'}/{k}"] = np.asarray(v)
        refshim.PARTITIONABLE = True
        print(key, {k: np.asarray(v).shape for k, v in ll.items()}, {k: np.asarray(v).shape for k, v in pred.items()})
    np.savez_compressed(os.path.join(HERE, "refpredict.npz"), **out)
    print("wrote refpredict.npz:", len(out), "arrays")


if __name__ == "__main__":
    main()
