"""Write ``reftoys.npz``: the Beta-Bernoulli and Dirichlet-Categorical models of the REFERENCE's own
test-suite (``test/infer/test_mcmc.py:231-275`` ``test_beta_bernoulli_x64``, ``:278-305``
``test_dirichlet_categorical_x64``: Dirichlet log-density, ``StickBreakingTransform`` + log|J|,
``Categorical(probs)`` through clipped log-probabilities), evaluated by the reference's
code from /root/reference over the NumPy stand-in (``refshim_full``): ``potential_energy`` at fixed
unconstrained points (Beta through the Dirichlet log-density, ``SigmoidTransform`` + log|J|,
``Bernoulli(probs)`` with clamped probabilities), ``initialize_model`` draws for three keys in both
Threefry layouts, ``postprocess_fn`` at one point.

    python tests/golden/make_reftoys.py     (needs /root/reference; not for the GPU box)
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, HERE)

import refshim  # noqa: E402
import refshim_full  # noqa: E402

W = refshim._wrap


def make_data():
    rng = np.random.default_rng(1)
    return (rng.random((1000, 2)) < np.array([0.9, 0.1])).astype(np.float32)


def make_cat_data():
    rng = np.random.default_rng(2)
    return rng.choice(3, size=2000, p=[0.1, 0.6, 0.3]).astype(np.int32)


def main():
    numpyro = refshim_full.install_full(True)
    import jax.numpy as jnp
    import numpyro.distributions as dist
    from numpyro.infer.util import initialize_model, potential_energy
    from oracle.threefry import PRNGKey

    jax = sys.modules["jax"]
    zeros_grad = lambda f, **kw: (lambda x: (f(x), refshim.tree_map(lambda v: v * 0, x)))  # noqa: E731
    jax.value_and_grad = zeros_grad
    sys.modules["numpyro.infer.util"].value_and_grad = zeros_grad

    def model(data):  # test_mcmc.py:239-244
        alpha = jnp.array([1.1, 1.1])
        beta = jnp.array([1.1, 1.1])
        p_latent = numpyro.sample("p_latent", dist.Beta(alpha, beta))
        numpyro.sample("obs", dist.Bernoulli(p_latent), obs=data)
        return p_latent

    data = make_data()
    out = {"data": data}
    args = (W(data),)
    with numpyro.util.control_flow_prims_disabled(), np.errstate(all="ignore"):
        for part in (True, False):
            refshim.PARTITIONABLE = part
            for seed in (0, 1, 2):
                info = initialize_model(W(PRNGKey(seed)), model, model_args=args)
                out[f"init_{'part' if part else 'legacy'}_{seed}"] = np.asarray(
                    refshim.ravel_pytree(info.param_info.z)[0])
        refshim.PARTITIONABLE = True
        zs = np.array([[2.0, -2.0], [0.3, 0.7], [-1.5, 1.0], [2.4, -2.3], [5.0, -6.0]], np.float32)
        out["z"] = zs
        out["U"] = np.stack([np.asarray(potential_energy(model, args, {}, {"p_latent": W(z)})) for z in zs])
        out["constrained/p_latent"] = np.asarray(info.postprocess_fn({"p_latent": W(zs[0])})["p_latent"])
    print("U", out["U"], "p", out["constrained/p_latent"])

    def dc_model(data):  # test_mcmc.py:283-288
        concentration = jnp.array([1.0, 1.0, 1.0])
        p_latent = numpyro.sample("p_latent", dist.Dirichlet(concentration))
        numpyro.sample("obs", dist.Categorical(p_latent), obs=data)
        return p_latent

    jnp.take_along_axis = lambda a, i, axis=-1: W(np.take_along_axis(np.asarray(a), np.asarray(i), axis=axis))
    jnp.expand_dims = lambda a, axis: W(np.expand_dims(np.asarray(a), axis))
    jnp.pad = lambda a, pw, mode="constant", constant_values=0.0: W(
        np.pad(np.asarray(a), pw, mode=mode, constant_values=constant_values))
    cat = make_cat_data()
    out["dc/data"] = cat
    args = (W(cat),)
    with numpyro.util.control_flow_prims_disabled(), np.errstate(all="ignore"):
        for part in (True, False):
            refshim.PARTITIONABLE = part
            for seed in (0, 1, 2):
                info = initialize_model(W(PRNGKey(seed)), dc_model, model_args=args)
                out[f"dc/init_{'part' if part else 'legacy'}_{seed}"] = np.asarray(
                    refshim.ravel_pytree(info.param_info.z)[0])
        refshim.PARTITIONABLE = True
        zs = np.array([[-1.2, 0.8], [0.0, 0.0], [1.5, -1.0], [-2.0, 1.9], [4.0, -5.0]], np.float32)
        out["dc/z"] = zs
        out["dc/U"] = np.stack([np.asarray(potential_energy(dc_model, args, {}, {"p_latent": W(z)})) for z in zs])
        out["dc/constrained/p_latent"] = np.asarray(info.postprocess_fn({"p_latent": W(zs[0])})["p_latent"])
    print("dc U", out["dc/U"], "p", out["dc/constrained/p_latent"])
    np.savez_compressed(os.path.join(HERE, "reftoys.npz"), **out)


if __name__ == "__main__":
    main()
