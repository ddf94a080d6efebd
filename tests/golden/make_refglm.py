"""Write ``refglm.npz``: the two GLMs of the registry written with the REFERENCE's
primitives and evaluated by the reference's own ``potential_energy`` / ``initialize_model`` /
``constrain_fn`` from /root/reference over the NumPy stand-in (``refshim_full``).

    python tests/golden/make_refglm.py     (needs /root/reference; not for the GPU box)
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


def make_data(seed=3, N=400, D=6):
    rng = np.random.default_rng(seed)
    X = (rng.normal(size=(N, D)) * 0.5).astype(np.float32)
    w = np.array([0.8, -0.5, 0.3, 0.0, 0.2, -0.1], np.float32)[:D]
    counts = rng.poisson(np.exp(X @ w + 0.3)).astype(np.float32)
    y = (X @ w + 0.4 * rng.normal(size=N)).astype(np.float32)
    return X, counts, y


def main():
    numpyro = refshim_full.install_full(True)
    import numpyro.distributions as dist
    from numpyro.infer.util import constrain_fn, initialize_model, potential_energy
    from oracle.threefry import PRNGKey

    jax = sys.modules["jax"]
    jnp = sys.modules["jax.numpy"]

    def poisson_regression(data, counts):
        dim = data.shape[1]
        coefs = numpyro.sample("coefs", dist.Normal(jnp.zeros(dim), jnp.ones(dim)))
        return numpyro.sample("obs", dist.Poisson(jnp.exp(jnp.dot(data, coefs))), obs=counts)

    def linear_regression(data, y):
        dim = data.shape[1]
        coefs = numpyro.sample("coefs", dist.Normal(jnp.zeros(dim), jnp.ones(dim)))
        prec_obs = numpyro.sample("prec_obs", dist.Gamma(3.0, 1.0))
        sigma_obs = 1.0 / jnp.sqrt(prec_obs)
        return numpyro.sample("obs", dist.Normal(jnp.dot(data, coefs), sigma_obs), obs=y)

    zeros_grad = lambda f, **kw: (lambda x: (f(x), refshim.tree_map(lambda v: v * 0, x)))  # noqa: E731
    jax.value_and_grad = zeros_grad
    sys.modules["numpyro.infer.util"].value_and_grad = zeros_grad
    X, counts, y = make_data()
    out = {"X": X, "counts": counts, "y": y}
    rng = np.random.default_rng(11)
    with numpyro.util.control_flow_prims_disabled(), np.errstate(all="ignore"):
        for name, model, obs, D in (("poisson", poisson_regression, counts, 6), ("normal", linear_regression, y, 7)):
            args = (W(X), W(obs))
            pts = (rng.normal(size=(4, D)) * 0.4).astype(np.float32)
            U = []
            for z in pts:
                sites = {"coefs": W(z[:6])}
                if name == "normal":
                    sites["prec_obs"] = W(z[6])
                U.append(np.float32(np.asarray(potential_energy(model, args, {}, sites))))
            out[f"{name}/z"], out[f"{name}/U"] = pts, np.asarray(U, np.float32)
            sites = {"coefs": W(pts[0][:6])}
            if name == "normal":
                sites["prec_obs"] = W(pts[0][6])
            c = constrain_fn(model, args, {}, sites, return_deterministic=True)
            for k, v in c.items():
                out[f"{name}/constrained/{k}"] = np.asarray(v, np.float32)
            if name == "normal":  # log_likelihood (infer/util.py:1094-1148) over the four points as draws
                from numpyro.infer.util import log_likelihood

                samples = {"coefs": W(pts[:, :6]), "prec_obs": W(np.exp(pts[:, 6]).astype(np.float32))}
                ll = log_likelihood(model, samples, *args)
                out["normal/loglik"] = np.asarray(ll["obs"], np.float32)
            for part in (True, False):
                refshim.PARTITIONABLE = part
                for seed in (0, 1, 2):
                    info = initialize_model(W(PRNGKey(seed)), model, model_args=args)
                    zi = info.param_info.z
                    out[f"{name}/init_{'part' if part else 'legacy'}_{seed}"] = np.concatenate(
                        [np.ravel(np.asarray(zi[k])) for k in sorted(zi)]).astype(np.float32)
            print(name, out[f"{name}/U"], out[f"{name}/init_part_0"][:3])
    np.savez_compressed(os.path.join(HERE, "refglm.npz"), **out)


if __name__ == "__main__":
    main()
