"""Write ``reflocscale.npz``: eight schools under the REFERENCE's ``LocScaleReparam``
(``numpyro/infer/reparam.py:64-145``), executed from /root/reference over the NumPy stand-in
(``refshim_full``): the centred model of ``README.md`` (``theta ~ Normal(mu, tau)``) wrapped in
``reparam(config={"theta": LocScaleReparam(centered)})`` for ``centered`` in {0, 0.3, 1}.

    python tests/golden/make_reflocscale.py     (needs /root/reference; not for the GPU box)

Per value of ``centered``: ``potential_energy`` at four unconstrained points, the init points of
``initialize_model`` (three keys, both Threefry layouts: site order and key consumption) and
``postprocess_fn`` (deterministic ``theta``) at one point.
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
CENTERED = (0.0, 0.3, 1.0)


def main():
    numpyro = refshim_full.install_full(True)
    import numpyro.distributions as dist
    from numpyro.infer.reparam import LocScaleReparam
    from numpyro.infer.util import initialize_model, potential_energy
    from oracle.threefry import PRNGKey

    jax = sys.modules["jax"]
    zeros_grad = lambda f, **kw: (lambda x: (f(x), refshim.tree_map(lambda v: v * 0, x)))  # noqa: E731
    jax.value_and_grad = zeros_grad
    sys.modules["numpyro.infer.util"].value_and_grad = zeros_grad
    f32 = lambda a: W(np.asarray(a, np.float32))  # noqa: E731

    def model_for(c):
        def eight_schools(J, sigma, y=None):
            mu = numpyro.sample("mu", dist.Normal(0, 5))
            tau = numpyro.sample("tau", dist.HalfCauchy(5))
            with numpyro.plate("J", J):
                with numpyro.handlers.reparam(config={"theta": LocScaleReparam(c)}):
                    theta = numpyro.sample("theta", dist.Normal(mu, tau))
                numpyro.sample("obs", dist.Normal(theta, sigma), obs=y)
        return eight_schools

    out = {}
    args, kwargs = (8, f32(rc.S8)), {"y": f32(rc.Y8)}
    for c in CENTERED:
        model, tag = model_for(c), f"c{c:g}"
        site = "theta" if c == 1.0 else "theta_decentered"
        with numpyro.util.control_flow_prims_disabled():
            for part in (True, False):
                refshim.PARTITIONABLE = part
                for seed in (0, 1, 2):
                    info = initialize_model(W(PRNGKey(seed)), model, model_args=args, model_kwargs=kwargs)
                    z0 = info.param_info.z
                    assert sorted(z0) == sorted(["mu", "tau", site]), sorted(z0)
                    out[f"{tag}/init_{'part' if part else 'legacy'}_{seed}"] = np.asarray(
                        refshim.ravel_pytree(z0)[0])
            refshim.PARTITIONABLE = True
            zs, Us = [], []
            for seed in range(4):
                z = (np.random.default_rng(seed).normal(size=10) * np.r_[3, 1, np.ones(8)]).astype(np.float32)
                zd = {"mu": W(z[0]), "tau": W(z[1]), site: W(z[2:])}
                zs.append(z)
                Us.append(np.asarray(potential_energy(model, args, kwargs, zd)))
                if seed == 0:
                    for k, v in info.postprocess_fn(zd).items():
                        out[f"{tag}/constrained/{k}"] = np.asarray(v)
        out[f"{tag}/z"], out[f"{tag}/U"] = np.stack(zs), np.stack(Us)
        print(tag, "U", [float(u) for u in Us], sorted(k for k in out if k.startswith(tag + "/constrained")))
    np.savez_compressed(os.path.join(HERE, "reflocscale.npz"), **out)
    print("wrote reflocscale.npz")


if __name__ == "__main__":
    main()
