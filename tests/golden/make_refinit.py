"""Write ``refinit.npz``: ``initialize_model`` of the REFERENCE with ``init_to_value`` given only SOME
sites (``numpyro/infer/initialization.py`` ``init_to_value`` defers to ``init_to_uniform`` for the
others; ``infer/util.py:396-424``, the branch that traces the seeded model), executed from
/root/reference over the NumPy stand-in (``refshim_full``), eight schools non-centred.

    python tests/golden/make_refinit.py     (needs /root/reference; not for the GPU box)
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
CASES = {"mu": {"mu": 1.5}, "tau": {"tau": 2.0}, "mu_theta": {"mu": -0.5, "theta_base": np.linspace(-1, 1, 8)}}


def main():
    numpyro = refshim_full.install_full(True)
    import make_refmodels as mr
    from numpyro.infer.initialization import init_to_value
    from numpyro.infer.util import initialize_model
    from oracle.threefry import PRNGKey

    jax = sys.modules["jax"]
    zeros_grad = lambda f, **kw: (lambda x: (f(x), refshim.tree_map(lambda v: v * 0, x)))  # noqa: E731
    jax.value_and_grad = zeros_grad
    sys.modules["numpyro.infer.util"].value_and_grad = zeros_grad
    model, args, kwargs = mr.models(numpyro, rc.make_data())["es"]
    out = {}
    with numpyro.util.control_flow_prims_disabled(), np.errstate(all="ignore"):
        for name, vals in CASES.items():
            strat = init_to_value(values={k: W(np.asarray(v, np.float32)) for k, v in vals.items()})
            for part in (True, False):
                refshim.PARTITIONABLE = part
                for seed in (0, 1, 2):
                    info = initialize_model(W(PRNGKey(seed)), model, init_strategy=strat, model_args=args,
                                            model_kwargs=kwargs)
                    z = info.param_info.z
                    out[f"{name}/{'part' if part else 'legacy'}_{seed}"] = np.concatenate(
                        [np.ravel(np.asarray(z[k])) for k in sorted(z)]).astype(np.float32)
            print(name, out[f"{name}/part_0"])
    np.savez_compressed(os.path.join(HERE, "refinit.npz"), **out)


if __name__ == "__main__":
    main()
