"""Write ``refgibbs.npz``: the REFERENCE's ``HMCGibbs`` (``numpyro/infer/hmc_gibbs.py:36-190``)
around the reference's ``NUTS(model)`` on the non-centred eight schools with the Gibbs site ``mu``,
executed from /root/reference over the NumPy stand-in (``refshim_full``): ``init`` and every
``sample`` state; plus the reference's ``potential_energy`` of the wrapped model (Gibbs sites
conditioned + substituted, ``:30-34``) at fixed points, for ``mu`` and for the positive site
``tau``.  The one substitution (as in ``make_refmodels.py``): ``value_and_grad(potential_fn)`` ->
the oracle's closed form of the conditioned family; its VALUE is checked here against the
reference's own model code at every call.

    python tests/golden/make_refgibbs.py     (needs /root/reference; not for the GPU box)
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, HERE)

import refgibbs_cases as gc  # noqa: E402
import refpin_cases as rc  # noqa: E402
import refshim  # noqa: E402
import refshim_full  # noqa: E402

W = refshim._wrap


def main():
    numpyro = refshim_full.install_full(True)
    import make_refmodels as mr
    import numpyro.infer.hmc_gibbs as hg
    from numpyro.infer import NUTS
    from numpyro.infer.util import potential_energy
    from oracle.gibbs import Conditioned
    from oracle.threefry import PRNGKey

    jax = sys.modules["jax"]
    fam_full = rc.families(rc.make_data())["es"]
    cur = {"gibbs": None, "fam": None, "max_rel": 0.0, "calls": 0}

    real_wrap = hg._wrap_model

    def spy_wrap(model, *args, **kwargs):  # every run of the wrapped model tells the current values
        cur["gibbs"] = {k: np.asarray(v, np.float32) for k, v in kwargs.get("_gibbs_sites", {}).items()}
        return real_wrap(model, *args, **kwargs)

    hg._wrap_model = spy_wrap

    def vag(f, **kw):
        def run(x):
            flat, unravel = refshim.ravel_pytree(x)
            with np.errstate(all="ignore"):
                v_ref = float(np.asarray(f(x)))  # the reference's model code (sets cur["gibbs"])
            cur["fam"].set_values(cur["gibbs"])
            U, g = cur["fam"].potential_and_grad(np.asarray(flat, np.float32))
            if np.isfinite(v_ref) and np.isfinite(U):
                cur["max_rel"] = max(cur["max_rel"], abs(v_ref - float(U)) / max(1.0, abs(v_ref)))
            cur["calls"] += 1
            return W(np.float32(U)), unravel(np.asarray(g, np.float32))
        return run

    jax.value_and_grad = vag
    for m in ("numpyro.infer.util", "numpyro.infer.hmc_util", "numpyro.infer.hmc_gibbs"):
        sys.modules[m].value_and_grad = vag
    model, args, kwargs = mr.models(numpyro, rc.make_data())["es"]
    out = {}
    gibbs_fn = gc.make_gibbs_fn(lambda key: jax.random.normal(key), W)
    with numpyro.util.control_flow_prims_disabled(), np.errstate(all="ignore"):
        for part in (True, False):
            refshim.PARTITIONABLE = part
            for seed in gc.SEEDS:
                cur["fam"] = Conditioned(fam_full, ["mu"], positive=("tau",))
                kernel = hg.HMCGibbs(NUTS(model, **gc.KERNEL_KW), gibbs_fn=gibbs_fn, gibbs_sites=["mu"])
                state = kernel.init(W(PRNGKey(seed)), gc.NUM_WARMUP, None, args, kwargs)
                rows = {k: [] for k in ("mu", "z", "pe", "num_steps", "rng_key", "step_size", "accept_prob", "i")}

                def rec(st):
                    h = st.hmc_state
                    rows["mu"].append(np.asarray(st.z["mu"], np.float32))
                    rows["z"].append(np.concatenate([np.ravel(np.asarray(h.z[k], np.float32)) for k in sorted(h.z)]))
                    rows["pe"].append(np.float32(np.asarray(h.potential_energy)))
                    rows["num_steps"].append(int(np.asarray(h.num_steps)))
                    rows["rng_key"].append(np.asarray(st.rng_key, np.uint32))
                    rows["step_size"].append(np.float32(np.asarray(h.adapt_state.step_size)))
                    rows["accept_prob"].append(np.float32(np.asarray(h.accept_prob)))
                    rows["i"].append(int(np.asarray(h.i)))

                rec(state)
                for _ in range(gc.NUM_STEPS):
                    state = kernel.sample(state, args, kwargs)
                    rec(state)
                tag = f"mu/{'part' if part else 'legacy'}_{seed}"
                for k, v in rows.items():
                    out[f"{tag}/{k}"] = np.asarray(v)
                print(tag, "tree sizes", rows["num_steps"], "mu[-1]", rows["mu"][-1])
        # HMC as the inner kernel
        from numpyro.infer import HMC

        refshim.PARTITIONABLE = True
        cur["fam"] = Conditioned(fam_full, ["mu"], positive=("tau",))
        kernel = hg.HMCGibbs(HMC(model, **gc.HMC_KW), gibbs_fn=gibbs_fn, gibbs_sites=["mu"])
        state = kernel.init(W(PRNGKey(gc.HMC_SEED)), gc.HMC_WARMUP, None, args, kwargs)
        rows = {k: [] for k in ("mu", "z", "pe", "num_steps", "rng_key", "step_size", "accept_prob")}

        def rec2(st):
            h = st.hmc_state
            rows["mu"].append(np.asarray(st.z["mu"], np.float32))
            rows["z"].append(np.concatenate([np.ravel(np.asarray(h.z[k], np.float32)) for k in sorted(h.z)]))
            rows["pe"].append(np.float32(np.asarray(h.potential_energy)))
            rows["num_steps"].append(int(np.asarray(h.num_steps)))
            rows["rng_key"].append(np.asarray(st.rng_key, np.uint32))
            rows["step_size"].append(np.float32(np.asarray(h.adapt_state.step_size)))
            rows["accept_prob"].append(np.float32(np.asarray(h.accept_prob)))

        rec2(state)
        for _ in range(gc.HMC_STEPS):
            state = kernel.sample(state, args, kwargs)
            rec2(state)
        for k, v in rows.items():
            out[f"hmc/{k}"] = np.asarray(v)
        print("hmc inner: accept", rows["accept_prob"], "mu[-1]", rows["mu"][-1])
        # the conditioned potential itself, reference model code only
        refshim.PARTITIONABLE = True
        wrapped = hg.partial(real_wrap, model) if hasattr(hg, "partial") else None
        from functools import partial
        wrapped = partial(real_wrap, model)
        rng = np.random.default_rng(5)
        for name, vals in (("mu", [0.7, -3.0, 4.4]), ("tau", [0.3, 1.7, 6.0])):
            pts, U = [], []
            for v in vals:
                full = rng.normal(size=10).astype(np.float32)
                sites = {"mu": full[0], "tau": full[1], "theta_base": full[2:]}
                free = {k: W(np.asarray(x, np.float32)) for k, x in sites.items() if k != name}
                kw = dict(kwargs)
                kw["_gibbs_sites"] = {name: W(np.float32(v))}
                U.append(np.float32(np.asarray(potential_energy(wrapped, args, kw, free))))
                pts.append(np.concatenate([np.ravel(np.asarray(free[k])) for k in sorted(free)]))
            out[f"cond/{name}/values"] = np.asarray(vals, np.float32)
            out[f"cond/{name}/z"] = np.asarray(pts, np.float32)
            out[f"cond/{name}/U"] = np.asarray(U, np.float32)
            print("cond", name, out[f"cond/{name}/U"])
    print("value_and_grad calls", cur["calls"], "max rel. diff oracle vs reference model code", cur["max_rel"])
    assert cur["max_rel"] < 5e-6
    out["meta/max_rel_value_diff"] = np.float64(cur["max_rel"])
    np.savez_compressed(os.path.join(HERE, "refgibbs.npz"), **out)


if __name__ == "__main__":
    main()
