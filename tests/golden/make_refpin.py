"""Write ``refpin.npz``: traces produced by the REFERENCE's own sampler source.

    python tests/golden/make_refpin.py          (needs /root/reference; not for the GPU box)

``refshim.install()`` executes the unmodified ``numpyro/util.py``, ``numpyro/infer/hmc_util.py``
and ``numpyro/infer/hmc.py`` from /root/reference over a NumPy stand-in for the slice of JAX
they use (see refshim.py for exactly what is and is not reference code in such a run).  Every
case of ``refpin_cases.CASES`` is driven through the reference's ``HMC``/``NUTS`` classes:
``kernel.init(key, num_warmup, init_params=z0)`` then ``kernel.sample`` once per transition, in
the reference's eager mode (``numpyro.util.control_flow_prims_disabled``).  The recorded traces
are what ``tests/test_refpin.py`` holds the oracle (CPU) and the CUDA path (GPU) to.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, HERE)

import refpin_cases as rc  # noqa: E402
import refshim  # noqa: E402


def run_case(hmc, util, fam, algo, kwargs, seed, num_warmup, num_samples):
    from oracle.threefry import PRNGKey

    model = refshim.FamilyModel(fam)
    kernel = getattr(hmc, algo)(model, **kwargs)
    z0 = rc.init_point(fam, seed)
    zd = {k: refshim._wrap(np.array(v)) for k, v in fam.unravel(z0).items()}
    out = {f: [] for f in rc.FIELDS}
    with util.control_flow_prims_disabled():
        state = kernel.init(refshim._wrap(PRNGKey(seed)), num_warmup, init_params=zd)
        for _ in range(num_warmup + num_samples):
            state = kernel.sample(state, (), {})
            out["z"].append(np.asarray(refshim.ravel_pytree(state.z)[0]))
            for f in rc.FIELDS[1:-1]:
                out[f].append(np.asarray(getattr(state, f)))
            out["step_size"].append(np.asarray(state.adapt_state.step_size))
    res = {f: np.stack(v) for f, v in out.items()}
    a = state.adapt_state
    for names, m in a.inverse_mass_matrix.items():
        res["imm/" + rc.block_tag(names)] = np.asarray(m)
        res["sqrt/" + rc.block_tag(names)] = np.asarray(a.mass_matrix_sqrt[names])
    res["blocks"] = np.array([rc.block_tag(n) for n in a.inverse_mass_matrix])
    res["z0"] = z0
    res["grad_evals"] = np.int64(model.potential.num_evals)
    res["rng_key"] = np.asarray(state.rng_key)
    return res


def diagnostics_inputs():
    """AR(1) chains [4 chains, 300 draws, 3 params] with different autocorrelations, plus a
    short anti-correlated one."""
    rng = np.random.default_rng(77)
    x = np.zeros((4, 300, 3))
    for j, rho in enumerate((0.0, 0.7, 0.95)):
        e = rng.normal(size=(4, 300))
        for t in range(1, 300):
            x[:, t, j] = rho * x[:, t - 1, j] + e[:, t]
    x[2] += 0.3  # one shifted chain: r_hat > 1
    y = np.zeros((2, 64, 2))
    e = rng.normal(size=(2, 64, 2))
    for t in range(1, 64):
        y[:, t] = -0.5 * y[:, t - 1] + e[:, t]
    return {"ar": x, "anti": y}


def diagnostics_fixture():
    """Known answers from the reference's own numpyro/diagnostics.py (executed unmodified)."""
    D = refshim.load_diagnostics()
    out = {}
    for name, x in diagnostics_inputs().items():
        out[f"diag/{name}/x"] = x
        out[f"diag/{name}/gelman_rubin"] = np.asarray(D.gelman_rubin(x))
        out[f"diag/{name}/split_gelman_rubin"] = np.asarray(D.split_gelman_rubin(x))
        out[f"diag/{name}/autocorrelation"] = np.asarray(D.autocorrelation(x, axis=1))
        out[f"diag/{name}/autocorrelation_unbiased"] = np.asarray(D.autocorrelation(x, axis=1, bias=False))
        out[f"diag/{name}/autocovariance"] = np.asarray(D.autocovariance(x, axis=1))
        out[f"diag/{name}/ess"] = np.asarray(D.effective_sample_size(x))
        out[f"diag/{name}/ess_unbiased"] = np.asarray(D.effective_sample_size(x, bias=False))
        out[f"diag/{name}/hpdi"] = np.asarray(D.hpdi(x.reshape(-1, x.shape[-1]), 0.9))
        s = D.summary({"p": x}, prob=0.9, group_by_chain=True)["p"]
        for k, v in s.items():
            out[f"diag/{name}/summary/{k}"] = np.asarray(v)
    import contextlib
    import io

    x = diagnostics_inputs()["ar"]
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):  # the table MCMC.print_summary prints
        D.print_summary({"a": x, "b": x[..., 0] * 2.0}, prob=0.9, group_by_chain=True)
    out["diag/print_summary"] = np.array(buf.getvalue())
    return out


def collect_fixture():
    """The reference's own ``fori_collect`` (numpyro/util.py, executed unmodified) over a
    counter: entry k of the result is 1 + the index of the iteration whose value was kept."""
    _, _, util = refshim.install(True)
    out = {}
    for c in rc.COLLECT_CASES:
        lower, upper, thin, csize = c
        with util.control_flow_prims_disabled():
            col, last = util.fori_collect(lower, upper, lambda v: v + 1, refshim._wrap(np.int32(0)),
                                          progbar=False, return_last_val=True,
                                          collection_size=csize, thinning=thin)
        assert int(last) == upper
        out[rc.collect_tag(c)] = np.asarray(col)
    return out


def main():
    data = rc.make_data()
    fams = rc.families(data)
    out = {"data/" + k: v for k, v in data.items()}
    for part in (True, False):
        hmc, hu, util = refshim.install(part)
        for name, (fk, algo, kwargs, seed, nw, ns, p) in rc.CASES.items():
            if p != part:
                continue
            res = run_case(hmc, util, fams[fk], algo, kwargs, seed, nw, ns)
            for k, v in res.items():
                out[f"{name}/{k}"] = v
            print(f"{name:28s} grad evals {int(res['grad_evals']):6d}  mean tree "
                  f"{res['num_steps'].mean():7.2f}  step {float(res['step_size'][-1]):.4f}  "
                  f"diverging {int(res['diverging'].sum())}")
    out.update(diagnostics_fixture())
    out.update(collect_fixture())
    np.savez_compressed(os.path.join(HERE, "refpin.npz"), **out)
    print("wrote refpin.npz", os.path.getsize(os.path.join(HERE, "refpin.npz")), "bytes")


if __name__ == "__main__":
    main()
