"""Oracle and CUDA path against traces of the REFERENCE's own sampler source.

``tests/golden/refpin.npz`` was produced by ``tests/golden/make_refpin.py``: the unmodified
``numpyro/infer/hmc.py`` + ``hmc_util.py`` (+ ``numpyro/util.py``) from /root/reference,
executed over a NumPy stand-in for JAX (``tests/golden/refshim.py`` says exactly what is and
is not reference code in such a run), driven through the reference's ``HMC``/``NUTS``
classes.  28 cases: fixed-parameter NUTS (both Threefry layouts), diagonal / dense /
block-structured mass-matrix warm-ups with window ends, dual averaging, ``max_tree_depth``
tuples, ``find_heuristic_step_size``, user mass matrices, HMC with ``trajectory_length`` /
``num_steps``, on eight schools, logistic, horseshoe (both likelihoods) and stochastic
volatility.

CPU: the oracle reproduces every recorded field of every transition BIT FOR BIT (same NumPy
arithmetic, so no tolerance) -- this is what pins the oracle to the reference.
GPU: the CUDA path through the C ABI -- identical tree sizes / divergences for the
fixed-parameter cases (tier 3), and statistically/numerically close adaptation.
"""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import refpin_cases as rc  # noqa: E402

G = np.load(os.path.join(HERE, "golden", "refpin.npz"))
DATA = {k[5:]: G[k] for k in G.files if k.startswith("data/")}
pytestmark = pytest.mark.filterwarnings("ignore")


def _oracle_trace(name):
    from oracle import hmc as ohmc
    from oracle.hmc_util import Config
    from oracle.threefry import PRNGKey

    fk, algo, kw, seed, nw, ns, part = rc.CASES[name]
    fam = rc.families(DATA)[fk]
    k = ohmc.HMCKernel(fam, algo=algo, cfg=Config(np.float32, part), **kw)
    s = k.init(PRNGKey(seed), nw, init_params=G[name + "/z0"])
    out = {f: [] for f in rc.FIELDS}
    for _ in range(nw + ns):
        s = k.sample(s)
        out["z"].append(s.z.copy())
        for f in rc.FIELDS[1:-1]:
            out[f].append(getattr(s, f))
        out["step_size"].append(s.adapt_state.step_size)
    return {f: np.asarray(v) for f, v in out.items()}, s


def test_fixture_data_is_the_case_table_data():
    new = rc.make_data()
    for k, v in new.items():
        assert np.array_equal(v, DATA[k]), k


@pytest.mark.parametrize("name", list(rc.CASES))
def test_oracle_matches_reference_source_bit_for_bit(name):
    res, last = _oracle_trace(name)
    for f in rc.FIELDS:
        want = G[f"{name}/{f}"]
        got = np.asarray(res[f], want.dtype)
        assert np.array_equal(got, want, equal_nan=True), (name, f)
    imm = last.adapt_state.inverse_mass_matrix
    blocks = [str(b) for b in G[name + "/blocks"]]
    if isinstance(imm, dict):
        assert [rc.block_tag(k) for k in imm] == blocks  # block ORDER = key-split order
        for k, m in imm.items():
            assert np.array_equal(m, G[f"{name}/imm/{rc.block_tag(k)}"]), (name, k)
            assert np.array_equal(last.adapt_state.mass_matrix_sqrt[k],
                                  G[f"{name}/sqrt/{rc.block_tag(k)}"]), (name, k)
    else:
        assert len(blocks) == 1
        assert np.array_equal(imm, G[f"{name}/imm/{blocks[0]}"])
        assert np.array_equal(last.adapt_state.mass_matrix_sqrt, G[f"{name}/sqrt/{blocks[0]}"])
    assert np.array_equal(np.asarray(last.rng_key, np.uint32), G[name + "/rng_key"])


@pytest.mark.skipif(not os.path.isdir("/root/reference/numpyro"),
                    reason="the reference checkout only exists in the build container")
@pytest.mark.parametrize("name", ["es_adapt_dense", "es_block_two", "es_hmc_traj", "logit_fixed"])
def test_fixture_regenerates_from_the_reference_source(name):
    """Re-executes the reference's hmc.py/hmc_util.py (never on the GPU box)."""
    import make_refpin
    import refshim

    fk, algo, kw, seed, nw, ns, part = rc.CASES[name]
    hmc, _, util = refshim.install(part)
    res = make_refpin.run_case(hmc, util, rc.families(DATA)[fk], algo, kw, seed, nw, ns)
    for f in rc.FIELDS:
        assert np.array_equal(res[f], G[f"{name}/{f}"], equal_nan=True), (name, f)
