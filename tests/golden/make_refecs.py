"""Write ``refecs.npz``: the pieces of the REFERENCE's HMCECS that run over the NumPy stand-in
(``refshim_full``), executed from /root/reference:

* ``numpyro.contrib.ecs_proxies._update_block`` (``:58-73``) -- the block update of the subsample
  (key splitting, block choice, padding, clamped slice update), both Threefry layouts;
* ``numpyro.primitives._subsample_fn`` (``:443-462``, CPU branch) -- the Fisher-Yates initial
  subsample;
* ``numpyro.contrib.ecs_proxies.perturbed_method`` (``:23-52``) -- the difference estimator with
  its variance correction, fed with the reference's own ``Bernoulli(logits=...)`` log-probs and
  a proxy stand-in.

    python tests/golden/make_refecs.py      (needs /root/reference; not for the GPU box)

``jax.random.randint`` is third party (jax is not installable here): the stand-in uses the
restatement in ``oracle/ecs.py`` -- so these fixtures pin the reference's control flow around
``randint``, not ``randint``'s bit stream (UNPINNED, like Threefry: SURVEY 8c).
"""
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, HERE)

import refshim  # noqa: E402
import refshim_full  # noqa: E402

W = refshim._wrap


def main():
    numpyro = refshim_full.install_full()
    import jax
    from jax import lax, random
    import jax.numpy as jnp

    from oracle import ecs as oecs
    from oracle import hmc_util as ohu
    from oracle.threefry import PRNGKey

    state = {"cfg": ohu.Config(partitionable=True)}

    def randint(key, shape=(), minval=0, maxval=1, dtype=None):
        return W(oecs.randint(state["cfg"], np.asarray(key), tuple(shape), int(minval), int(maxval)))

    random.randint = randint
    jnp.pad = lambda a, pw, **k: W(np.pad(np.asarray(a), pw))

    def dus(operand, update, start, axis):
        operand, update = np.asarray(operand).copy(), np.asarray(update)
        n = update.shape[axis]
        s = int(min(max(int(start), 0), operand.shape[axis] - n))  # XLA clamps the start index
        sl = [slice(None)] * operand.ndim
        sl[axis] = slice(s, s + n)
        operand[tuple(sl)] = update
        return W(operand)

    lax.dynamic_update_slice_in_dim = dus

    def scan(f, init, xs):
        carry = init
        for x in np.asarray(xs):
            carry, _ = f(carry, x)
        return carry, None

    lax.scan = scan
    jax.default_backend = lambda: "cpu"

    import numpyro.contrib.ecs_proxies as ep
    import numpyro.primitives as prim

    for mod in (ep, prim):
        mod.random = random
        mod.lax = lax
        mod.jnp = jnp
    out = {}
    cases = [(1000, 100, 581012, 1), (1000, 7, 581012, 2), (37, 5, 400, 3), (64, 64, 5000, 4), (10, 1, 50, 5)]
    for part in (True, False):
        state["cfg"] = ohu.Config(partitionable=part)
        refshim.PARTITIONABLE = part
        for m, nb, size, seed in cases:
            idx0 = np.random.default_rng(seed).integers(0, size, m).astype(np.int32)
            key = PRNGKey(seed)
            rk, new, pad, new_idx, start = ep._update_block(W(key), nb, W(idx0), (size, m))
            tag = f"update_block/{int(part)}/{m}_{nb}_{size}"
            out[tag + "/idx0"] = idx0
            out[tag + "/key"] = np.asarray(key)
            out[tag + "/rng_key"] = np.asarray(rk)
            out[tag + "/idx"] = np.asarray(new)
            out[tag + "/new_idx"] = np.asarray(new_idx)
            out[tag + "/start"] = np.int32(start)
            out[tag + "/pad"] = np.int32(pad)
        for size, m, seed in [(500, 20, 1), (4000, 200, 2)]:
            class _At:  # val.at[idx].set(v)
                def __init__(self, a): self.a = a
                def __getitem__(self, i):
                    a = self.a
                    return types.SimpleNamespace(set=lambda v: _arr(_put(a, i, v)))
            def _put(a, i, v):
                b = np.asarray(a).copy(); b[np.asarray(i)] = np.asarray(v); return b
            class _arr(np.ndarray):
                def __new__(cls, a):
                    return np.asarray(a).view(cls)
                @property
                def at(self): return _At(self)
            jnp.arange = lambda n, *a, **k: _arr(np.arange(n, dtype=np.int32))
            jnp.array = lambda a, *x, **k: np.asarray([int(v) for v in a])
            key = PRNGKey(100 + seed)
            val = prim._subsample_fn(size, m, W(key))
            tag = f"subsample_fn/{int(part)}/{size}_{m}"
            out[tag + "/key"] = np.asarray(key)
            out[tag + "/idx"] = np.asarray(val).astype(np.int32)
    # perturbed_method: reference Bernoulli log-probs + estimator
    import numpyro.distributions as dist
    rng = np.random.default_rng(7)
    n, m, D = 5000, 60, 4
    X = rng.standard_normal((m, D)).astype(np.float32)
    y = (rng.random(m) < 0.5).astype(np.float32)
    th = (rng.standard_normal(D) * 0.5).astype(np.float32)
    proxy_sub = (rng.standard_normal(m) * 0.3 - 0.7).astype(np.float32)
    proxy_all = np.float32(-3100.25)
    est = ep.perturbed_method({"N": (n, m)},
                              lambda params, names, gs: ({"N": W(proxy_all)}, {"N": W(proxy_sub)}))
    fn = dist.Bernoulli(logits=W(X @ th))
    val = est({"obs": (fn, W(y), "N", -1)}, {"coefs": W(th)}, ())
    out["perturbed/X"], out["perturbed/y"], out["perturbed/theta"] = X, y, th
    out["perturbed/proxy_sub"], out["perturbed/proxy_all"] = proxy_sub, proxy_all
    out["perturbed/n_m"] = np.array([n, m], np.int32)
    out["perturbed/value"] = np.float32(np.asarray(val))
    np.savez_compressed(os.path.join(HERE, "refecs.npz"), **out)
    print("wrote refecs.npz:", len(out), "arrays; perturbed value", out["perturbed/value"])


if __name__ == "__main__":
    main()
