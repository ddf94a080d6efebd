"""CPU restatement of HMC with energy-conserving subsampling.

TEST INFRASTRUCTURE: imported only by ``tests/`` (and the golden-fixture scripts).  Follows

* ``numpyro/infer/hmc_gibbs.py:38-190``  (``HMCGibbs``: state, init key handling, sample),
* ``numpyro/infer/hmc_gibbs.py:502-665`` (``HMCECS``: init, block pseudo-marginal update of the
  subsample, Metropolis test ``exp(pe - pe_new)``, gradient refresh on acceptance),
* ``numpyro/contrib/ecs_proxies.py:23-52``  (``perturbed_method``: difference estimator with the
  variance correction ``- n^2 / (2 m) var(diff)``),
* ``numpyro/contrib/ecs_proxies.py:55-91``  (``_update_block`` / ``block_update``),
* ``numpyro/contrib/ecs_proxies.py:94-278`` (``taylor_proxy``: first/second-order Taylor control
  variate around reference parameters; the gradients/Hessians the reference gets from autodiff
  are written in closed form for the Bernoulli-logits likelihood),
* ``numpyro/primitives.py:443-462``       (``_subsample_fn``, CPU branch: Fisher-Yates with one
  ``randint`` per drawn index),
* ``numpyro/handlers.py`` ``seed.process_message`` (one ``split`` per unobserved sample site and
  per subsampling plate, in trace order),

for the registered logistic-regression model with a subsampled plate
(``examples/covtype.py:66-71`` with ``subsample_size``).

``randint`` restates ``jax._src.random._randint`` (third party, not under ``/root/reference``;
jax is not installable here): parity of its bit stream is UNPINNED against a live JAX, like
Threefry itself; the arithmetic is uint32 with wrap-around as in XLA.
"""
from collections import namedtuple

import numpy as np

from . import hmc as ohmc
from .potentials import LOG_SQRT_2PI, Family

_U32 = np.uint32

HMCECSState = namedtuple("HMCECSState", "z hmc_state rng_key gibbs_state accept_prob")


# --------------------------------------------------------------------------- jax.random.randint
def randint(cfg, key, shape, minval, maxval):
    """``jax.random.randint(key, shape, minval, maxval)`` for int32 with in-range bounds:
    ``k1, k2 = split(key)``; ``hi = bits(k1)``, ``lo = bits(k2)``; ``span = maxval - minval``;
    ``mult = (2^16 % span)^2 % span``; ``off = ((hi % span) * mult + lo % span) % span`` in uint32
    (products wrap); result ``minval + off``."""
    ks = cfg.split(key, 2)
    hi = np.asarray(cfg.random_bits(ks[0], shape), _U32)
    lo = np.asarray(cfg.random_bits(ks[1], shape), _U32)
    span = _U32(maxval - minval) if maxval > minval else _U32(1)
    mult = _U32(65536) % span
    with np.errstate(over="ignore"):
        mult = _U32(mult * mult) % span
        off = ((hi % span).astype(_U32) * mult + (lo % span).astype(_U32)).astype(_U32) % span
    return (np.int32(minval) + off.astype(np.int32)).astype(np.int32)


# --------------------------------------------------------------------------- primitives.py:443-462
def subsample_fn(cfg, size, subsample_size, rng_key):
    """Fisher-Yates partial shuffle of ``arange(size)``; returns the last ``subsample_size``
    entries (``val[-subsample_size:]``)."""
    keys = cfg.split(rng_key, subsample_size)
    val = np.arange(size, dtype=np.int32)
    for idx in range(subsample_size):
        i_p1 = size - idx
        i = i_p1 - 1
        j = int(randint(cfg, keys[idx], (), 0, i_p1))
        val[i], val[j] = val[j], val[i]
    return val[size - subsample_size:].copy()


def prototype_subsample(cfg, key_u, size, subsample_size, sites_before_plate=1):
    """The subsample indices of ``trace(substitute(seed(model, key_u), init_to_sample))``
    (``hmc_gibbs.py:547-552``): the seed handler splits its key once per unobserved sample site
    and once for the plate, in trace order (``coefs`` precedes plate ``N`` in
    ``examples/covtype.py:66-71``)."""
    k = np.asarray(key_u, _U32)
    sub = None
    for _ in range(sites_before_plate + 1):
        ks = cfg.split(k, 2)
        k, sub = ks[0], ks[1]
    return subsample_fn(cfg, size, subsample_size, sub)


# --------------------------------------------------------------------------- ecs_proxies.py:58-73
def update_block(cfg, rng_key, num_blocks, subsample_idx, plate_size):
    size, subsample_size = plate_size
    ks = cfg.split(rng_key, 3)
    rng_key, subkey, block_key = ks[0], ks[1], ks[2]
    block_size = (subsample_size - 1) // num_blocks + 1
    pad = block_size - (subsample_size - 1) % block_size - 1
    chosen_block = int(randint(cfg, block_key, (), 0, num_blocks))
    new_idx = randint(cfg, subkey, (block_size,), 0, size)
    padded = np.concatenate([np.asarray(subsample_idx, np.int32), np.zeros(pad, np.int32)])
    # lax.dynamic_update_slice_in_dim clamps the start so that the update fits
    start = min(max(chosen_block * block_size, 0), padded.shape[0] - block_size)
    padded[start:start + block_size] = new_idx
    return rng_key, padded[:subsample_size].copy(), pad, new_idx, start


# --------------------------------------------------------------------------- the potential
def _bce(l, y):
    e = np.exp(-np.abs(l))
    return np.maximum(l, 0) + np.log1p(e) - l * y, e


def _sigmoid(l, e):
    return np.where(l >= 0, 1 / (1 + e), e / (1 + e))


def perturbed_loglik(loglik_rows, proxy_rows, proxy_sum, n, m, f=np.float32):
    """``perturbed_method.estimator`` (``contrib/ecs_proxies.py:23-52``) for one subsample site:
    ``proxy_sum + n mean(diff) - 0.5 n^2/m var(diff)``.  Returns (estimate, diff, mean)."""
    n, m = f(n), f(m)
    diff = (loglik_rows - proxy_rows).astype(f)
    mean = diff.mean(dtype=f)
    var = ((diff - mean) ** 2).mean(dtype=f)  # jnp.var: population variance
    return f(proxy_sum + n * mean - f(0.5) * (n * n / m) * var), diff, mean


class LogisticECS(Family):
    """``coefs ~ Normal(0, 1)[D]``; plate ``N`` of size n subsampled to m rows;
    ``obs ~ Bernoulli(logits = X[idx] @ coefs)``.

    ``degree`` 0: no proxy -- the plate scales the subsample log-likelihood by n/m
    (``hmc_gibbs.py:584-588``, plate ``scale``).  ``degree`` 1/2: ``estimate_likelihood`` +
    ``perturbed_method`` with ``taylor_proxy(ref, degree)``: the likelihood is masked and the factor
    ``proxy_sum(theta) + n mean(diff) - n^2/(2m) var(diff)`` is added, ``diff_i = loglik_i(theta) -
    proxy_i(theta)`` over the current subsample.
    """

    name = "logistic_ecs"

    def __init__(self, X, y, subsample_size, degree=0, ref=None, dtype=np.float32):
        super().__init__(dtype)
        self.X = np.ascontiguousarray(X, dtype=self.dtype)
        self.y = np.ascontiguousarray(y).astype(self.dtype)
        self.n, self.m = self.X.shape[0], int(subsample_size)
        self.degree = int(degree)
        self.trace_sites = (("coefs", self.X.shape[1]),)
        self.idx = np.arange(self.m, dtype=np.int32)
        if self.degree:
            self.ref = np.asarray(ref, self.dtype)
            # ecs_proxies.py:165-169: sums over ALL rows at the reference point
            l = self.X @ self.ref
            bce, e = _bce(l, self.y)
            sig = _sigmoid(l, e).astype(self.dtype)
            self.ref_sum = self.dtype.type((-bce).sum(dtype=self.dtype))
            self.ref_grad = (self.X.T @ (self.y - sig)).astype(self.dtype)
            w = (sig * (1 - sig)).astype(self.dtype)
            self.ref_hess = (-(self.X.T * w) @ self.X).astype(self.dtype) if self.degree == 2 else None

    def set_subsample(self, idx):
        self.idx = np.asarray(idx, np.int32).copy()

    def potential_and_grad(self, z):
        f = self.dtype.type
        z = np.asarray(z, self.dtype)
        Xs, ys = self.X[self.idx], self.y[self.idx]
        n, m = f(self.n), f(self.m)
        prior = (f(0.5) * z * z + f(LOG_SQRT_2PI)).sum(dtype=self.dtype)
        l = Xs @ z
        bce, e = _bce(l, ys)
        sig = _sigmoid(l, e).astype(self.dtype)
        if self.degree == 0:
            U = prior + (n / m) * bce.sum(dtype=self.dtype)
            grad = z + (n / m) * (Xs.T @ (sig - ys))
            return f(U), grad.astype(self.dtype)
        # taylor proxy (ecs_proxies.py:232-274)
        dz = z - self.ref
        ls = Xs @ self.ref
        bce_s, e_s = _bce(ls, ys)
        sig_s = _sigmoid(ls, e_s).astype(self.dtype)
        delta = Xs @ dz
        proxy = -bce_s + (ys - sig_s) * delta
        dproxy = ys - sig_s
        proxy_sum = self.ref_sum + self.ref_grad @ dz
        gsum = self.ref_grad.copy()
        if self.degree == 2:
            w = sig_s * (1 - sig_s)
            proxy = proxy - f(0.5) * w * delta * delta
            dproxy = dproxy - w * delta
            Hd = self.ref_hess @ dz
            proxy_sum = proxy_sum + f(0.5) * (Hd @ dz)
            gsum = gsum + Hd
        loglik, diff, mean = perturbed_loglik(-bce, proxy, proxy_sum, self.n, self.m, f)
        U = prior - loglik
        cdiff = (ys - sig) - dproxy
        coef = cdiff * (n / m - (n * n / (m * m)) * (diff - mean))
        grad = z - gsum - Xs.T @ coef
        return f(U), grad.astype(self.dtype)

    def constrain(self, z):
        return {"coefs": np.asarray(z)}


# --------------------------------------------------------------------------- hmc_gibbs.py:502-665
class HMCECSKernel:
    """``HMCECS(NUTS(model), num_blocks=..., proxy=...)`` on a ``LogisticECS`` family."""

    def __init__(self, family, num_blocks=1, cfg=None, **kernel_kwargs):
        self.family = family
        self.num_blocks = int(num_blocks)
        self.inner = ohmc.HMCKernel(family, cfg=cfg, **kernel_kwargs)
        self.cfg = self.inner.cfg
        self.trace = None  # optional list of (proposed idx, pe_new, accept_prob, accepted)

    def init(self, rng_key, num_warmup, init_params=None):
        cfg, fam = self.cfg, self.family
        ks = cfg.split(rng_key, 2)  # hmc_gibbs.py:547
        rng_key, key_u = ks[0], ks[1]
        idx = prototype_subsample(cfg, key_u, fam.n, fam.m)
        if fam.degree:  # hmc_gibbs.py:594-595: one more split for gibbs_init (its key is unused)
            rng_key = cfg.split(rng_key, 2)[0]
        fam.set_subsample(idx)
        ks = cfg.split(rng_key, 2)  # HMCGibbs.init, hmc_gibbs.py:135
        rng_key, key_z = ks[0], ks[1]
        hmc_state = self.inner.init(key_z, num_warmup, init_params)
        return HMCECSState({"N": idx, "coefs": hmc_state.z}, hmc_state, rng_key, (), np.float32(0))

    def sample(self, state):
        cfg, fam, f = self.cfg, self.family, self.cfg.f
        rng_key = cfg.split(state.rng_key, 2)[0]  # hmc_gibbs.py:619 (rng_gibbs is unused)
        idx = state.z["N"]
        _, idx_new, _, _, _ = update_block(cfg, rng_key, self.num_blocks, idx, (fam.n, fam.m))
        pe = state.hmc_state.potential_energy
        fam.set_subsample(idx_new)
        with np.errstate(all="ignore"):
            pe_new, grad_new = fam.potential_and_grad(state.hmc_state.z)
            accept_prob = f(min(np.exp(f(pe - pe_new)), f(1.0))) if not np.isnan(pe - pe_new) else f(np.nan)
        transition = bool(cfg.bernoulli(rng_key, accept_prob))  # hmc_gibbs.py:640
        if self.trace is not None:
            self.trace.append((idx_new.copy(), f(pe_new), accept_prob, transition))
        if transition:
            idx, hmc_state = idx_new, state.hmc_state._replace(z_grad=grad_new, potential_energy=f(pe_new))
        else:
            fam.set_subsample(idx)
            hmc_state = state.hmc_state
        hmc_state = self.inner.sample(hmc_state)
        return HMCECSState({"N": idx, "coefs": hmc_state.z}, hmc_state, rng_key, (), accept_prob)


def run_chain(kernel, rng_key, num_warmup, num_samples):
    state = kernel.init(rng_key, num_warmup)
    zs, acc, steps = [], [], []
    for i in range(num_warmup + num_samples):
        state = kernel.sample(state)
        if i >= num_warmup:
            zs.append(state.hmc_state.z.copy())
            acc.append(state.accept_prob)
            steps.append(state.hmc_state.num_steps)
    return {"z": np.asarray(zs), "accept_prob": np.asarray(acc), "num_steps": np.asarray(steps)}, state
