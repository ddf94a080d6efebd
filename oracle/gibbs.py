"""Oracle restatement of ``HMCGibbs`` (``numpyro/infer/hmc_gibbs.py:30-190``) over the oracle's
HMC/NUTS kernel (``oracle/hmc.py``).

TEST INFRASTRUCTURE (see ``oracle/__init__.py``).

Pinned to the reference: ``tests/golden/make_refgibbs.py`` runs the reference's own ``HMCGibbs``
(``init`` + ``sample``, around the reference's ``NUTS(model)``) from /root/reference over the NumPy
stand-in and stores every state; ``tests/test_refgibbs.py`` holds this module to them bit for bit.
"""
from collections import namedtuple

import numpy as np

from . import hmc as ohmc
from .potentials import Family

HMCGibbsState = namedtuple("HMCGibbsState", "z, hmc_state, rng_key")  # hmc_gibbs.py:22


class Conditioned(Family):
    """A family with the latent sites ``names`` conditioned (``_wrap_model``,
    ``hmc_gibbs.py:30-34``: ``condition`` makes them observed -- no init draw, no key split, no
    Jacobian term -- and ``substitute`` feeds their CONSTRAINED values): the flat vector covers
    the remaining sites; ``positive`` lists the full family's sites under ``ExpTransform``."""

    def __init__(self, full, names, positive=()):
        super().__init__(full.dtype)
        self.full, self.names, self.positive = full, tuple(names), tuple(positive)
        self.name = full.name
        offs = full.offsets()
        self.trace_sites = tuple((k, s) for k, s in full.trace_sites if k not in self.names)
        self.cond_sites = tuple((k, s, off) for k, s, off in full.flat_sites if k in self.names)
        self.cond_index = np.concatenate([np.arange(off, off + s) for _, s, off in self.cond_sites])
        self.free_index = np.concatenate([np.arange(offs[k][0], offs[k][0] + s)
                                          for k, s, _ in self.flat_sites])
        self._zfull = np.zeros(full.dim, self.dtype)
        self._offset = self.dtype.type(0)

    def set_values(self, constrained):
        """Constrained values of the conditioned sites."""
        off = self.dtype.type(0)
        for k, s, o in self.cond_sites:
            v = np.asarray(constrained[k], self.dtype).reshape(s)
            if k in self.positive:  # the full potential carries -log|J| = -z for this site; the
                v = np.log(v)       # reference's conditioned density does not
                off = off + v.sum(dtype=self.dtype)
            self._zfull[o:o + s] = v
        self._offset = self.dtype.type(off)

    def potential_and_grad(self, z):
        zf = self._zfull.copy()
        zf[self.free_index] = np.asarray(z, self.dtype)
        U, g = self.full.potential_and_grad(zf)
        return self.dtype.type(U + self._offset), g[self.free_index]

    def constrain(self, z):
        """The inner kernel's ``postprocess_fn`` under the substituted values: constrained HMC
        sites + deterministic sites (``hmc_gibbs.py:109-119``)."""
        z = np.asarray(z)
        zf = np.broadcast_to(self._zfull, z.shape[:-1] + (self.full.dim,)).copy()
        zf[..., self.free_index] = z
        out = self.full.constrain(zf)
        return {k: v for k, v in out.items() if k not in self.names}


class HMCGibbs:
    """``hmc_gibbs.py:36-190``.  ``prior_sample(name, key, cfg)`` draws a Gibbs site from its
    prior as ``init_to_sample`` does (``initialization.py:88-93``)."""

    def __init__(self, family, gibbs_fn, gibbs_sites, prior_sample, positive=(), cfg=None, **kernel_kwargs):
        self.full = family
        self.gibbs_sites = list(gibbs_sites)
        self.gibbs_fn = gibbs_fn
        self.prior_sample = prior_sample
        self.cond = Conditioned(family, self.gibbs_sites, positive)
        self.inner = ohmc.HMCKernel(self.cond, cfg=cfg, **kernel_kwargs)
        self.cfg = self.inner.cfg
        self.prototype = None

    def _sites(self, flat):
        return {k: flat[off:off + s] for k, s, off in self.cond.flat_sites}

    def init(self, rng_key, num_warmup, init_params=None):
        cfg = self.cfg
        init_params = None if init_params is None else dict(init_params)
        if self.prototype is None:  # :123-130
            rng_key, key_u = cfg.split(rng_key)
            k, proto = key_u, {}
            for name, _ in self.full.trace_sites:  # seed handler: one split per latent site
                k, ks = cfg.split(k)
                if name in self.gibbs_sites:
                    proto[name] = self.prior_sample(name, ks, cfg)
            self.prototype = proto
        rng_key, key_z = cfg.split(rng_key)  # :132
        gibbs = {}
        for name, _ in self.full.trace_sites:  # :136-142
            if init_params and name in init_params and name in self.gibbs_sites:
                gibbs[name] = np.asarray(init_params.pop(name), cfg.dtype)
            elif name in self.gibbs_sites:
                gibbs[name] = self.prototype[name]
        self.cond.set_values(gibbs)
        ip = None
        if init_params:
            ip = np.concatenate([np.ravel(np.asarray(init_params[k], cfg.dtype))
                                 for k, _, _ in self.cond.flat_sites])
        hmc_state = self.inner.init(key_z, num_warmup, ip)
        return HMCGibbsState({**gibbs, **self._sites(hmc_state.z)}, hmc_state, rng_key)

    def sample(self, state):
        cfg = self.cfg
        rng_key, rng_gibbs = cfg.split(state.rng_key)  # :156
        z_gibbs = {k: v for k, v in state.z.items() if k in self.gibbs_sites}
        z_hmc = self.cond.constrain(state.hmc_state.z)  # :166-168 (under the CURRENT values)
        z_gibbs = self.gibbs_fn(rng_key=rng_gibbs, gibbs_sites=z_gibbs, hmc_sites=z_hmc)  # :170
        z_gibbs = {k: np.asarray(v, cfg.dtype) for k, v in z_gibbs.items()}
        self.cond.set_values(z_gibbs)
        pe, z_grad = self.inner.potential_fn(state.hmc_state.z)  # :178-181
        hmc_state = state.hmc_state._replace(z_grad=z_grad, potential_energy=cfg.f(pe))
        hmc_state = self.inner.sample(hmc_state)  # :185
        return HMCGibbsState({**z_gibbs, **self._sites(hmc_state.z)}, hmc_state, rng_key)
