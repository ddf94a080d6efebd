"""Oracle restatement of ``numpyro/infer/hmc.py`` (+ the init search of
``numpyro/infer/util.py`` and the collection loop of ``numpyro/util.py``).

TEST INFRASTRUCTURE (see ``oracle/__init__.py``).
"""
from collections import namedtuple
import copy
import math

import numpy as np

from . import hmc_util as hu
from .hmc_util import Config, IntegratorState

HMCState = namedtuple(  # hmc.py:31-48
    "HMCState",
    ["i", "z", "z_grad", "potential_energy", "energy", "r", "trajectory_length", "num_steps",
     "accept_prob", "mean_accept_prob", "diverging", "adapt_state", "rng_key"],
)


def momentum_generator(cfg, mass_matrix_sqrt, rng_key):
    """hmc.py:92-110.  A model-built kernel always carries a dict mass matrix with a
    single block (``hmc.py:760-769``), so the key is first ``split(key, 1)``."""
    if isinstance(mass_matrix_sqrt, dict):  # structured mass matrix: one key per block
        rng_keys = cfg.split(rng_key, len(mass_matrix_sqrt))
        r = np.zeros(sum(cfg.block_index(k).size for k in mass_matrix_sqrt), cfg.dtype)
        for (names, mm_sqrt), key in zip(mass_matrix_sqrt.items(), rng_keys):
            eps = cfg.normal(key, (mm_sqrt.shape[0],))
            r[cfg.block_index(names)] = (mm_sqrt * eps if mm_sqrt.ndim == 1
                                         else mm_sqrt @ eps).astype(cfg.dtype)
        return r
    rng_key = cfg.split(rng_key, 1)[0]
    eps = cfg.normal(rng_key, (mass_matrix_sqrt.shape[0],))
    if mass_matrix_sqrt.ndim == 1:
        return (mass_matrix_sqrt * eps).astype(cfg.dtype)
    return (mass_matrix_sqrt @ eps).astype(cfg.dtype)


def find_valid_initial_params(cfg, rng_key, family, init_values=None, radius=2.0):
    """``infer/util.py:386-477`` fast branch ``:423-432``: up to 100 tries; per site
    in TRACE order ``uniform(subkey, shape, -radius, radius)`` then
    ``key, subkey = split(key)``; valid iff U and grad are finite."""
    init_values = {} if init_values is None else init_values
    key = rng_key
    offs = family.offsets()
    i, is_valid = 0, False
    z = pe = grad = None
    while i < 100 and not is_valid:
        key, subkey = cfg.split(key)
        z = np.zeros(family.dim, cfg.dtype)
        for name, size in family.trace_sites:
            off, _ = offs[name]
            if name in init_values:
                z[off:off + size] = np.asarray(init_values[name], cfg.dtype).ravel()
            else:
                z[off:off + size] = cfg.uniform(subkey, (size,), -radius, radius)
                key, subkey = cfg.split(key)
        with np.errstate(all="ignore"):
            pe, grad = family.potential_and_grad(z)
        is_valid = bool(np.isfinite(pe) and np.all(np.isfinite(grad)))
        i += 1
    if not is_valid:
        raise RuntimeError("Cannot find valid initial parameters. Please check your model again.")
    return (z, pe, grad), is_valid


class HMCKernel:
    """Functional ``hmc(potential_fn, algo)`` closure pair (hmc.py:113-538) bound to one
    registered family, plus the ``HMC.init`` key handling (hmc.py:740-799)."""

    def __init__(self, family, algo="NUTS", step_size=1.0, inverse_mass_matrix=None,
                 adapt_step_size=True, adapt_mass_matrix=True, dense_mass=False,
                 target_accept_prob=0.8, num_steps=None, trajectory_length=2 * math.pi,
                 max_tree_depth=10, regularize_mass_matrix=True, find_heuristic_step_size=False,
                 cfg=None):
        self.cfg = copy.copy(cfg) if cfg is not None else Config()
        self.cfg.site_index = {k: np.arange(off, off + s) for k, (off, s) in
                               family.offsets().items()}
        self.family = family
        if isinstance(dense_mass, (list, tuple)) or isinstance(inverse_mass_matrix, dict):
            # hmc.py:760-769: structural form; a bool becomes one block over the sorted sites
            if isinstance(dense_mass, bool):
                dense_mass = [tuple(sorted(self.cfg.site_index))] if dense_mass else []
            dense_mass = [tuple(b) for b in dense_mass]
        self.algo = algo
        self.step_size = step_size
        self.inverse_mass_matrix = inverse_mass_matrix
        self.adapt_step_size = adapt_step_size
        self.adapt_mass_matrix = adapt_mass_matrix
        self.dense_mass = dense_mass
        self.target_accept_prob = target_accept_prob
        self.fixed_num_steps = num_steps
        self.trajectory_length = None if algo == "NUTS" else trajectory_length
        self.max_treedepth = (max_tree_depth if isinstance(max_tree_depth, tuple)
                              else (max_tree_depth, max_tree_depth))
        self.regularize_mass_matrix = regularize_mass_matrix
        self.find_heuristic_step_size = find_heuristic_step_size
        self.max_delta_energy = 1000.0
        self.num_grad_evals = 0
        self.trace = None  # optional list collecting every decision (tier-3 parity)

        def potential_fn(z):
            self.num_grad_evals += 1
            with np.errstate(all="ignore"):
                return family.potential_and_grad(z)

        self.potential_fn = potential_fn
        self.vv_init, self.vv_update = hu.velocity_verlet(self.cfg, potential_fn)

    # -- HMC.init (hmc.py:740-799) + init_kernel (hmc.py:193-362)
    def init(self, rng_key, num_warmup, init_params=None, init_values=None):
        cfg, f = self.cfg, self.cfg.f
        rng_key, rng_key_init_model = cfg.split(rng_key)
        if init_params is None:
            (z, pe, z_grad), _ = find_valid_initial_params(cfg, rng_key_init_model, self.family,
                                                           init_values)
        else:
            z = np.asarray(init_params, cfg.dtype)
            pe, z_grad = self.potential_fn(z)
        self.wa_steps = num_warmup
        frss = None
        if self.find_heuristic_step_size:  # hmc.py:319-323
            def frss(step_size, imm, z_info, key):
                return hu.find_reasonable_step_size(
                    cfg, self.vv_update, lambda m, k: momentum_generator(cfg, m, k), step_size, imm,
                    z_info, key)
        self.wa_init, self.wa_update = hu.warmup_adapter(
            cfg, num_warmup, find_reasonable_step_size=frss, adapt_step_size=self.adapt_step_size,
            adapt_mass_matrix=self.adapt_mass_matrix, dense_mass=self.dense_mass,
            target_accept_prob=self.target_accept_prob,
            regularize_mass_matrix=self.regularize_mass_matrix)
        keys = cfg.split(rng_key, 3)
        rng_key_hmc, rng_key_wa, rng_key_momentum = keys[0], keys[1], keys[2]
        z_info = IntegratorState(z, None, pe, z_grad)
        wa_state = self.wa_init(z_info, rng_key_wa, self.step_size, self.inverse_mass_matrix)
        r = momentum_generator(cfg, wa_state.mass_matrix_sqrt, rng_key_momentum)
        energy = f(pe + hu.kinetic_energy(cfg, wa_state.inverse_mass_matrix, r))
        tl = None if self.trajectory_length is None else f(self.trajectory_length)
        return HMCState(0, z, z_grad, f(pe), energy, None, tl, 0, f(0), f(0), False, wa_state,
                        rng_key_hmc)

    # -- _hmc_next (hmc.py:364-414)
    def _hmc_next(self, step_size, imm, vv_state, rng_key, trajectory_length):
        cfg, f = self.cfg, self.cfg.f
        if self.fixed_num_steps is not None:
            num_steps = self.fixed_num_steps
        else:
            num_steps = int(np.ceil(f(trajectory_length) / f(step_size)))  # hmc.py:85-89
        if trajectory_length is not None:
            step_size = f(trajectory_length) / f(num_steps)
        vv_state_new = vv_state
        for _ in range(num_steps):
            vv_state_new = self.vv_update(step_size, imm, vv_state_new)
        energy_old = f(vv_state.potential_energy + hu.kinetic_energy(cfg, imm, vv_state.r))
        with np.errstate(all="ignore"):
            energy_new = f(vv_state_new.potential_energy
                           + hu.kinetic_energy(cfg, imm, vv_state_new.r))
            delta_energy = f(energy_new - energy_old)
            if np.isnan(delta_energy):
                delta_energy = f(np.inf)
            accept_prob = f(min(np.exp(-delta_energy), f(1.0)))
        diverging = bool(delta_energy > f(self.max_delta_energy))
        transition = cfg.bernoulli(rng_key, accept_prob)
        if self.trace is not None:
            self.trace.append(("mh", bool(transition)))
        if transition:
            return vv_state_new, energy_new, num_steps, accept_prob, diverging
        return vv_state, energy_old, num_steps, accept_prob, diverging

    # -- _nuts_next (hmc.py:416-455)
    def _nuts_next(self, step_size, imm, vv_state, rng_key, max_treedepth_current):
        cfg, f = self.cfg, self.cfg.f
        tree = hu.build_tree(cfg, self.vv_update, vv_state, imm, step_size, rng_key,
                             max_delta_energy=self.max_delta_energy,
                             max_tree_depth=(max_treedepth_current, max(self.max_treedepth)),
                             trace=self.trace)
        with np.errstate(all="ignore"):
            accept_prob = f(tree.sum_accept_probs / f(tree.num_proposals))
        vv_state = IntegratorState(tree.z_proposal, vv_state.r, tree.z_proposal_pe,
                                   tree.z_proposal_grad)
        return vv_state, tree.z_proposal_energy, tree.num_proposals, accept_prob, tree.diverging

    # -- sample_kernel (hmc.py:459-530)
    def sample(self, hmc_state):
        cfg, f = self.cfg, self.cfg.f
        keys = cfg.split(hmc_state.rng_key, 3)
        rng_key, rng_key_momentum, rng_key_transition = keys[0], keys[1], keys[2]
        r = momentum_generator(cfg, hmc_state.adapt_state.mass_matrix_sqrt, rng_key_momentum)
        vv_state = IntegratorState(hmc_state.z, r, hmc_state.potential_energy, hmc_state.z_grad)
        if self.algo == "HMC":
            vv_state, energy, num_steps, accept_prob, diverging = self._hmc_next(
                hmc_state.adapt_state.step_size, hmc_state.adapt_state.inverse_mass_matrix,
                vv_state, rng_key_transition, hmc_state.trajectory_length)
        else:
            depth = self.max_treedepth[0] if hmc_state.i < self.wa_steps else self.max_treedepth[1]
            vv_state, energy, num_steps, accept_prob, diverging = self._nuts_next(
                hmc_state.adapt_state.step_size, hmc_state.adapt_state.inverse_mass_matrix,
                vv_state, rng_key_transition, depth)
        if hmc_state.i < self.wa_steps:
            adapt_state = self.wa_update(hmc_state.i, accept_prob, vv_state, hmc_state.adapt_state)
        else:
            adapt_state = hmc_state.adapt_state
        itr = hmc_state.i + 1
        n = itr if hmc_state.i < self.wa_steps else itr - self.wa_steps
        with np.errstate(all="ignore"):
            mean_accept_prob = f(hmc_state.mean_accept_prob
                                 + (accept_prob - hmc_state.mean_accept_prob) / f(n))
        return HMCState(itr, vv_state.z, vv_state.z_grad, f(vv_state.potential_energy), f(energy),
                        None, hmc_state.trajectory_length, num_steps, accept_prob,
                        mean_accept_prob, bool(diverging), adapt_state, rng_key)


FIELDS = ("z", "potential_energy", "energy", "num_steps", "accept_prob", "mean_accept_prob",
          "diverging", "step_size")


def run_chain(kernel, rng_key, num_warmup, num_samples, thinning=1, init_params=None,
              init_values=None, collect_warmup=False):
    """One chain of ``MCMC.run`` (mcmc.py:461-513) through ``fori_collect``
    (util.py:277-407): collected from iteration ``lower`` with thinning
    ``start_idx = lower + (upper - lower) % thinning``."""
    state = kernel.init(rng_key, num_warmup, init_params, init_values)
    lower = 0 if collect_warmup else num_warmup
    upper = num_warmup + num_samples
    start_idx = lower + (upper - lower) % thinning
    size = (upper - lower) // thinning
    out = {k: [None] * size for k in FIELDS}
    for i in range(upper):
        state = kernel.sample(state)
        idx = (i - start_idx) // thinning
        # util.py:334-351: every i >= start_idx writes slot idx, so the LAST draw of each
        # thinning block is the one kept.
        if i >= start_idx and 0 <= idx < size:
            out["z"][idx] = state.z.copy()
            for k in FIELDS[1:-1]:
                out[k][idx] = getattr(state, k)
            out["step_size"][idx] = state.adapt_state.step_size
    res = {k: np.asarray(v) for k, v in out.items()}
    return res, state


def run_mcmc(family, rng_key, num_warmup, num_samples, num_chains=1, cfg=None, **kernel_kwargs):
    """``MCMC.run`` key handling (mcmc.py:661-662): split into per-chain keys when
    ``num_chains > 1``; returns arrays stacked ``[chain, draw, ...]``."""
    cfg = cfg if cfg is not None else Config()
    keys = cfg.split(rng_key, num_chains) if num_chains > 1 else [np.asarray(rng_key)]
    results, states, grad_evals = [], [], 0
    for c in range(num_chains):
        kernel = HMCKernel(family, cfg=cfg, **kernel_kwargs)
        res, st = run_chain(kernel, keys[c], num_warmup, num_samples)
        results.append(res)
        states.append(st)
        grad_evals += kernel.num_grad_evals
    stacked = {k: np.stack([r[k] for r in results]) for k in results[0]}
    return stacked, states, grad_evals
