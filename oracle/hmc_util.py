"""Oracle restatement of ``numpyro/infer/hmc_util.py`` in NumPy.

TEST INFRASTRUCTURE (see ``oracle/__init__.py``).  Every function cites the
reference lines it follows.  State is held as flat vectors (``ravel_pytree`` of
the site dict in sorted-name order); the mass matrix is one block, 1-D
(diagonal) or 2-D (dense), or -- structured ``dense_mass`` lists / dict-valued
``inverse_mass_matrix`` (``hmc_util.py:439-486``) -- a dict keyed by site-name tuples whose
blocks address the flat vector through ``cfg.site_index``.  Arithmetic is done in ``dtype`` (float32 by default,
like the reference with x64 disabled); Python loops play the role of
``lax.while_loop``/``cond`` exactly as the reference's own eager mode does
(``numpyro/util.py:104-151`` ``control_flow_prims_disabled``).
"""
from collections import namedtuple

import numpy as np

from . import threefry as rnd

AdaptWindow = namedtuple("AdaptWindow", ["start", "end"])  # hmc_util.py:16
HMCAdaptState = namedtuple(  # hmc_util.py:18-30
    "HMCAdaptState",
    ["step_size", "inverse_mass_matrix", "mass_matrix_sqrt", "mass_matrix_sqrt_inv",
     "ss_state", "mm_state", "window_idx", "rng_key"],
)
IntegratorState = namedtuple("IntegratorState", ["z", "r", "potential_energy", "z_grad"])
TreeInfo = namedtuple(  # hmc_util.py:36-57
    "TreeInfo",
    ["z_left", "r_left", "z_left_grad", "z_right", "r_right", "z_right_grad",
     "z_proposal", "z_proposal_pe", "z_proposal_grad", "z_proposal_energy",
     "depth", "weight", "r_sum", "turning", "diverging", "sum_accept_probs",
     "num_proposals"],
)


class Config:
    """Numerics switches shared by the oracle functions."""

    def __init__(self, dtype=np.float32, partitionable=True):
        self.dtype = np.dtype(dtype)
        self.f = self.dtype.type
        self.partitionable = partitionable
        self.site_index = {}  # site name -> flat indices (set by oracle.hmc.HMCKernel)

    def block_index(self, site_names):
        """Flat indices of a mass-matrix block, in the block's OWN site order
        (``tuple(z[k] for k in site_names)`` then ``ravel_pytree``, hmc_util.py:182-186)."""
        return np.concatenate([self.site_index[k] for k in site_names])

    def split(self, key, n=2):
        return rnd.split(key, n, partitionable=self.partitionable)

    def random_bits(self, key, shape=()):
        return rnd.random_bits(key, shape, partitionable=self.partitionable)

    def uniform(self, key, shape=(), lo=0.0, hi=1.0):
        return rnd.uniform(key, shape, lo, hi, partitionable=self.partitionable).astype(self.dtype)

    def normal(self, key, shape=()):
        return rnd.normal(key, shape, partitionable=self.partitionable).astype(self.dtype)

    def bernoulli(self, key, p=0.5):
        # jax.random.bernoulli draws uniform in the dtype of p (float32 here)
        u = rnd.uniform(key, (), partitionable=self.partitionable)
        return bool(u < np.float32(p))


# ---------------------------------------------------------------- dual averaging
def dual_averaging(cfg, t0=10, kappa=0.75, gamma=0.05):
    """hmc_util.py:60-130."""
    f = cfg.f

    def init_fn(prox_center=0.0):
        return f(0), f(0), f(0), 0, f(prox_center)

    def update_fn(g, state):
        x_t, x_avg, g_avg, t, prox_center = state
        g = f(g)
        t = t + 1
        inv = f(1) / f(t + t0)
        g_avg = (f(1) - inv) * g_avg + g / f(t + t0)
        x_t = prox_center - (np.sqrt(f(t)) / f(gamma)) * g_avg
        weight_t = f(np.power(f(t), f(-kappa)))
        x_avg = (f(1) - weight_t) * x_avg + weight_t * x_t
        return f(x_t), f(x_avg), f(g_avg), t, prox_center

    return init_fn, update_fn


def _cholesky(a):
    """``jnp.linalg.cholesky``: NaN-filled instead of raising when not positive definite."""
    try:
        return np.linalg.cholesky(a)
    except np.linalg.LinAlgError:
        return np.full_like(a, np.nan)


# ---------------------------------------------------------------- welford
def welford_covariance(cfg, diagonal=True):
    """hmc_util.py:133-239."""
    f, dt = cfg.f, cfg.dtype

    def init_fn(size):
        if isinstance(size, dict):  # hmc_util.py:156-160: block -> SHAPE of its mass matrix
            return {k: init_fn(v) for k, v in size.items()}
        if isinstance(size, (int, np.integer)):
            shape = (size,) if diagonal else (size, size)
        else:
            shape = tuple(size)
        return np.zeros(shape[-1], dt), np.zeros(shape, dt), 0

    def update_fn(sample, state):
        if isinstance(state, dict):  # hmc_util.py:176-182
            return {k: update_fn(sample[cfg.block_index(k)], st) for k, st in state.items()}
        mean, m2, n = state
        n = n + 1
        delta_pre = sample - mean
        mean = mean + delta_pre / f(n)
        delta_post = sample - mean
        if m2.ndim == 1:
            m2 = m2 + delta_pre * delta_post
        else:
            m2 = m2 + np.outer(delta_post, delta_pre)
        return mean.astype(dt), m2.astype(dt), n

    def final_fn(state, regularize=False):
        if isinstance(state, dict):  # hmc_util.py:205-215
            out = {k: final_fn(st, regularize=regularize) for k, st in state.items()}
            return tuple({k: v[i] for k, v in out.items()} for i in range(3))
        mean, m2, n = state
        with np.errstate(divide="ignore", invalid="ignore"):
            cov = m2 / f(n - 1)
        if regularize:
            scaled_cov = (f(n) / f(n + 5)) * cov
            shrinkage = f(1e-3) * (f(5) / f(n + 5))
            if scaled_cov.ndim == 1:
                cov = scaled_cov + shrinkage
            else:
                cov = scaled_cov + shrinkage * np.identity(mean.shape[0], dt)
        cov = cov.astype(dt)
        if cov.ndim == 2:
            # in ``dt`` like the reference (LAPACK potrf/trtrs in f32 with x64 off)
            tril_inv = np.ascontiguousarray(np.swapaxes(
                _cholesky(cov[::-1, ::-1])[::-1, ::-1], -2, -1))
            from scipy.linalg import solve_triangular

            cov_inv_sqrt = solve_triangular(tril_inv, np.identity(cov.shape[-1], dt), lower=True)
            tril_inv, cov_inv_sqrt = tril_inv.astype(dt), cov_inv_sqrt.astype(dt)
        else:
            tril_inv = np.sqrt(cov)
            cov_inv_sqrt = (f(1) / tril_inv).astype(dt)
        return cov, cov_inv_sqrt, tril_inv

    return init_fn, update_fn, final_fn


# ---------------------------------------------------------------- kinetic energy
def kinetic_grad(inverse_mass_matrix, r, cfg=None):
    """hmc_util.py:1203-1220 (``M^-1 r``)."""
    if isinstance(inverse_mass_matrix, dict):
        v = np.empty_like(r)
        for names, imm in inverse_mass_matrix.items():
            idx = cfg.block_index(names)
            v[idx] = kinetic_grad(imm, r[idx])
        return v
    if inverse_mass_matrix.ndim == 2:
        return inverse_mass_matrix @ r
    return inverse_mass_matrix * r


def kinetic_energy(cfg, inverse_mass_matrix, r):
    """hmc_util.py:1183-1200."""
    if isinstance(inverse_mass_matrix, dict):
        ke = cfg.f(0)
        for names, imm in inverse_mass_matrix.items():
            ke = cfg.f(ke + kinetic_energy(cfg, imm, r[cfg.block_index(names)]))
        return ke
    v = kinetic_grad(inverse_mass_matrix, r)
    return cfg.f(0.5) * cfg.f(np.dot(v, r))


# ---------------------------------------------------------------- velocity verlet
def velocity_verlet(cfg, potential_fn):
    """hmc_util.py:262-311.  ``potential_fn(z) -> (U, grad)``."""
    f = cfg.f

    def init_fn(z, r, potential_energy=None, z_grad=None):
        if potential_energy is None or z_grad is None:
            potential_energy, z_grad = potential_fn(z)
        return IntegratorState(z, r, potential_energy, z_grad)

    def update_fn(step_size, inverse_mass_matrix, state):
        z, r, _, z_grad = state
        step_size = f(step_size)
        r = r - f(0.5) * step_size * z_grad
        r_grad = kinetic_grad(inverse_mass_matrix, r, cfg)
        z = z + step_size * r_grad
        potential_energy, z_grad = potential_fn(z)
        r = r - f(0.5) * step_size * z_grad
        return IntegratorState(z, r, potential_energy, z_grad)

    return init_fn, update_fn


# ---------------------------------------------------------------- step-size heuristic
def find_reasonable_step_size(cfg, vv_update, momentum_generator, init_step_size,
                              inverse_mass_matrix, z_info, rng_key):
    """hmc_util.py:314-384.  ``momentum_generator(inverse_mass_matrix, key)``: the reference
    hands the INVERSE mass matrix to its momentum generator here (``:362``), restated as is.
    ``z_info`` = IntegratorState with potential_energy and z_grad set."""
    f = cfg.f
    target = f(np.log(f(0.8)))
    finfo = np.finfo(cfg.dtype)
    z, _, potential_energy, z_grad = z_info
    step_size, last_direction, direction = f(init_step_size), 0, 0

    def cond(step_size, last_direction, direction):
        not_small = (step_size > finfo.tiny) or (direction >= 0)
        not_large = (step_size < finfo.max) or (direction <= 0)
        return not_small and not_large and (last_direction == 0 or direction == last_direction)

    while cond(step_size, last_direction, direction):
        rng_key, rng_key_momentum = cfg.split(rng_key)
        with np.errstate(all="ignore"):
            step_size = f(f(2.0) ** direction) * step_size
            r = momentum_generator(inverse_mass_matrix, rng_key_momentum)
            new = vv_update(step_size, inverse_mass_matrix,
                            IntegratorState(z, r, potential_energy, z_grad))
            energy_current = f(kinetic_energy(cfg, inverse_mass_matrix, r) + f(potential_energy))
            energy_new = f(kinetic_energy(cfg, inverse_mass_matrix, new.r)
                           + f(new.potential_energy))
            delta_energy = f(energy_new - energy_current)
        direction_new = 1 if target < -delta_energy else -1
        last_direction, direction = direction, direction_new
    return step_size


# ---------------------------------------------------------------- schedule
def build_adaptation_schedule(num_steps):
    """hmc_util.py:387-436 (integer, host)."""
    schedule = []
    if num_steps < 20:
        schedule.append(AdaptWindow(0, num_steps - 1))
        return schedule
    start_buffer_size, end_buffer_size, init_window_size = 75, 50, 25
    if (start_buffer_size + end_buffer_size + init_window_size) > num_steps:
        start_buffer_size = int(0.15 * num_steps)
        end_buffer_size = int(0.1 * num_steps)
        init_window_size = num_steps - start_buffer_size - end_buffer_size
    schedule.append(AdaptWindow(0, start_buffer_size - 1))
    end_window_start = num_steps - end_buffer_size
    next_window_size = init_window_size
    next_window_start = start_buffer_size
    while next_window_start < end_window_start:
        cur_window_start, cur_window_size = next_window_start, next_window_size
        if 3 * cur_window_size <= end_window_start - cur_window_start:
            next_window_size = 2 * cur_window_size
        else:
            cur_window_size = end_window_start - cur_window_start
        next_window_start = cur_window_start + cur_window_size
        schedule.append(AdaptWindow(cur_window_start, next_window_start - 1))
    schedule.append(AdaptWindow(end_window_start, num_steps - 1))
    return schedule


# ---------------------------------------------------------------- mass matrix init
def initialize_mass_matrix(cfg, size, inverse_mass_matrix, dense_mass):
    """hmc_util.py:439-515.  ``dense_mass`` a list of site-name tuples -> block dicts, in the
    reference's insertion order: the dense blocks, then the user's other blocks (diagonal),
    then ONE diagonal block of the remaining sites in sorted order."""
    dt, f = cfg.dtype, cfg.f
    if isinstance(dense_mass, list):
        sites = sorted(cfg.site_index)
        if inverse_mass_matrix is None:
            imm = {}
        elif not isinstance(inverse_mass_matrix, dict):
            imm = {tuple(sites): inverse_mass_matrix}
        else:
            imm = dict(inverse_mass_matrix)
        sqrt, sqrt_inv = {}, {}
        for names in dense_mass:
            imm[names], sqrt[names], sqrt_inv[names] = initialize_mass_matrix(
                cfg, cfg.block_index(names).size, imm.get(names), True)
        for names in list(imm):
            if names in dense_mass:
                continue
            imm[names], sqrt[names], sqrt_inv[names] = initialize_mass_matrix(
                cfg, cfg.block_index(names).size, imm[names], False)
        remaining = tuple(sorted(set(sites) - set().union(*imm) if imm else set(sites)))
        if len(remaining) > 0:
            imm[remaining], sqrt[remaining], sqrt_inv[remaining] = initialize_mass_matrix(
                cfg, cfg.block_index(remaining).size, None, False)
        assert sorted(k for names in imm for k in names) == sites, (
            "There seems to be a conflict of sites names specified in the initial"
            " `inverse_mass_matrix` and in `dense_mass` argument.")
        return imm, sqrt, sqrt_inv
    if inverse_mass_matrix is None:
        imm = np.identity(size, dt) if dense_mass else np.ones(size, dt)
        return imm, imm, imm
    imm = np.asarray(inverse_mass_matrix, dt)
    if dense_mass:
        if imm.ndim == 1:
            imm = np.diag(imm)
        sqrt_inv = np.ascontiguousarray(np.swapaxes(
            _cholesky(imm[::-1, ::-1])[::-1, ::-1], -2, -1))
        from scipy.linalg import solve_triangular

        sqrt = solve_triangular(sqrt_inv, np.identity(imm.shape[-1], dt), lower=True)
        return imm, sqrt.astype(dt), sqrt_inv.astype(dt)
    if imm.ndim == 2:
        imm = np.diag(imm).copy()
    sqrt_inv = np.sqrt(imm)
    return imm, (f(1) / sqrt_inv).astype(dt), sqrt_inv


# ---------------------------------------------------------------- warmup adapter
def warmup_adapter(cfg, num_adapt_steps, find_reasonable_step_size=None, adapt_step_size=True,
                   adapt_mass_matrix=True, dense_mass=False, target_accept_prob=0.8,
                   regularize_mass_matrix=True):
    """hmc_util.py:518-707."""
    f = cfg.f
    if find_reasonable_step_size is None:
        find_reasonable_step_size = lambda step_size, imm, z_info, key: step_size  # noqa: E731
    ss_init, ss_update = dual_averaging(cfg)
    mm_init, mm_update, mm_final = welford_covariance(cfg, diagonal=not dense_mass)
    schedule = build_adaptation_schedule(num_adapt_steps)
    num_windows = len(schedule)

    def init_fn(z_info, rng_key, step_size=1.0, inverse_mass_matrix=None):
        rng_key, rng_key_ss = cfg.split(rng_key)
        size = np.size(z_info[0])
        imm, mm_sqrt, mm_sqrt_inv = initialize_mass_matrix(cfg, size, inverse_mass_matrix,
                                                           dense_mass)
        step_size = f(step_size)
        if adapt_step_size:
            step_size = f(find_reasonable_step_size(step_size, imm, z_info, rng_key_ss))
        with np.errstate(over="ignore"):
            ss_state = ss_init(np.log(f(10) * step_size))
        if isinstance(imm, dict):  # hmc_util.py:577-580
            size = {k: v.shape for k, v in imm.items()}
        mm_state = mm_init(size)
        return HMCAdaptState(step_size, imm, mm_sqrt, mm_sqrt_inv, ss_state, mm_state, 0, rng_key)

    def _update_at_window_end(z_info, rng_key_ss, state):
        step_size, imm, mm_sqrt, mm_sqrt_inv, ss_state, mm_state, window_idx, rng_key = state
        if adapt_mass_matrix:
            imm, mm_sqrt, mm_sqrt_inv = mm_final(mm_state, regularize=regularize_mass_matrix)
            mm_state = mm_init({k: v.shape for k, v in imm.items()} if isinstance(imm, dict)
                               else imm.shape[-1])
        if adapt_step_size:
            step_size = f(find_reasonable_step_size(step_size, imm, z_info, rng_key_ss))
            ss_state = ss_init(f(np.log(f(10))) + np.log(step_size))
        return HMCAdaptState(step_size, imm, mm_sqrt, mm_sqrt_inv, ss_state, mm_state,
                             window_idx, rng_key)

    def update_fn(t, accept_prob, z_info, state):
        step_size, imm, mm_sqrt, mm_sqrt_inv, ss_state, mm_state, window_idx, rng_key = state
        if rng_key is not None:
            rng_key, rng_key_ss = cfg.split(rng_key)
        else:
            rng_key_ss = None
        if adapt_step_size:
            ss_state = ss_update(f(target_accept_prob) - f(accept_prob), ss_state)
            log_step_size, log_step_size_avg = ss_state[0], ss_state[1]
            with np.errstate(over="ignore"):
                step_size = np.exp(log_step_size_avg) if t == (num_adapt_steps - 1) else np.exp(
                    log_step_size)
            finfo = np.finfo(cfg.dtype)
            step_size = f(np.clip(step_size, finfo.tiny, finfo.max))
        is_middle_window = (0 < window_idx) and (window_idx < (num_windows - 1))
        if adapt_mass_matrix and is_middle_window:
            mm_state = mm_update(z_info[0], mm_state)
        t_at_window_end = t == schedule[window_idx].end
        window_idx = window_idx + 1 if t_at_window_end else window_idx
        state = HMCAdaptState(step_size, imm, mm_sqrt, mm_sqrt_inv, ss_state, mm_state,
                              window_idx, rng_key)
        if t_at_window_end and is_middle_window:
            state = _update_at_window_end(z_info, rng_key_ss, state)
        return state

    return init_fn, update_fn


# ---------------------------------------------------------------- U-turn
def momentum_angle(cfg, inverse_mass_matrix, r_left, r_right, r_sum):
    """hmc_util.py:710-737."""
    f = cfg.f
    if isinstance(inverse_mass_matrix, dict):  # per block, then summed (:711-720)
        left_angle, right_angle = f(0), f(0)
        for names, imm in inverse_mass_matrix.items():
            idx = cfg.block_index(names)
            la, ra = momentum_angle(cfg, imm, r_left[idx], r_right[idx], r_sum[idx])
            left_angle, right_angle = f(left_angle + la), f(right_angle + ra)
        return left_angle, right_angle
    v_left = kinetic_grad(inverse_mass_matrix, r_left)
    v_right = kinetic_grad(inverse_mass_matrix, r_right)
    r_sum = r_sum - (r_left + r_right) / f(2)
    return f(np.dot(v_left, r_sum)), f(np.dot(v_right, r_sum))


def is_turning(cfg, inverse_mass_matrix, r_left, r_right, r_sum):
    """hmc_util.py:740-746 (generalised criterion, ``<= 0``)."""
    left_angle, right_angle = momentum_angle(cfg, inverse_mass_matrix, r_left, r_right, r_sum)
    return bool((left_angle <= 0) | (right_angle <= 0))


def _expit(cfg, x):
    f = cfg.f
    with np.errstate(over="ignore"):
        return f(1) / (f(1) + np.exp(-f(x)))


def uniform_transition_kernel(cfg, current_tree, new_tree):
    """hmc_util.py:749-753."""
    return _expit(cfg, new_tree.weight - current_tree.weight)


def biased_transition_kernel(cfg, current_tree, new_tree):
    """hmc_util.py:756-764."""
    f = cfg.f
    with np.errstate(over="ignore", invalid="ignore"):
        p = np.exp(f(new_tree.weight - current_tree.weight))
    if new_tree.turning or new_tree.diverging:
        return f(0.0)
    return f(min(p, f(1.0))) if not np.isnan(p) else f(np.nan)


def combine_tree(cfg, current_tree, new_tree, inverse_mass_matrix, going_right, rng_key,
                 biased_transition, trace=None):
    """hmc_util.py:767-848."""
    f = cfg.f
    if going_right:
        z_left, r_left, z_left_grad = current_tree.z_left, current_tree.r_left, current_tree.z_left_grad
        z_right, r_right, z_right_grad = new_tree.z_right, new_tree.r_right, new_tree.z_right_grad
    else:
        z_left, r_left, z_left_grad = new_tree.z_left, new_tree.r_left, new_tree.z_left_grad
        z_right, r_right, z_right_grad = (current_tree.z_right, current_tree.r_right,
                                          current_tree.z_right_grad)
    r_sum = current_tree.r_sum + new_tree.r_sum
    if biased_transition:
        transition_prob = biased_transition_kernel(cfg, current_tree, new_tree)
        turning = new_tree.turning | is_turning(cfg, inverse_mass_matrix, r_left, r_right, r_sum)
    else:
        transition_prob = uniform_transition_kernel(cfg, current_tree, new_tree)
        turning = current_tree.turning
    transition = cfg.bernoulli(rng_key, transition_prob)
    if trace is not None:
        trace.append(("biased" if biased_transition else "uniform", bool(transition)))
    src = new_tree if transition else current_tree
    with np.errstate(over="ignore", invalid="ignore"):
        tree_weight = f(np.logaddexp(f(current_tree.weight), f(new_tree.weight)))
    return TreeInfo(
        z_left, r_left, z_left_grad, z_right, r_right, z_right_grad,
        src.z_proposal, src.z_proposal_pe, src.z_proposal_grad, src.z_proposal_energy,
        current_tree.depth + 1, tree_weight, r_sum, bool(turning), bool(new_tree.diverging),
        f(current_tree.sum_accept_probs + new_tree.sum_accept_probs),
        current_tree.num_proposals + new_tree.num_proposals,
    )


def build_basetree(cfg, vv_update, z, r, z_grad, inverse_mass_matrix, step_size, going_right,
                   energy_current, max_delta_energy):
    """hmc_util.py:851-894: one leapfrog = one gradient evaluation."""
    f = cfg.f
    step_size = f(step_size) if going_right else -f(step_size)
    z_new, r_new, pe_new, z_new_grad = vv_update(step_size, inverse_mass_matrix,
                                                 (z, r, energy_current, z_grad))
    with np.errstate(over="ignore", invalid="ignore"):
        energy_new = f(pe_new + kinetic_energy(cfg, inverse_mass_matrix, r_new))
        delta_energy = f(energy_new - energy_current)
        if np.isnan(delta_energy):
            delta_energy = f(np.inf)
        tree_weight = -delta_energy
        diverging = bool(delta_energy > f(max_delta_energy))
        accept_prob = f(min(np.exp(-delta_energy), f(1.0)))
    return TreeInfo(z_new, r_new, z_new_grad, z_new, r_new, z_new_grad, z_new, pe_new,
                    z_new_grad, energy_new, 0, tree_weight, r_new, False, diverging,
                    accept_prob, 1)


def leaf_idx_to_ckpt_idxs(n):
    """hmc_util.py:941-958 (integer)."""
    n = int(n)
    idx_max = bin(n >> 1).count("1")
    num_subtrees = bin(((~n) & (n + 1)) - 1).count("1")
    idx_min = idx_max - num_subtrees + 1
    return idx_min, idx_max


def is_iterative_turning(cfg, inverse_mass_matrix, r, r_sum, r_ckpts, r_sum_ckpts, idx_min,
                         idx_max):
    """hmc_util.py:961-981."""
    i, turning = idx_max, False
    while i >= idx_min and not turning:
        subtree_r_sum = r_sum - r_sum_ckpts[i] + r_ckpts[i]
        turning = is_turning(cfg, inverse_mass_matrix, r_ckpts[i], r, subtree_r_sum)
        i -= 1
    return turning


def iterative_build_subtree(cfg, prototype_tree, vv_update, inverse_mass_matrix, step_size,
                            going_right, rng_key, energy_current, max_delta_energy, r_ckpts,
                            r_sum_ckpts, trace=None):
    """hmc_util.py:984-1085 (LOOP B)."""
    max_num_proposals = 2 ** prototype_tree.depth
    tree = prototype_tree._replace(num_proposals=0)
    turning = False
    # NB: the reference's while_loop carries copies of the checkpoint arrays; the
    # caller's (all-zero) arrays are never written (hmc_util.py:1063-1065).
    r_ckpts = r_ckpts.copy()
    r_sum_ckpts = r_sum_ckpts.copy()
    while tree.num_proposals < max_num_proposals and not turning and not tree.diverging:
        rng_key, transition_rng_key = cfg.split(rng_key)
        if going_right:
            z, r, z_grad = tree.z_right, tree.r_right, tree.z_right_grad
        else:
            z, r, z_grad = tree.z_left, tree.r_left, tree.z_left_grad
        new_leaf = build_basetree(cfg, vv_update, z, r, z_grad, inverse_mass_matrix, step_size,
                                  going_right, energy_current, max_delta_energy)
        if tree.num_proposals == 0:
            new_tree = new_leaf
        else:
            new_tree = combine_tree(cfg, tree, new_leaf, inverse_mass_matrix, going_right,
                                    transition_rng_key, False, trace)
        leaf_idx = tree.num_proposals
        ckpt_idx_min, ckpt_idx_max = leaf_idx_to_ckpt_idxs(leaf_idx)
        r = new_leaf.r_right
        r_sum = new_tree.r_sum
        if leaf_idx % 2 == 0:
            r_ckpts[ckpt_idx_max] = r
            r_sum_ckpts[ckpt_idx_max] = r_sum
        turning = is_iterative_turning(cfg, inverse_mass_matrix, new_leaf.r_right, r_sum, r_ckpts,
                                       r_sum_ckpts, ckpt_idx_min, ckpt_idx_max)
        tree = new_tree
    return tree._replace(depth=prototype_tree.depth, turning=turning)


def double_tree(cfg, current_tree, vv_update, inverse_mass_matrix, step_size, going_right,
                rng_key, energy_current, max_delta_energy, r_ckpts, r_sum_ckpts, trace=None):
    """hmc_util.py:907-938."""
    key, transition_key = cfg.split(rng_key)
    new_tree = iterative_build_subtree(cfg, current_tree, vv_update, inverse_mass_matrix,
                                       step_size, going_right, key, energy_current,
                                       max_delta_energy, r_ckpts, r_sum_ckpts, trace)
    return combine_tree(cfg, current_tree, new_tree, inverse_mass_matrix, going_right,
                        transition_key, True, trace)


def build_tree(cfg, verlet_update, verlet_state, inverse_mass_matrix, step_size, rng_key,
               max_delta_energy=1000.0, max_tree_depth=10, trace=None):
    """hmc_util.py:1088-1180 (LOOP A)."""
    f = cfg.f
    if isinstance(max_tree_depth, tuple):
        max_tree_depth_current, max_tree_depth = max_tree_depth
    else:
        max_tree_depth_current = max_tree_depth
    z, r, potential_energy, z_grad = verlet_state
    energy_current = f(potential_energy + kinetic_energy(cfg, inverse_mass_matrix, r))
    latent_size = np.size(r)
    r_ckpts = np.zeros((max_tree_depth, latent_size), cfg.dtype)
    r_sum_ckpts = np.zeros((max_tree_depth, latent_size), cfg.dtype)
    tree = TreeInfo(z, r, z_grad, z, r, z_grad, z, potential_energy, z_grad, energy_current,
                    0, f(0), r, False, False, f(0), 0)
    key = rng_key
    while tree.depth < max_tree_depth_current and not tree.turning and not tree.diverging:
        keys = cfg.split(key, 3)
        key, direction_key, doubling_key = keys[0], keys[1], keys[2]
        going_right = cfg.bernoulli(direction_key)
        if trace is not None:
            trace.append(("direction", bool(going_right)))
        tree = double_tree(cfg, tree, verlet_update, inverse_mass_matrix, step_size, going_right,
                           doubling_key, energy_current, max_delta_energy, r_ckpts, r_sum_ckpts,
                           trace)
    return tree
