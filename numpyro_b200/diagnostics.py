"""Host-side convergence diagnostics with the surface of ``numpyro/diagnostics.py``
(NumPy; not on the accelerated path).  ``MCMC.print_summary`` uses these."""
from collections import OrderedDict
from itertools import product

import numpy as np


def _chain_variance_stats(x):  # diagnostics.py:29-42
    n = x.shape[1]
    var_within = x.var(axis=1, ddof=1).mean(axis=0)
    var_estimator = var_within * (n - 1) / n
    if x.shape[0] > 1:
        var_estimator = var_estimator + x.mean(axis=1).var(axis=0, ddof=1)
    else:
        var_within = var_estimator
    return var_within, var_estimator


def gelman_rubin(x):  # diagnostics.py:45-61
    assert x.ndim >= 2 and x.shape[0] >= 2 and x.shape[1] >= 2
    var_within, var_estimator = _chain_variance_stats(x)
    with np.errstate(invalid="ignore", divide="ignore"):
        return np.sqrt(var_estimator / var_within)


def split_gelman_rubin(x):  # diagnostics.py:64-80
    assert x.ndim >= 2 and x.shape[1] >= 4
    h = x.shape[1] // 2
    return gelman_rubin(np.concatenate([x[:, :h], x[:, -h:]], axis=0))


def _fft_next_fast_len(target):  # diagnostics.py:83-98
    if target <= 2:
        return target
    while True:
        m = target
        for p in (2, 3, 5):
            while m % p == 0:
                m //= p
        if m == 1:
            return target
        target += 1


def autocorrelation(x, axis=0, bias=True):  # diagnostics.py:101-142
    n = x.shape[axis]
    m2 = 2 * _fft_next_fast_len(n)
    x = np.swapaxes(x, axis, -1)
    centered = x - x.mean(axis=-1, keepdims=True)
    freq = np.fft.rfft(centered, n=m2, axis=-1)
    ac = np.fft.irfft(freq * np.conjugate(freq), n=m2, axis=-1)[..., :n]
    if not bias:
        ac = ac / np.arange(n, 0.0, -1)
    with np.errstate(invalid="ignore", divide="ignore"):
        ac = (ac / ac[..., :1]).astype(np.float64)
    return np.swapaxes(ac, axis, -1)


def autocovariance(x, axis=0, bias=True):  # diagnostics.py:145-155
    return autocorrelation(x, axis, bias) * x.var(axis=axis, keepdims=True)


def effective_sample_size(x, bias=True):  # diagnostics.py:158-203
    x = np.asarray(x)
    assert x.ndim >= 2 and x.shape[1] >= 2
    gamma_k_c = autocovariance(x, axis=1, bias=bias)
    var_within, var_estimator = _chain_variance_stats(x)
    rho_k = 1.0 - (var_within - gamma_k_c.mean(axis=0)) / var_estimator
    rho_k[0] = 1.0
    big_rho = rho_k[:-1:2, ...] + rho_k[1::2, ...]
    big_rho = np.concatenate(
        [big_rho[:1], np.minimum.accumulate(np.clip(big_rho[1:, ...], 0, None), axis=0)], axis=0)
    tau = -1.0 + 2.0 * big_rho.sum(axis=0)
    return np.prod(x.shape[:2]) / tau


def bulk_effective_sample_size(x):
    """Rank-normalised split-chain bulk ESS (Vehtari et al. 2021) of ``x[chain, draw, ...]``.

    Not in the reference (its ``effective_sample_size`` is the Geyer estimator on raw draws,
    diagnostics.py:158-203); this is the harness metric of SURVEY.md 8(d): the draws are split
    in halves, replaced by normal scores of their pooled fractional ranks, and fed through
    the same Geyer estimator.
    """
    from scipy.special import ndtri
    from scipy.stats import rankdata

    x = np.asarray(x, np.float64)
    n_half = x.shape[1] // 2
    halves = np.concatenate([x[:, :n_half], x[:, x.shape[1] - n_half:]], axis=0)
    flat = halves.reshape((-1,) + halves.shape[2:])
    ranks = rankdata(flat, method="average", axis=0)
    z = ndtri((ranks - 0.375) / (flat.shape[0] + 0.25)).reshape(halves.shape)
    return effective_sample_size(z)


def hpdi(x, prob=0.90, axis=0):  # diagnostics.py:206-231
    x = np.swapaxes(x, axis, 0)
    sorted_x = np.sort(x, axis=0)
    mass = x.shape[0]
    index_length = int(prob * mass)
    lengths = sorted_x[index_length:] - sorted_x[: (mass - index_length)]
    start = lengths.argmin(axis=0)
    left = np.swapaxes(np.take_along_axis(sorted_x, start[None, ...], axis=0), axis, 0)
    right = np.swapaxes(np.take_along_axis(sorted_x, (start + index_length)[None, ...], axis=0),
                        axis, 0)
    return np.concatenate([left, right], axis=axis)


def summary(samples, prob=0.90, group_by_chain=True):  # diagnostics.py:234-285
    if not isinstance(samples, dict):
        samples = {"Param:0": samples}
    if not group_by_chain:
        samples = {k: np.asarray(v)[None, ...] for k, v in samples.items()}
    out = {}
    for name, value in samples.items():
        value = np.asarray(value)
        if len(value) == 0:
            continue
        flat = np.reshape(value, (-1,) + value.shape[2:])
        hpd = hpdi(flat, prob=prob)
        out[name] = OrderedDict([
            ("mean", flat.mean(axis=0)), ("std", flat.std(axis=0, ddof=1)),
            ("median", np.median(flat, axis=0)),
            ("{:.1f}%".format(50 * (1 - prob)), hpd[0]), ("{:.1f}%".format(50 * (1 + prob)), hpd[1]),
            ("n_eff", effective_sample_size(value)), ("r_hat", split_gelman_rubin(value))])
    return out


def print_summary(samples, prob=0.90, group_by_chain=True):  # diagnostics.py:288-342
    if not isinstance(samples, dict):
        samples = {"Param:0": samples}
    if not group_by_chain:
        samples = {k: np.asarray(v)[None, ...] for k, v in samples.items()}
    summary_dict = summary(samples, prob, group_by_chain=True)
    if not summary_dict:
        return
    max_len = max(max(len(k) + 2 + 3 * (np.ndim(v) - 2) for k, v in samples.items()), 10)
    name_format = "{:>" + str(max_len) + "}"
    header_format = name_format + " {:>9}" * 7
    columns = [""] + list(list(summary_dict.values())[0].keys())
    print()
    print(header_format.format(*columns))
    row_format = name_format + " {:>9.2f}" * 7
    for name, stats in summary_dict.items():
        shape = stats["mean"].shape
        if len(shape) == 0:
            print(row_format.format(name, *stats.values()))
        else:
            for idx in product(*map(range, shape)):
                idx_str = "[{}]".format(",".join(map(str, idx)))
                print(row_format.format(name + idx_str, *[v[idx] for v in stats.values()]))
    print()


# ---------------------------------------------------------------------------- on the device
def device_ess_rhat(x, first=0, n=None, layout="chain_draw"):
    """``effective_sample_size`` / ``gelman_rubin`` / ``split_gelman_rubin``
    (``numpyro/diagnostics.py:45-80,158-203``) computed by the CUDA library
    (``b200mc_diag_ess_rhat``) -- no host NumPy on the path, so that diagnostics over 16,384
    chains x D are not the bottleneck.

    ``x``: a NumPy array ``[chain, draw, D]`` (``layout="chain_draw"``, what ``get_samples(
    group_by_chain=True)`` returns; uploaded as float32) or a ``_lib.DeviceArray`` in the
    samplers' collection layout ``[draw, chain, D]`` (``layout="draw_chain"``; nothing is copied).
    Returns ``{"n_eff", "r_hat", "split_r_hat"}`` as float64 arrays of length D
    (``r_hat`` is NaN for a single chain, as the reference's function requires two).
    """
    import ctypes as C

    from . import _lib

    lib = _lib.lib()
    if isinstance(x, _lib.DeviceArray):
        if layout != "draw_chain":
            raise ValueError("a DeviceArray must be in the collection layout [draw, chain, D]")
        xd, (nt, c, d) = x, x.shape
    else:
        x = np.asarray(x, np.float32)
        if x.ndim == 2:
            x = x[..., None]
        if layout == "chain_draw":
            x = np.swapaxes(x, 0, 1)
        x = np.ascontiguousarray(x.reshape(x.shape[0], x.shape[1], -1))
        nt, c, d = x.shape
        xd = _lib.DeviceArray.from_numpy(x)
    n = nt - first if n is None else n
    outs = [_lib.DeviceArray((d,), np.float64) for _ in range(3)]
    _lib.check(lib.b200mc_diag_ess_rhat(xd.ptr, nt, first, n, c, d, outs[0].ptr,
                                        outs[1].ptr if c > 1 else None,
                                        outs[2].ptr if n >= 4 else None, None))
    res = {"n_eff": outs[0].numpy(), "r_hat": outs[1].numpy() if c > 1 else np.full(d, np.nan),
           "split_r_hat": outs[2].numpy() if n >= 4 else np.full(d, np.nan)}
    return res
