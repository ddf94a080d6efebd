"""Oracle restatement of ``numpyro/diagnostics.py`` (NumPy) + rank-normalised bulk ESS.

TEST / MEASUREMENT INFRASTRUCTURE (see ``oracle/__init__.py``): used by the parity
tests (tier 4) and by ``bench.py`` to turn draws into the min-bulk-ESS/sec metric.
The reference has no bulk ESS: ``bulk_ess`` feeds rank-normalised
split chains (Vehtari et al. 2021) through the restated Geyer/Stan estimator.
"""
import numpy as np


def _compute_chain_variance_stats(x):
    """diagnostics.py:29-42."""
    C, N = x.shape[:2]
    chain_var = x.var(axis=1, ddof=1)
    var_within = chain_var.mean(axis=0)
    var_estimator = var_within * (N - 1) / N
    if x.shape[0] > 1:
        chain_mean = x.mean(axis=1)
        var_between = chain_mean.var(axis=0, ddof=1)
        var_estimator = var_estimator + var_between
    else:
        var_within = var_estimator
    return var_within, var_estimator


def gelman_rubin(x):
    """diagnostics.py:45-61."""
    assert x.ndim >= 2 and x.shape[0] >= 2 and x.shape[1] >= 2
    var_within, var_estimator = _compute_chain_variance_stats(x)
    with np.errstate(invalid="ignore", divide="ignore"):
        return np.sqrt(var_estimator / var_within)


def split_gelman_rubin(x):
    """diagnostics.py:64-80."""
    assert x.ndim >= 2 and x.shape[1] >= 4
    n_half = x.shape[1] // 2
    return gelman_rubin(np.concatenate([x[:, :n_half], x[:, -n_half:]], axis=0))


def _fft_next_fast_len(target):
    """diagnostics.py:83-98."""
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


def autocorrelation(x, axis=0, bias=True):
    """diagnostics.py:101-142."""
    N = x.shape[axis]
    M2 = 2 * _fft_next_fast_len(N)
    x = np.swapaxes(x, axis, -1)
    centered = x - x.mean(axis=-1, keepdims=True)
    freqvec = np.fft.rfft(centered, n=M2, axis=-1)
    gram = freqvec * np.conjugate(freqvec)
    autocorr = np.fft.irfft(gram, n=M2, axis=-1)[..., :N]
    if not bias:
        autocorr = autocorr / np.arange(N, 0.0, -1)
    with np.errstate(invalid="ignore", divide="ignore"):
        autocorr = (autocorr / autocorr[..., :1]).astype(np.float64)
    return np.swapaxes(autocorr, axis, -1)


def autocovariance(x, axis=0, bias=True):
    """diagnostics.py:145-155."""
    return autocorrelation(x, axis, bias) * x.var(axis=axis, keepdims=True)


def effective_sample_size(x, bias=True):
    """diagnostics.py:158-203 (Geyer initial monotone sequence, Stan 2.18)."""
    x = np.asarray(x)
    assert x.ndim >= 2 and x.shape[1] >= 2
    gamma_k_c = autocovariance(x, axis=1, bias=bias)
    var_within, var_estimator = _compute_chain_variance_stats(x)
    rho_k = 1.0 - (var_within - gamma_k_c.mean(axis=0)) / var_estimator
    rho_k[0] = 1.0
    Rho_k = rho_k[:-1:2, ...] + rho_k[1::2, ...]
    Rho_init = Rho_k[:1]
    Rho_k = np.concatenate(
        [Rho_init, np.minimum.accumulate(np.clip(Rho_k[1:, ...], 0, None), axis=0)], axis=0)
    tau = -1.0 + 2.0 * Rho_k.sum(axis=0)
    return np.prod(x.shape[:2]) / tau


def _rank_normalise(x):
    """z-scores of fractional ranks over all chains and draws (Vehtari et al. 2021, eq. 14)."""
    from scipy.special import ndtri
    from scipy.stats import rankdata

    shape = x.shape
    flat = x.reshape((shape[0] * shape[1],) + shape[2:])
    ranks = rankdata(flat, method="average", axis=0)
    z = ndtri((ranks - 0.375) / (flat.shape[0] + 0.25))
    return z.reshape(shape)


def bulk_ess(x):
    """Rank-normalised split-chain bulk ESS; ``x`` is ``[chain, draw, ...]``."""
    x = np.asarray(x, np.float64)
    n_half = x.shape[1] // 2
    split = np.concatenate([x[:, :n_half], x[:, -n_half:]], axis=0)
    return effective_sample_size(_rank_normalise(split))


def summary_stats(x):
    """mean / sd / n_eff / split r_hat / MCSE for ``x`` of shape [chain, draw, ...]."""
    x = np.asarray(x, np.float64)
    flat = x.reshape((-1,) + x.shape[2:])
    mean, sd = flat.mean(axis=0), flat.std(axis=0, ddof=1)
    n_eff = effective_sample_size(x)
    return {"mean": mean, "sd": sd, "n_eff": n_eff, "r_hat": split_gelman_rubin(x),
            "mcse": sd / np.sqrt(np.maximum(n_eff, 1.0)), "bulk_ess": bulk_ess(x)}
