"""``numpyro.infer.reparam`` markers for the registered families.  The reference
rewrites the traced program (``numpyro/infer/reparam.py``); here a reparametrisation selects a
closed-form potential (``numpyro_b200.models.reparam`` / ``numpyro_b200.handlers.reparam``)."""
import numpy as np


class LocScaleReparam:
    """``reparam.py:64-145``.  ``centered`` in [0, 1]: 0 fully decentres, 1 leaves the site as is.
    ``centered=None`` (a learnable per-element parameter, an SVI feature) is not available."""

    def __init__(self, centered=None, shape_params=()):
        assert centered is None or isinstance(centered, (int, float, np.generic, np.ndarray))
        assert isinstance(shape_params, (tuple, list))
        assert all(isinstance(name, str) for name in shape_params)
        if centered is None:
            raise NotImplementedError("LocScaleReparam(centered=None) learns `centered` as a parameter "
                                      "(numpyro.param); only fixed values are registered")
        if not np.all((np.asarray(centered) >= 0) & (np.asarray(centered) <= 1)):  # reparam.py:95-99
            raise ValueError("`centered` argument does not satisfy `0 <= centered <= 1`.")
        if np.ndim(centered) != 0:
            raise NotImplementedError("a per-element `centered` is not registered")
        self.centered = float(centered)
        self.shape_params = shape_params
