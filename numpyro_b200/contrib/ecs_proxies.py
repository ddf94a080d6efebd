"""``numpyro.contrib.ecs_proxies`` surface: the Taylor control variate of ``HMCECS``
(``contrib/ecs_proxies.py:94-278``).  The proxy itself is evaluated inside the CUDA potential
(``csrc/models.cuh`` ``logistic_ecs_potential``); this is the marker the wrapper takes."""
from ..infer.hmc_gibbs import TaylorProxy, taylor_proxy

__all__ = ["taylor_proxy", "TaylorProxy"]
