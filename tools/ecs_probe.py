"""Timing probe for the HMCECS device loop: transitions/s against the subsample size m."""
import sys, time, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as g
g.build()
import bench
from numpyro_b200 import models
from numpyro_b200.infer import HMCECS, NUTS
from numpyro_b200.random import PRNGKey

X, y = bench.make_covtype_shape(n=100000)
ref = np.zeros(X.shape[1], np.float32)
for m, deg in ((1000, 2), (250, 2), (50, 2), (1000, 0), (1000, 1)):
    proxy = HMCECS.taylor_proxy({"coefs": ref}, deg) if deg else None
    k = HMCECS(NUTS(models.logistic_regression), num_blocks=10, proxy=proxy)
    st = k.init(PRNGKey(1), 300, None, (X, y), {"subsample_size": m})
    _, st = k.run_loop(st, 300, 0, 1, 0, [], (X, y), {"subsample_size": m})
    t = time.perf_counter()
    raw, st = k.run_loop(st, 400, 0, 1, 400, ["hmc_state.num_steps"], (X, y), {"subsample_size": m})
    dt = time.perf_counter() - t
    ns = raw["hmc_state.num_steps"].mean()
    print(f"m={m} degree={deg}: {400 / dt:.0f} transitions/s, {dt / 400 * 1e6:.1f} us per transition, "
          f"mean leapfrogs {ns:.1f}, {dt / 400 / (ns + 1) * 1e6:.1f} us per potential", flush=True)
