"""README example timing: eight schools, 4 chains x (1000 warm-up + 1000 samples), run from the repo root."""
import sys, time
import numpy as np
sys.path.insert(0, ".")
import __graft_entry__ as g
g.build()
from numpyro_b200 import models
from numpyro_b200.infer import MCMC, NUTS
from numpyro_b200.random import PRNGKey
y = np.array([28.0, 8.0, -3.0, 7.0, -1.0, 1.0, 18.0, 12.0]); s = np.array([15.0, 10.0, 16.0, 11.0, 9.0, 11.0, 10.0, 18.0])
for i in range(3):
    m = MCMC(NUTS(models.eight_schools_noncentered), num_warmup=1000, num_samples=1000, num_chains=4, chain_method="vectorized")
    t0 = time.perf_counter(); m.run(PRNGKey(0), 8, s, y=y, extra_fields=("num_steps",)); dt = time.perf_counter() - t0
print("4 chains x 2000 transitions: %.1f ms, grad evals %d, E[mu]=%.2f" % (dt * 1e3, int(m.get_extra_fields()["num_steps"].sum()), m.get_samples()["mu"].mean()))
