import sys, time, ctypes as C
import numpy as np
sys.path.insert(0, ".")
import __graft_entry__ as g
g.build()
import bench
from numpyro_b200 import _lib, models
from numpyro_b200.infer import MCMC, NUTS
from numpyro_b200.random import PRNGKey
lib = _lib.lib()
X, y = bench.make_covtype_shape()
key = PRNGKey(1)
kern = NUTS(models.logistic_regression)
mc = MCMC(kern, num_warmup=300, num_samples=100)
mc.warmup(key, X, y, extra_fields=("num_steps",))
state = mc.post_warmup_state
adapted_step = float(np.asarray(state.adapt_state.step_size))
imm = np.asarray(next(iter(state.adapt_state.inverse_mass_matrix.values())), np.float32)
z_last = {k: np.asarray(v) for k, v in kern._last_state.z.items()}
print("adapted", adapted_step, imm[:4], z_last["coefs"][:3])
for i in range(5):
    t0 = time.perf_counter()
    k2 = NUTS(models.logistic_regression, step_size=adapted_step, inverse_mass_matrix=imm,
              adapt_step_size=False, adapt_mass_matrix=False)
    m2 = MCMC(k2, num_warmup=0, num_samples=100)
    m2.run(key, X, y, init_params=z_last, extra_fields=("num_steps", "accept_prob", "potential_energy"))
    ns = m2.get_extra_fields()["num_steps"]
    ex = m2.get_extra_fields()
    dt = time.perf_counter() - t0
    st = m2.last_state
    print(i, "ms", round(dt * 1e3, 2), "grads", int(ns.sum()), "acc", float(ex["accept_prob"].mean()), "pe0", float(ex["potential_energy"][0]),
          "step", float(st.adapt_state.step_size), "imm", np.asarray(next(iter(st.adapt_state.inverse_mass_matrix.values())))[:3])
    k2._destroy()
