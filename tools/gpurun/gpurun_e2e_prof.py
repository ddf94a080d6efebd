"""Where does the end-to-end (host buffers -> draws) time of one 100-transition step go?"""
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
def pinned_copy(a):
    ptr = lib.b200mc_host_alloc(a.nbytes)
    buf = (C.c_char * a.nbytes).from_address(ptr)
    out = np.frombuffer(buf, dtype=a.dtype).reshape(a.shape); out[...] = a
    return out
X, y = pinned_copy(X), pinned_copy(y)
T = {}
def timed(name, fn):
    def w(*a, **k):
        t0 = time.perf_counter(); r = fn(*a, **k); T[name] = T.get(name, 0.0) + time.perf_counter() - t0
        return r
    return w
_lib.DeviceArray.from_numpy = classmethod(timed("from_numpy", _lib.DeviceArray.from_numpy.__func__))
orig_create, orig_init, orig_destroy = lib.b200mc_create, lib.b200mc_init, lib.b200mc_destroy
class L:  # proxy to time the C calls
    def __getattr__(self, n):
        f = getattr(lib, n)
        return timed(n, f) if n in ("b200mc_create", "b200mc_init", "b200mc_destroy", "b200mc_sample", "b200mc_get_vector", "b200mc_get_scalars", "b200mc_free", "b200mc_malloc") else f
_lib._lib = L()
for i in range(6):
    if i == 2: T.clear(); t_all = time.perf_counter()
    k2 = NUTS(models.logistic_regression, step_size=0.25, adapt_step_size=False, adapt_mass_matrix=False)
    m2 = MCMC(k2, num_warmup=0, num_samples=100)
    t0 = time.perf_counter()
    m2.run(PRNGKey(1), X, y, extra_fields=("num_steps",))
    T["run_total"] = T.get("run_total", 0) + time.perf_counter() - t0
    ns = m2.get_extra_fields()["num_steps"]; _ = m2.get_samples()["coefs"].mean()
    t0 = time.perf_counter(); k2._destroy(); T["destroy_total"] = T.get("destroy_total", 0) + time.perf_counter() - t0
print("per step (ms), 4 steps:", {k: round(1e3 * v / 4, 3) for k, v in T.items()}, "wall", round(1e3 * (time.perf_counter() - t_all) / 4, 3), "grad evals last", int(ns.sum()))
