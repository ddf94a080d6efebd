"""Per-phase timing of bench.py's end-to-end leg over many steps (run from the repo root):
which host-side phase carries the occasional 5-20 ms spikes?"""
import sys, time, gc, ctypes as C
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
kern = NUTS(models.logistic_regression)
mc = MCMC(kern, num_warmup=300, num_samples=100)
mc.warmup(PRNGKey(1), X, y)
st = mc.post_warmup_state
step = float(np.asarray(st.adapt_state.step_size)); imm = np.asarray(next(iter(st.adapt_state.inverse_mass_matrix.values())), np.float32)
z_last = {k: np.asarray(v) for k, v in kern._last_state.z.items()}
T = {}
cur = {}
def timed(name, fn):
    def w(*a, **k):
        t0 = time.perf_counter(); r = fn(*a, **k); cur[name] = cur.get(name, 0.0) + time.perf_counter() - t0
        return r
    return w
_lib.DeviceArray.from_numpy = classmethod(timed("from_numpy", _lib.DeviceArray.from_numpy.__func__))
_lib.DeviceArray.numpy = timed("d2h", _lib.DeviceArray.numpy)
class L:
    def __getattr__(self, n):
        f = getattr(lib, n)
        return timed(n, f) if n in ("b200mc_create", "b200mc_init", "b200mc_sample", "b200mc_get_vector", "b200mc_get_scalars", "b200mc_set_vector", "b200mc_set_scalars") else f
_lib._lib = L()
gc.collect(); gc.disable()
rows = []
for i in range(43):
    cur.clear()
    t0 = time.perf_counter()
    k2 = NUTS(models.logistic_regression, step_size=step, inverse_mass_matrix=imm, adapt_step_size=False, adapt_mass_matrix=False)
    m2 = MCMC(k2, num_warmup=0, num_samples=100)
    m2.run(PRNGKey(1), X, y, init_params=z_last, extra_fields=("num_steps",))
    _ = m2.get_samples()["coefs"].mean()
    cur["total"] = time.perf_counter() - t0
    k2._destroy()
    if i >= 3: rows.append(dict(cur))
keys = sorted({k for r in rows for k in r})
for k in keys:
    v = np.array([r.get(k, 0.0) for r in rows]) * 1e3
    print(f"{k:22s} median {np.median(v):7.3f} ms  max {v.max():7.3f}  p90 {np.percentile(v, 90):7.3f}")
tot = np.array([r["total"] for r in rows]) * 1e3
print("steps over 18 ms:", [(i, round(t, 1), {k: round(rows[i].get(k, 0) * 1e3, 1) for k in keys if k != "total" and rows[i].get(k, 0) * 1e3 > 1.0}) for i, t in enumerate(tot) if t > 18][:6])
