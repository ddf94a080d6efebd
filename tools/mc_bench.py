#!/usr/bin/env python
"""Quick device-timed throughput of C logistic chains at covtype shape (dev aid)."""
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    chains = int(sys.argv[1]) if len(sys.argv) > 1 else 4
    import __graft_entry__ as g

    g.build()
    from bench import make_covtype_shape, N_COLS
    from numpyro_b200 import _lib, models
    from numpyro_b200.infer import MCMC, NUTS
    from numpyro_b200.random import PRNGKey

    lib = _lib.lib()
    X, y = make_covtype_shape()
    kern = NUTS(models.logistic_regression)
    mc = MCMC(kern, num_warmup=300, num_samples=100, num_chains=chains, chain_method="vectorized")
    mc.warmup(PRNGKey(1), X, y)
    T_ = 100
    ns = _lib.DeviceArray((T_, chains), np.int32)
    col = _lib.Collect(n_keep=T_, num_steps_dev=ns.ptr)
    ev0, ev1 = lib.b200mc_event_create(), lib.b200mc_event_create()
    prof = (C.c_uint64 * 16)()
    out = []
    for i in range(6):
        if i == 2:
            _lib.check(lib.b200mc_get_profile(kern._handle, prof, 1))
        _lib.check(lib.b200mc_event_record(ev0, None))
        _lib.check(lib.b200mc_sample(kern._handle, T_, 0, 1, C.byref(col), None))
        _lib.check(lib.b200mc_event_record(ev1, None))
        ms = C.c_float()
        _lib.check(lib.b200mc_event_elapsed_ms(ev0, ev1, C.byref(ms)))
        n = ns.numpy()
        out.append((float(ms.value), int(n.sum()), n.sum(0).tolist()))
    _lib.check(lib.b200mc_get_profile(kern._handle, prof, 0))
    lay = (C.c_int * 8)()
    _lib.check(lib.b200mc_glm_layout(kern._handle, lay))
    passes = max(1, int(prof[4]))
    res = {"chains": chains, "layout": list(lay),
           "steps": [{"ms": round(m, 3), "grads": gsum, "kgrads_per_s": round(gsum / m, 1), "per_chain": pc}
                     for m, gsum, pc in out],
           "cycles_per_pass": {k: round(float(prof[i]) / passes, 1) for i, k in
                               enumerate(("stream", "arrival_wait", "grid_reduce", "sampler_step"))},
           "passes": passes}
    print(json.dumps(res))


if __name__ == "__main__":
    main()
