"""One adaptation launch + one 50-transition sampling launch of the headline kernel, printing
the gradient evaluations of each, for `ncu --set full -k regex:glm_kernel` (run from the repo root)."""
import sys
import numpy as np
sys.path.insert(0, ".")
import __graft_entry__ as g
g.build()
import bench
from numpyro_b200 import models
from numpyro_b200.infer import MCMC, NUTS
from numpyro_b200.random import PRNGKey

X, y = bench.make_covtype_shape()
k = NUTS(models.logistic_regression)
m = MCMC(k, num_warmup=100, num_samples=50)
m.warmup(PRNGKey(1), X, y, extra_fields=("num_steps",), collect_warmup=True)
print("warmup launch grad evals", int(m.get_extra_fields()["num_steps"].sum()))
m.run(PRNGKey(1), X, y, extra_fields=("num_steps",))
print("sampling launch grad evals", int(m.get_extra_fields()["num_steps"].sum()), "total incl. init", k.grad_evals())
