"""Run the reference's OWN ``numpyro/infer/hmc.py`` and ``hmc_util.py`` without JAX.

TEST INFRASTRUCTURE.  ``jax`` / ``jaxlib`` are not installable in this image, so the reference
cannot be imported as a package.  Its sampler, however, is plain Python over a small slice of
the ``jax`` API.  ``install()`` puts a NumPy-backed stand-in for that slice into
``sys.modules`` (32-bit default dtypes like JAX with x64 off, ``.at[].set``, pytrees,
``ravel_pytree``, eager ``lax`` control flow) and then executes the UNMODIFIED source files

    /root/reference/numpyro/util.py
    /root/reference/numpyro/infer/hmc_util.py
    /root/reference/numpyro/infer/hmc.py
    /root/reference/numpyro/diagnostics.py        (``load_diagnostics()``)

as ``numpyro.util`` / ``numpyro.infer.hmc_util`` / ``numpyro.infer.hmc``.  What is NOT the
reference's code in such a run:

* ``jax.random`` -> ``oracle/threefry.py`` (third-party Threefry-2x32, restated);
* ``jax.value_and_grad(f)`` -> ``f.value_and_grad`` (closed-form gradients of the registered
  families from ``oracle/potentials.py``; JAX's autodiff is third-party);
* ``numpyro.infer.util.initialize_model`` -> a stub returning the family's potential (model
  tracing needs the whole distributions/handlers stack); callers pass ``init_params``.

Everything else -- ``velocity_verlet``, ``build_tree`` and the whole iterative NUTS machinery,
``warmup_adapter``, ``dual_averaging``, ``welford_covariance``, ``_initialize_mass_matrix``
(block-structured mass matrices included), ``momentum_generator``, ``sample_kernel``,
``HMC``/``NUTS`` classes -- is the reference's code, line for line.  Used by
``make_refpin.py`` to write ``refpin.npz``; never on the GPU box (no /root/reference there).
"""
from collections import OrderedDict, namedtuple
import importlib.util
import os
import sys
import types

import numpy as np

REFERENCE = os.environ.get("NUMPYRO_REFERENCE", "/root/reference")
PARTITIONABLE = True  # jax_threefry_partitionable (default True from jax 0.5.0)

_DEMOTE = {np.dtype(np.float64): np.float32, np.dtype(np.int64): np.int32,
           np.dtype(np.uint64): np.uint32, np.dtype(np.complex128): np.complex64}


class _At:
    def __init__(self, arr):
        self.arr = arr

    def __getitem__(self, idx):
        return _AtIdx(self.arr, idx)


class _AtIdx:
    def __init__(self, arr, idx):
        self.arr, self.idx = arr, idx

    def set(self, v):
        out = np.array(self.arr, copy=True)
        out[self.idx] = v
        return _wrap(out)

    def add(self, v):
        out = np.array(self.arr, copy=True)
        out[self.idx] += v
        return _wrap(out)


def _is_np(x):
    return isinstance(x, (np.ndarray, np.generic))


def _demote_in(x):
    """Argument -> plain ndarray/scalar with JAX's x64-off dtypes (no 64-bit types exist)."""
    if isinstance(x, Arr):
        x = np.asarray(x)
    if _is_np(x):
        t = _DEMOTE.get(x.dtype)
        return x.astype(t) if t is not None else x
    if isinstance(x, (list, tuple)) and any(_is_np(e) for e in x):
        return type(x)(_demote_in(e) for e in x)
    return x


def _ints_to_f32(args):
    return [a.astype(np.float32) if _is_np(a) and a.dtype.kind in "iu" else a for a in args]


def _call(fn, args, kwargs):
    """Evaluate a NumPy function the way JAX would type it: integer arrays meeting floats
    (f32 arrays or weak Python floats) are converted to float32 BEFORE the arithmetic
    (NumPy would compute int32 (+) float32 in float64), and an all-integer call whose result
    is floating (true_divide, sqrt, exp, ...) runs in float32."""
    args = [_demote_in(a) for a in args]
    kwargs = {k: _demote_in(v) for k, v in kwargs.items()}
    has_float = any((_is_np(a) and a.dtype.kind == "f") or isinstance(a, float) for a in args)
    has_int = any(_is_np(a) and a.dtype.kind in "iu" for a in args)
    if has_float and has_int and fn is not np.where:
        args = _ints_to_f32(args)
    with np.errstate(all="ignore"):
        out = fn(*args, **kwargs)
        probe = out[0] if isinstance(out, tuple) else out
        if _is_np(probe) and probe.dtype == np.float64 and has_int and not has_float:
            out = fn(*_ints_to_f32(args), **kwargs)
    if isinstance(out, tuple):
        return tuple(_wrap(o) for o in out)
    return _wrap(out)


class Arr(np.ndarray):
    """ndarray with JAX's ``.at`` and JAX's (x64 off) result types for every ufunc."""

    @property
    def at(self):
        return _At(self)

    def __array_ufunc__(self, ufunc, method, *inputs, out=None, **kwargs):
        return _call(getattr(ufunc, method), inputs, kwargs)

    def __bool__(self):
        return bool(np.asarray(self))


def _wrap(x):
    """NumPy value -> what JAX would hand back with x64 off (a 32-bit array)."""
    if _is_np(x):
        a = np.asarray(x)
        t = _DEMOTE.get(a.dtype)
        if t is not None:
            a = a.astype(t)
        return a.view(Arr)
    return x


def _unwrap(x):
    return _demote_in(x)


def _lift(fn):
    def wrapped(*args, **kwargs):
        out = _call(fn, args, kwargs)
        if isinstance(out, (float, int)) and not isinstance(out, bool):
            out = _wrap(np.asarray(out))
        return out
    wrapped.__name__ = getattr(fn, "__name__", "fn")
    return wrapped


# ---------------------------------------------------------------- pytrees (jax.tree_util)
def _is_namedtuple(x):
    return isinstance(x, tuple) and hasattr(x, "_fields")


def tree_flatten(tree):
    leaves = []

    def rec(t):
        if t is None:
            return ("none",)
        if isinstance(t, OrderedDict):
            return ("odict", list(t.keys()), [rec(v) for v in t.values()])
        if isinstance(t, dict):
            ks = sorted(t.keys())
            return ("dict", ks, [rec(t[k]) for k in ks])
        if _is_namedtuple(t):
            return ("ntuple", type(t), [rec(v) for v in t])
        if isinstance(t, tuple):
            return ("tuple", None, [rec(v) for v in t])
        if isinstance(t, list):
            return ("list", None, [rec(v) for v in t])
        leaves.append(t)
        return ("leaf",)

    return leaves, rec(tree)


def tree_unflatten(treedef, leaves):
    it = iter(leaves)

    def rec(d):
        kind = d[0]
        if kind == "none":
            return None
        if kind == "leaf":
            return next(it)
        kids = [rec(c) for c in d[2]]
        if kind == "odict":
            return OrderedDict(zip(d[1], kids))
        if kind == "dict":
            return dict(zip(d[1], kids))
        if kind == "ntuple":
            return d[1](*kids)
        if kind == "tuple":
            return tuple(kids)
        return kids

    return rec(treedef)


def tree_map(f, tree, *rest):
    leaves, treedef = tree_flatten(tree)
    others = [tree_flatten(r)[0] for r in rest]
    return tree_unflatten(treedef, [f(*xs) for xs in zip(leaves, *others)])


def ravel_pytree(tree):
    leaves, treedef = tree_flatten(tree)
    arrs = [np.asarray(l) for l in leaves]
    shapes = [a.shape for a in arrs]
    sizes = [a.size for a in arrs]
    dtype = np.result_type(*arrs) if arrs else np.dtype(np.float32)
    dtype = _DEMOTE.get(np.dtype(dtype), dtype)
    flat = (np.concatenate([a.ravel().astype(dtype) for a in arrs]) if arrs
            else np.zeros((0,), np.float32))

    def unravel(v):
        v = np.asarray(v)
        out, off = [], 0
        for shp, sz in zip(shapes, sizes):
            out.append(_wrap(v[off:off + sz].reshape(shp)))
            off += sz
        return tree_unflatten(treedef, out)

    return _wrap(flat), unravel


# ---------------------------------------------------------------- jax.numpy
def _make_jnp():
    jnp = types.ModuleType("jax.numpy")

    def _default(dtype, kind=float):
        if dtype is None:
            return np.float32 if kind is float else np.int32
        return result_type(dtype)

    def result_type(*args):
        args = [np.asarray(a) if isinstance(a, Arr) else a for a in args]
        if len(args) == 1 and args[0] is float:
            return np.dtype(np.float32)
        if len(args) == 1 and args[0] is int:
            return np.dtype(np.int32)
        if len(args) == 1 and args[0] is bool:
            return np.dtype(np.bool_)
        dt = np.result_type(*args)
        return np.dtype(_DEMOTE.get(np.dtype(dt), dt))

    jnp.result_type = result_type
    jnp.zeros = lambda shape, dtype=None: _wrap(np.zeros(shape, _default(dtype)))
    jnp.ones = lambda shape, dtype=None: _wrap(np.ones(shape, _default(dtype)))
    jnp.identity = lambda n, dtype=None: _wrap(np.identity(n, _default(dtype)))
    jnp.eye = lambda n, dtype=None: _wrap(np.eye(n, dtype=_default(dtype)))

    def array(x, dtype=None):
        a = np.array(_unwrap(x), dtype=None if dtype is None else result_type(dtype))
        return _wrap(a)

    jnp.array = array
    jnp.asarray = array
    for name in ("where", "log", "exp", "sqrt", "matmul", "multiply", "sum", "dot", "clip",
                 "swapaxes", "reciprocal", "mean", "var", "stack", "isnan", "diag", "cov", "outer",
                 "logaddexp", "ceil", "add", "einsum", "abs", "maximum", "minimum", "log1p",
                 "expand_dims", "concatenate", "broadcast_to", "all", "any", "isfinite", "square",
                 "arange", "take", "reshape", "cumsum", "floor", "power", "sign", "tanh",
                 "argsort", "linspace", "triu", "tril"):
        setattr(jnp, name, _lift(getattr(np, name)))
    jnp.bitwise_count = _lift(np.bitwise_count)
    jnp.ndim = lambda x: np.ndim(x)
    jnp.size = lambda x, axis=None: int(np.size(x, axis))
    jnp.shape = lambda x: np.shape(x)
    jnp.inf = np.inf
    jnp.nan = np.nan
    jnp.pi = np.pi
    jnp.finfo = np.finfo
    jnp.iinfo = np.iinfo
    jnp.float32, jnp.float64, jnp.int32, jnp.int64 = np.float32, np.float64, np.int32, np.int64
    jnp.uint32, jnp.bool_ = np.uint32, np.bool_
    jnp.ndarray = np.ndarray
    jnp.dtype = np.dtype
    jnp.issubdtype = np.issubdtype
    jnp.floating, jnp.integer = np.floating, np.integer
    linalg = types.ModuleType("jax.numpy.linalg")
    linalg.cholesky = _lift(np.linalg.cholesky)
    linalg.inv = _lift(np.linalg.inv)
    jnp.linalg = linalg
    return jnp


# ---------------------------------------------------------------- jax.random
def _make_random():
    from oracle import threefry as tf

    rnd = types.ModuleType("jax.random")
    rnd.PRNGKey = lambda seed: _wrap(tf.PRNGKey(seed))
    rnd.key = rnd.PRNGKey
    rnd.split = lambda key, num=2: _wrap(tf.split(np.asarray(key), num, partitionable=PARTITIONABLE))

    def _dtype(dtype):
        return np.float32 if dtype in (None, float) else np.dtype(dtype).type

    def normal(key, shape=(), dtype=None):
        assert _dtype(dtype) is np.float32
        return _wrap(tf.normal(np.asarray(key), tuple(shape), partitionable=PARTITIONABLE))

    def uniform(key, shape=(), dtype=None, minval=0.0, maxval=1.0):
        assert _dtype(dtype) is np.float32
        return _wrap(tf.uniform(np.asarray(key), tuple(shape), minval, maxval,
                                partitionable=PARTITIONABLE))

    def bernoulli(key, p=np.float32(0.5), shape=None):
        p = np.asarray(p)
        if p.dtype.kind != "f":
            p = p.astype(np.float32)
        shape = p.shape if shape is None else tuple(shape)
        return _wrap(np.asarray(tf.bernoulli(np.asarray(key), p, shape, partitionable=PARTITIONABLE)))

    rnd.normal, rnd.uniform, rnd.bernoulli = normal, uniform, bernoulli
    return rnd


# ---------------------------------------------------------------- jax.lax / transforms
def _make_lax():
    lax = types.ModuleType("jax.lax")

    def while_loop(cond_fun, body_fun, init_val):
        val = init_val
        while cond_fun(val):
            val = body_fun(val)
        return val

    def fori_loop(lower, upper, body_fun, init_val):
        val = init_val
        for i in range(int(lower), int(upper)):
            val = body_fun(i, val)
        return val

    def cond(pred, *args):
        if len(args) == 4:  # old signature: (true_operand, true_fun, false_operand, false_fun)
            t_op, t_fn, f_op, f_fn = args
            return t_fn(t_op) if pred else f_fn(f_op)
        t_fn, f_fn, *ops = args
        return t_fn(*ops) if pred else f_fn(*ops)

    lax.while_loop, lax.fori_loop, lax.cond = while_loop, fori_loop, cond
    lax.convert_element_type = lambda x, dt: _wrap(np.asarray(x).astype(_make_jnp().result_type(dt)))
    lax.stop_gradient = lambda x: x
    return lax


def _value_and_grad(f, argnums=0, has_aux=False):
    """AD is third-party: closed-form gradients travel with the potential as
    ``f.value_and_grad`` (see ``Potential`` below)."""
    if not hasattr(f, "value_and_grad"):
        raise NotImplementedError("refshim: no autodiff; potential needs .value_and_grad")
    return f.value_and_grad


def _no(name):
    def fail(*a, **k):
        raise NotImplementedError(f"refshim: jax.{name} is not emulated")
    return fail


def _build_jax():
    jax = types.ModuleType("jax")
    jax.__path__ = []  # package
    jnp, rnd, lax = _make_jnp(), _make_random(), _make_lax()
    jax.numpy, jax.random, jax.lax = jnp, rnd, lax
    jax.value_and_grad = _value_and_grad
    jax.grad = _no("grad")
    jax.jacfwd = _no("jacfwd")
    jax.vmap = _no("vmap")
    jax.pmap = _no("pmap")
    jax.jit = lambda fn, *a, **k: fn
    jax.device_get = lambda x: x
    jax.device_put = lambda x, *a, **k: x
    jax.Array = np.ndarray
    tree = types.ModuleType("jax.tree")
    tree.map, tree.flatten, tree.unflatten = tree_map, tree_flatten, tree_unflatten
    tree.leaves = lambda t: tree_flatten(t)[0]
    jax.tree = tree
    tu = types.ModuleType("jax.tree_util")
    tu.tree_map, tu.tree_flatten, tu.tree_unflatten = tree_map, tree_flatten, tree_unflatten
    tu.register_pytree_node = lambda *a, **k: None
    jax.tree_util = tu
    jax.tree_map, jax.tree_flatten = tree_map, tree_flatten
    fu = types.ModuleType("jax.flatten_util")
    fu.ravel_pytree = ravel_pytree
    jax.flatten_util = fu
    core = types.ModuleType("jax.core")
    core.Tracer = type("Tracer", (), {})
    jax.core = core
    exp = types.ModuleType("jax.experimental")
    exp.io_callback = _no("experimental.io_callback")
    jax.experimental = exp
    dtypes = types.ModuleType("jax.dtypes")
    dtypes.prng_key = type("prng_key", (), {})
    dtypes.issubdtype = lambda a, b: False
    jax.dtypes = dtypes
    cfg = types.SimpleNamespace(update=lambda *a, **k: None, jax_enable_x64=False)
    jax.config = cfg
    scipy = types.ModuleType("jax.scipy")
    scipy.__path__ = []
    sl = types.ModuleType("jax.scipy.linalg")
    import scipy.linalg as _sl

    def solve_triangular(a, b, lower=False, trans=0):
        a, b = np.asarray(a), np.asarray(b)
        return _wrap(_sl.solve_triangular(a, b, lower=lower, trans=trans).astype(a.dtype))

    sl.solve_triangular = solve_triangular
    sp = types.ModuleType("jax.scipy.special")

    def expit(x):
        x = np.asarray(x)
        with np.errstate(all="ignore"):
            return _wrap((1 / (1 + np.exp(-x))).astype(x.dtype if x.dtype.kind == "f" else np.float32))

    sp.expit = expit
    scipy.linalg, scipy.special = sl, sp
    jax.scipy = scipy
    mods = {"jax": jax, "jax.numpy": jnp, "jax.numpy.linalg": jnp.linalg, "jax.random": rnd,
            "jax.lax": lax, "jax.tree": tree, "jax.tree_util": tu, "jax.flatten_util": fu,
            "jax.core": core, "jax.experimental": exp, "jax.dtypes": dtypes, "jax.scipy": scipy,
            "jax.scipy.linalg": sl, "jax.scipy.special": sp}
    return mods


ParamInfo = namedtuple("ParamInfo", ["z", "potential_energy", "z_grad"])  # infer/util.py:52


class Potential:
    """``potential_fn`` over the site dict of a registered family (closed-form gradient)."""

    def __init__(self, family):
        self.family = family
        self.num_evals = 0

    def _flat(self, z):
        return np.asarray(ravel_pytree(z)[0], self.family.dtype)

    def __call__(self, z):
        return self.value_and_grad(z)[0]

    def value_and_grad(self, z):
        flat, unravel = ravel_pytree(z)
        self.num_evals += 1
        with np.errstate(all="ignore"):
            pe, g = self.family.potential_and_grad(np.asarray(flat, self.family.dtype))
        return _wrap(np.asarray(pe, self.family.dtype)), unravel(np.asarray(g, self.family.dtype))


class FamilyModel:
    """Stands where a NumPyro model callable would: ``initialize_model`` (stubbed below)
    resolves it to the family's potential."""

    def __init__(self, family):
        self.family = family
        self.potential = Potential(family)


def _load(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def install(partitionable=True):
    """Returns ``(hmc_module, hmc_util_module, numpyro_util_module)`` of the reference,
    executed over the NumPy stand-in."""
    global PARTITIONABLE
    PARTITIONABLE = bool(partitionable)
    if "numpyro.infer.hmc" in sys.modules and getattr(sys.modules["numpyro"], "_refshim", False):
        m = sys.modules
        return m["numpyro.infer.hmc"], m["numpyro.infer.hmc_util"], m["numpyro.util"]
    if "jax" in sys.modules and not getattr(sys.modules["jax"], "_refshim", False):
        raise RuntimeError("a real jax is imported; run the reference itself instead of the shim")
    root = os.path.join(REFERENCE, "numpyro")
    if not os.path.isdir(root):
        raise FileNotFoundError(root)
    mods = _build_jax()
    mods["jax"]._refshim = True
    sys.modules.update(mods)
    pkg = types.ModuleType("numpyro")
    pkg.__path__ = []
    pkg._refshim = True
    dist = types.ModuleType("numpyro.distributions")  # only used by consensus helpers
    infer = types.ModuleType("numpyro.infer")
    infer.__path__ = []
    mcmc = types.ModuleType("numpyro.infer.mcmc")
    mcmc.MCMCKernel = type("MCMCKernel", (), {})  # abstract base, mcmc.py:32-158
    iutil = types.ModuleType("numpyro.infer.util")
    iutil.ParamInfo = ParamInfo

    def initialize_model(rng_key, model, dynamic_args=False, init_strategy=None, model_args=(),
                         model_kwargs=None, forward_mode_differentiation=False, **kw):
        assert isinstance(model, FamilyModel)
        fam = model.family
        z0 = {k: _wrap(np.zeros(s, fam.dtype)) for k, s in fam.trace_sites}
        pot = model.potential
        potential_fn_gen = (lambda *a, **k: pot) if dynamic_args else pot
        info = ParamInfo(z0, None, None)
        return (info, potential_fn_gen, (lambda *a, **k: (lambda z: z)), None)

    iutil.initialize_model = initialize_model
    iutil.init_to_uniform = lambda *a, **k: None
    iutil.find_stack_level = lambda: 2
    sys.modules.update({"numpyro": pkg, "numpyro.distributions": dist, "numpyro.infer": infer,
                        "numpyro.infer.mcmc": mcmc, "numpyro.infer.util": iutil})
    pkg.distributions, pkg.infer = dist, infer
    infer.mcmc, infer.util = mcmc, iutil
    util = _load("numpyro.util", os.path.join(root, "util.py"))
    pkg.util = util
    hmc_util = _load("numpyro.infer.hmc_util", os.path.join(root, "infer", "hmc_util.py"))
    infer.hmc_util = hmc_util
    hmc = _load("numpyro.infer.hmc", os.path.join(root, "infer", "hmc.py"))
    infer.hmc = hmc
    return hmc, hmc_util, util


def load_diagnostics():
    """The reference's ``numpyro/diagnostics.py`` (NumPy arithmetic; it only touches
    ``jax.device_get`` and ``jax.tree``), executed unmodified."""
    install(PARTITIONABLE)
    if "numpyro.diagnostics" not in sys.modules:
        _load("numpyro.diagnostics", os.path.join(REFERENCE, "numpyro", "diagnostics.py"))
    return sys.modules["numpyro.diagnostics"]
