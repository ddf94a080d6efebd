"""Import the WHOLE reference package (``/root/reference/numpyro``) without JAX.

TEST INFRASTRUCTURE, a wider variant of ``refshim.py``: the same NumPy stand-in for the slice of
JAX the sampler uses, plus NumPy/SciPy versions of the special functions the distributions
need, plus a permissive fallback (any other ``jax.*`` name resolves to an inert stub so that
module-level imports and decorators succeed; CALLING such a stub raises).  With it the
reference's own model code runs eagerly: handlers, ``biject_to`` transforms, reparam,
distributions' ``log_prob`` -- so ``numpyro.infer.util.initialize_model`` / ``potential_energy``
/ ``constrain_fn`` of the four registered models can be evaluated by the reference itself
(values only: there is no autodiff here).  Never used on the GPU box.
"""
import importlib.abc
import importlib.machinery
import sys
import types

import numpy as np
import scipy.special as sps

import refshim
from refshim import _lift, _wrap


STRICT = False  # set once the package is imported: calling an un-emulated name then raises


class _Stub:
    """Inert placeholder: importable, usable as decorator / annotation / base class."""

    def __init__(self, name):
        object.__setattr__(self, "_n", name)
        object.__setattr__(self, "__name__", name.rsplit(".", 1)[-1])
        object.__setattr__(self, "__qualname__", name)
        object.__setattr__(self, "__doc__", "")
        object.__setattr__(self, "__module__", "jax")

    def __call__(self, *a, **k):
        if len(a) == 1 and callable(a[0]) and not k:
            return a[0]  # used as a decorator
        if not a and k:
            return lambda f: f  # decorator factory
        if not STRICT:
            return _Stub(self._n + "()")  # module-level call while the package is being imported
        raise NotImplementedError(f"refshim_full: {self._n} is not emulated")

    def __getattr__(self, k):
        if k.startswith("__") and k.endswith("__"):
            raise AttributeError(k)
        return _Stub(self._n + "." + k)

    def __getitem__(self, k):
        return self

    def __mro_entries__(self, bases):
        return (object,)

    def __or__(self, o):
        return self

    __ror__ = __or__

    def __repr__(self):
        return f"<stub {self._n}>"


class _Permissive(types.ModuleType):
    def __getattr__(self, k):
        if k.startswith("__") and k.endswith("__"):
            raise AttributeError(k)
        return _Stub(self.__name__ + "." + k)


class _Finder(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    def find_spec(self, name, path=None, target=None):
        if name == "jax" or name.startswith("jax."):
            return importlib.machinery.ModuleSpec(name, self, is_package=True)
        return None

    def create_module(self, spec):
        m = _Permissive(spec.name)
        m.__path__ = []
        return m

    def exec_module(self, module):
        pass


def _permissive_copy(mod):
    """Known attributes first, stubs for the rest."""
    m = _Permissive(mod.__name__)
    m.__dict__.update({k: v for k, v in mod.__dict__.items() if not k.startswith("__")})
    m.__path__ = []
    return m


def install_full(partitionable=True):
    refshim.PARTITIONABLE = bool(partitionable)
    if "numpyro" in sys.modules and getattr(sys.modules["numpyro"], "_refshim", False):
        raise RuntimeError("refshim.install() was used in this process; use a fresh one")
    if "jax" in sys.modules and not getattr(sys.modules["jax"], "_refshim", False):
        raise RuntimeError("a real jax is imported")
    if getattr(sys.modules.get("jax"), "_refshim_full", False):
        import numpyro

        return numpyro
    mods = {k: _permissive_copy(v) for k, v in refshim._build_jax().items()}
    # re-link the submodules on the copies
    for k, m in mods.items():
        if "." in k:
            parent, _, leaf = k.rpartition(".")
            setattr(mods[parent], leaf, m)
    jax = mods["jax"]
    jax._refshim = True
    jax._refshim_full = True
    jnp = mods["jax.numpy"]
    for name in ("broadcast_arrays", "broadcast_shapes", "result_type", "zeros_like", "ones_like",
                 "full", "full_like", "atleast_1d", "moveaxis", "squeeze", "ravel", "prod", "max",
                 "min", "cumprod", "log2", "exp2", "expm1", "less", "greater", "equal", "not_equal",
                 "logical_and", "logical_or", "logical_not", "isinf", "round", "clip", "sort",
                 "argmax", "argmin", "tile", "repeat", "flip", "diagonal", "trace", "vectorize",
                 "floor_divide", "remainder", "mod", "arctan", "arctan2", "sin", "cos", "sinh",
                 "cosh", "arcsinh", "log10", "nan_to_num", "amax", "amin", "count_nonzero",
                 "searchsorted", "interp", "median", "quantile", "std", "matrix_transpose",
                 "tensordot", "vdot", "kron", "split", "array_split", "hstack", "vstack", "column_stack",
                 "tril_indices", "triu_indices", "indices", "meshgrid", "isscalar", "iscomplexobj",
                 "negative", "subtract", "divide", "true_divide", "absolute", "sum", "cumsum"):
        if hasattr(np, name) and name not in ("result_type",):
            setattr(jnp, name, _lift(getattr(np, name)))
    f32 = lambda dt: np.float32 if dt is None else jnp.result_type(dt)  # noqa: E731
    jnp.empty = lambda shape, dtype=None: _wrap(np.zeros(shape, f32(dtype)))
    jnp.empty_like = lambda x, dtype=None: _wrap(np.zeros_like(np.asarray(x), dtype=dtype))
    jnp.full = lambda shape, v, dtype=None: _wrap(np.full(shape, v, dtype=None if dtype is None else jnp.result_type(dtype)))
    jnp.broadcast_shapes = np.broadcast_shapes
    _clip = _lift(np.clip)
    jnp.clip = lambda x, min=None, max=None, **k: _clip(x, k.get("a_min", min), k.get("a_max", max))
    jnp.shape = np.shape
    jnp.ndim = np.ndim
    jnp.newaxis = None
    jnp.e = np.e
    jnp.euler_gamma = np.euler_gamma
    jnp.bool_ = np.bool_
    jnp.generic = np.generic
    jnp.number = np.number
    jnp.inexact = np.inexact
    lax = mods["jax.lax"]

    # lax.map / vmap over the leading axis as a Python loop + stack: by definition the batched
    # result equals the per-element calls, which is all the fixtures need (Predictive /
    # log_likelihood batch over posterior draws through numpyro.util.soft_vmap)
    def _loop_map(fn, xs):
        leaves, treedef = refshim.tree_flatten(xs)
        n = np.shape(leaves[0])[0]
        outs = [fn(refshim.tree_unflatten(treedef, [np.asarray(leaf)[i] for leaf in leaves])) for i in range(n)]
        o_leaves = [refshim.tree_flatten(o)[0] for o in outs]
        o_def = refshim.tree_flatten(outs[0])[1]
        return refshim.tree_unflatten(o_def, [_wrap(np.stack([np.asarray(ol[j]) for ol in o_leaves]))
                                              for j in range(len(o_leaves[0]))])

    lax.map = _loop_map

    def _vmap(fn, in_axes=0, out_axes=0, **kw):
        if in_axes != 0 or out_axes != 0:
            raise NotImplementedError("refshim_full: vmap is emulated for in_axes=0, out_axes=0 only")
        return lambda *xs: _loop_map(lambda t: fn(*t), tuple(xs))

    jax.vmap = _vmap
    lax.broadcast_shapes = np.broadcast_shapes
    lax.lgamma = _lift(sps.gammaln)
    lax.digamma = _lift(sps.digamma)
    lax.erf = _lift(sps.erf)
    lax.erf_inv = _lift(sps.erfinv)
    lax.exp = _lift(np.exp)
    lax.log = _lift(np.log)
    lax.sqrt = _lift(np.sqrt)
    lax.expm1 = _lift(np.expm1)
    lax.log1p = _lift(np.log1p)
    lax.tanh = _lift(np.tanh)
    lax.abs = _lift(np.abs)
    lax.sign = _lift(np.sign)
    lax.pow = _lift(np.power)
    lax.clamp = lambda lo, x, hi: _lift(np.clip)(x, lo, hi)
    lax.stop_gradient = lambda x: x
    lax.dynamic_slice_in_dim = _Stub("jax.lax.dynamic_slice_in_dim")
    special = mods["jax.scipy.special"]
    for name in ("gammaln", "digamma", "betaln", "erf", "erfc", "erfinv", "logsumexp", "xlogy",
                 "xlog1py", "logit", "ndtr", "ndtri", "log_ndtr", "gammainc", "gammaincc", "betainc",
                 "i0e", "i1e", "polygamma", "entr", "multigammaln", "zeta"):
        if hasattr(sps, name):
            setattr(special, name, _lift(getattr(sps, name)))
    nn = _Permissive("jax.nn")
    nn.__path__ = []
    nn.softplus = _lift(lambda x: np.logaddexp(x, 0.0))
    nn.sigmoid = special.expit
    nn.log_sigmoid = _lift(lambda x: -np.logaddexp(-x, 0.0))
    nn.relu = _lift(lambda x: np.maximum(x, 0))
    nn.softmax = _lift(lambda x, axis=-1: sps.softmax(x, axis=axis))
    nn.log_softmax = _lift(lambda x, axis=-1: sps.log_softmax(x, axis=axis))
    nn.one_hot = _Stub("jax.nn.one_hot")
    jax.nn = nn
    mods["jax.nn"] = nn
    tu = mods["jax.tree_util"]
    tu.register_pytree_node = lambda *a, **k: None
    tu.register_pytree_node_class = lambda cls: cls
    tu.tree_leaves = lambda t: refshim.tree_flatten(t)[0]
    tu.tree_structure = lambda t: refshim.tree_flatten(t)[1]
    tu.tree_all = lambda t: all(refshim.tree_flatten(t)[0])
    tu.tree_reduce = lambda f, t, *i: __import__("functools").reduce(f, refshim.tree_flatten(t)[0], *i)
    tu.Partial = __import__("functools").partial
    jax.tree.leaves = tu.tree_leaves
    jax.tree.structure = tu.tree_structure
    jax.tree.all = tu.tree_all
    jax.tree.reduce = tu.tree_reduce
    class custom_jvp:  # the primal function, with inert derivative registration
        def __init__(self, fun, nondiff_argnums=()):
            self.fun = fun
            self.__name__ = getattr(fun, "__name__", "fn")

        def __call__(self, *a, **k):
            return self.fun(*a, **k)

        def defjvp(self, f, *a, **k):
            return f

        def defjvps(self, *fs):
            return None

        def defvjp(self, *fs, **k):
            return None

    jax.custom_jvp = custom_jvp
    jax.custom_vjp = custom_jvp
    jax.eval_shape = _Stub("jax.eval_shape")
    jax.local_device_count = lambda: 1
    jax.device_count = lambda: 1
    jax.devices = lambda *a: [types.SimpleNamespace(platform="cpu", id=0)]
    jax.default_backend = lambda: "cpu"
    jax.__version__ = "0.6.0"
    jax.__version_info__ = (0, 6, 0)
    sys.modules.update(mods)
    sys.meta_path.insert(0, _Finder())
    md = types.ModuleType("multipledispatch")

    def dispatch(*types_, **kw):
        def deco(f):
            return f
        return deco

    md.dispatch = dispatch
    sys.modules["multipledispatch"] = md
    if refshim.REFERENCE not in sys.path:
        sys.path.insert(0, refshim.REFERENCE)
    global STRICT
    import numpyro

    STRICT = True
    return numpyro
