"""ctypes loader for the C restatement (``oracle/potentials.c``).  TEST INFRASTRUCTURE."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(_HERE, "_build", "liboracle.so")


def build(force=False):
    src = os.path.join(_HERE, "potentials.c")
    if force or not os.path.exists(SO) or os.path.getmtime(SO) < os.path.getmtime(src):
        os.makedirs(os.path.dirname(SO), exist_ok=True)
        subprocess.check_call(["gcc", "-O3", "-march=native", "-fopenmp", "-shared", "-fPIC", src,
                               "-o", SO, "-lm"])
    return SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        _lib.oracle_logistic_potential_and_grad.restype = C.c_float
        _lib.oracle_logistic_potential_and_grad.argtypes = [C.c_void_p, C.c_void_p, C.c_long, C.c_int,
                                                            C.c_void_p, C.c_void_p, C.c_void_p]
        _lib.oracle_num_threads.restype = C.c_int
        _lib.oracle_set_num_threads.argtypes = [C.c_int]
    return _lib


def use_all_cores():
    """All host cores this process may run on, whatever OMP_NUM_THREADS says (torchrun sets it
    to 1).  Returns the thread count now in effect."""
    try:
        n = len(os.sched_getaffinity(0))
    except AttributeError:
        n = os.cpu_count() or 1
    lib().oracle_set_num_threads(n)
    return lib().oracle_num_threads()


class CLogistic:
    """Drop-in for ``oracle.potentials.Logistic`` backed by the C code (all host cores)."""

    name = "logistic"

    def __init__(self, X, y):
        self.X = np.ascontiguousarray(X, np.float32)
        self.y = np.ascontiguousarray(y).astype(np.uint8)
        self.dtype = np.dtype(np.float32)
        self.trace_sites = (("coefs", self.X.shape[1]),)
        self.dim = self.X.shape[1]
        self._scratch = np.empty(self.X.shape[0], np.float32)
        self.cores = lib().oracle_num_threads()

    flat_sites = property(lambda self: [("coefs", self.dim, 0)])

    def offsets(self):
        return {"coefs": (0, self.dim)}

    def potential_and_grad(self, z):
        z = np.ascontiguousarray(z, np.float32)
        g = np.empty(self.dim, np.float32)
        U = lib().oracle_logistic_potential_and_grad(
            self.X.ctypes.data, self.y.ctypes.data, self.X.shape[0], self.dim, z.ctypes.data,
            g.ctypes.data, self._scratch.ctypes.data)
        return np.float32(U), g

    def constrain(self, z):
        return {"coefs": np.asarray(z)}
