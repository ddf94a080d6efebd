#!/usr/bin/env python
"""Measured memory roofs for the streaming kernel's access pattern (b200mc_membench):
L2-resident buffer (the covtype-shape image minus pinned tiles, ping-pong sweep) and an
out-of-L2 buffer (forward sweeps), at the kernel's own tile size and ring depth."""
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def measure(tile_bytes=7200, stages=20, verbose=False):
    from numpyro_b200 import _lib

    lib = _lib.lib()
    out = {}
    g = C.c_float()
    for name, mb, passes, rev in (("l2_pingpong_125MB", 125, 40, 1), ("l2_pingpong_100MB", 100, 40, 1),
                                  ("l2_forward_64MB", 64, 60, 0), ("forward_125MB", 125, 40, 0),
                                  ("hbm_forward_1024MB", 1024, 6, 0), ("hbm_forward_257MB", 257, 20, 0)):
        _lib.check(lib.b200mc_membench(mb << 20, tile_bytes, stages, passes, rev, C.byref(g), None))
        out[name] = round(float(g.value), 1)
        if verbose:
            print(name, out[name], "GB/s", flush=True)
    return out


if __name__ == "__main__":
    import __graft_entry__ as ge

    ge.build()
    print(json.dumps(measure(verbose=True)))
