"""Loader for the REFERENCE's own Python operators (TEST INFRASTRUCTURE ONLY, like everything under oracle/).

The reference modules (python/RAHT_param.py, RAHT.py, iRAHT.py, voxelize_pc.py, utils.py, crosscheck.py) have the
same names as this repo's drop-in shims, so they are imported under private names from

    /root/reference/python          when the checkout is present (the build container), else
    oracle/_ref/python              the copy staged by `make -C oracle ref` (git-ignored; travels to the GPU box)

without leaving anything of theirs in sys.modules / sys.path.  Only tests/ and bench.py's --impl reference /
cpu_baseline legs may use this; the product package never does (tests/test_abi.py checks).

    ref = refpy.load()              # namespace with .RAHT_param, .RAHT, .iRAHT, .voxelize_pc, .utils, .crosscheck
    ref.RAHT_param.RAHT_param_reorder_fast(V, minV, width, depth)
"""
import importlib
import os
import sys
import types

_HERE = os.path.dirname(os.path.abspath(__file__))
_NAMES = ("utils", "RAHT_param", "RAHT", "iRAHT", "voxelize_pc", "crosscheck")
_cache = None


def directory():
    for d in ("/root/reference/python", os.path.join(_HERE, "_ref", "python")):
        if os.path.isfile(os.path.join(d, "RAHT_param.py")):
            return d
    return None


def available():
    return directory() is not None


def load():
    """Import the reference modules (once) and hand them back in a namespace; sys.modules / sys.path are restored."""
    global _cache
    if _cache is not None:
        return _cache
    d = directory()
    if d is None:
        raise ImportError("reference python not found (neither /root/reference/python nor oracle/_ref/python: run `make -C oracle ref` "
                          "in the build container)")
    saved = {n: sys.modules.pop(n, None) for n in _NAMES}
    sys.path.insert(0, d)
    ns = types.SimpleNamespace(directory=d)
    try:
        for n in _NAMES:
            m = importlib.import_module(n)
            if not os.path.abspath(m.__file__).startswith(os.path.abspath(d)):
                raise ImportError(f"{n} resolved to {m.__file__}, not to the reference under {d}")
            setattr(ns, n, m)
    finally:
        sys.path.remove(d)
        for n in _NAMES:
            sys.modules.pop(n, None)
            if saved[n] is not None:
                sys.modules[n] = saved[n]
    _cache = ns
    return ns
