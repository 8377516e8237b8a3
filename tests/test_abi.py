"""CPU tests: the C-ABI library loads and exports every symbol include/raht_b200.h declares
(no compute calls without a GPU), and the product package never touches oracle/."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "raht-3dgs-codec_b200")
LIB = os.path.join(PKG, "libraht_b200.so")
HDR = os.path.join(ROOT, "include", "raht_b200.h")


def declared_symbols():
    src = open(HDR).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(raht_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    assert os.path.exists(LIB), "build the library first: __graft_entry__.build()"
    lib = ctypes.CDLL(LIB)
    syms = declared_symbols()
    assert len(syms) >= 30
    missing = [s for s in syms if not hasattr(lib, s)]
    assert not missing, f"declared in raht_b200.h but not exported: {missing}"


def test_ctypes_signatures_cover_header():
    from raht_b200 import _lib
    assert sorted(_lib.SIGNATURES) == declared_symbols()
    import re
    header = open(os.path.join(ROOT, "include", "raht_b200.h")).read()
    assert _lib.lib.raht_abi_version() == int(re.search(r"#define RAHT_B200_ABI_VERSION (\d+)", header).group(1)) == 18


def test_argument_validation_without_gpu():
    """Host-side validation returns RAHT_E_INVALID and a message before any CUDA call."""
    from raht_b200 import _lib
    rc = _lib.lib.raht_plan_create(None, 10, 30, None, None, 0, ctypes.byref(ctypes.c_void_p()), None)   # tile_rows too small
    assert rc == -1
    assert b"tile_rows" in _lib.lib.raht_last_error()
    rc = _lib.lib.raht_rlgr_encode(None, 10, 10, 1, 1, None, 12, None, None, None, None, 0, None)        # slot not multiple of 8
    assert rc == -1


def test_product_package_does_not_use_the_oracle():
    bad = []
    for dirpath, _, files in os.walk(PKG):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f), errors="replace").read()
                if re.search(r"\boracle\b", txt) and "no CPU" not in txt:
                    bad.append(os.path.join(dirpath, f))
                if re.search(r"^\s*(from|import)\s+oracle", txt, flags=re.M):
                    bad.append(os.path.join(dirpath, f))
    assert not bad, bad


def test_shims_refuse_cpu_tensors():
    import torch
    import RAHT_param
    import RAHT
    with pytest.raises(RuntimeError):
        RAHT_param.RAHT_param_reorder_fast(torch.zeros(4, 3, dtype=torch.float64), torch.zeros(3), 4, 2)
    with pytest.raises(RuntimeError):
        RAHT.RAHT2_optimized(torch.zeros(4, 3, dtype=torch.float64), [], [], [])
