"""The C-ABI library loads and exports every symbol include/rt3.h declares; record sizes match the numpy / ctypes mirrors.
No compute calls: this file runs without a GPU."""
import ctypes as C
import re
import subprocess

import numpy as np
import pytest

from raytracer3_b200 import _lib, types


def test_exports_match_header(built):
    L = _lib.load()
    declared = _lib.declared_symbols()
    assert len(declared) >= 30
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.SO_PATH], capture_output=True, text=True, check=True).stdout
    exported = sorted(set(re.findall(r" T (rt3_[a-z0-9_]+)", out)))
    assert exported == declared, (set(declared) ^ set(exported))
    for name in declared:
        assert hasattr(L, name)
        assert name in L._rt3_signatures, f"{name} has no ctypes signature"


def test_record_sizes(built):
    L = _lib.load()
    assert L.rt3_abi_version() == 1
    assert L.rt3_sizeof(0) == types.PARAMS.itemsize == 304          # GConst, src/renderer/mod.rs:47-63
    assert L.rt3_sizeof(1) == types.VERTEX.itemsize == 32           # Vertex, src/assets/mod.rs:127-133
    assert L.rt3_sizeof(2) == types.GEOMETRY.itemsize == 112
    assert L.rt3_sizeof(3) == types.RAY.itemsize == 32
    assert L.rt3_sizeof(4) == types.HIT.itemsize == 16
    assert L.rt3_sizeof(5) == C.sizeof(types.BuildStats)
    assert L.rt3_sizeof(6) == C.sizeof(types.Counters)


def test_no_cpu_fallback(built):
    """Without a CUDA device the library refuses to create a context (and says why); with one it succeeds."""
    import torch
    L = _lib.load()
    h = C.c_void_p()
    st = L.rt3_create(0, C.byref(h))
    if torch.cuda.is_available():
        assert st == 0 and h.value
        L.rt3_destroy(h)
    else:
        assert st == _lib.ERR_NO_DEVICE and not h.value
        assert b"no CUDA device" in L.rt3_last_error(None) and b"no CPU fallback" in L.rt3_last_error(None)
        from raytracer3_b200.host import RenderContext
        with pytest.raises(_lib.Rt3Error):
            RenderContext(0)
    assert L.rt3_create(0, None) == _lib.ERR_INVALID


def test_product_never_touches_the_oracle():
    """The shipped package and the CUDA sources must not import, link or call anything under oracle/."""
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for dirpath, _, files in os.walk(os.path.join(root, "raytracer3_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle_py" not in src and "liboracle" not in src and "orc_" not in src, f
    deps = subprocess.run(["ldd", _lib.SO_PATH], capture_output=True, text=True).stdout
    assert "oracle" not in deps and "emu" not in deps
