"""CPU tests of the drop-in boundary: the shared library loads, exports every symbol the
header declares, the ctypes binding covers them all, and argument validation fails loudly
without touching a GPU."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "ai4e_b200.h")


def _declared():
    src = open(HEADER).read()
    return re.findall(r"AI4E_API\s+(?:const\s+char\*|int)\s+(ai4e_\w+)\s*\(", src)


def test_header_declares_api():
    names = _declared()
    assert len(names) >= 20 and len(set(names)) == len(names)
    for must in ("ai4e_preprocess", "ai4e_conv_layer", "ai4e_engine_submit", "ai4e_engine_wait", "ai4e_pool_create"):
        assert must in names


def test_library_exports_every_declared_symbol():
    from ai4e_b200 import _lib as L
    lib = L.load()  # raises if the .so is missing or lacks a bound symbol
    for name in _declared():
        assert hasattr(lib, name), f"{name} declared in the header but not exported"
    assert set(L.SIGNATURES) == set(_declared()), "ctypes binding and header disagree"
    assert lib.ai4e_version() >= 100


def test_struct_layouts_match_header():
    from ai4e_b200 import _lib as L
    assert C.sizeof(L.LayerDesc) == 12 * 4 + 2 * 8
    assert C.sizeof(L.ModelDesc) == 8 * 4 + 8 * 4 + 8
    assert C.sizeof(L.EngineStats) == 8 * 13 + 8 + 3 * 8 + 2 * 4


def test_argument_validation_needs_no_gpu():
    from ai4e_b200 import _lib as L
    lib = L.load()
    mean = (C.c_float * 3)(0.5, 0.5, 0.5)
    std = (C.c_float * 3)(0.2, 0.2, 0.2)
    rc = lib.ai4e_preprocess(None, 1, 8, 8, 3, 0, 0, 0, 0, 8, 8, mean, std, 0, 3, None, None)
    assert rc == L.ERR_INVALID and b"null" in lib.ai4e_last_error()
    rc = lib.ai4e_conv_layer(None, 1, 8, 8, 64, None, None, 64, 1, 1, 1, 0, None, 1, None, 0, None)
    assert rc == L.ERR_INVALID
    rc = lib.ai4e_maxpool(None, 1, 8, 8, 64, 3, 2, 1, None, None)
    assert rc == L.ERR_INVALID
    out = C.c_void_p()
    rc = lib.ai4e_engine_create(0, None, None, 0, 1, 0, C.byref(out))
    assert rc == L.ERR_INVALID
    t = C.c_uint64()
    assert lib.ai4e_engine_submit(None, None, 1, 1, 3, 0, 0, C.byref(t)) == L.ERR_INVALID


def test_no_cpu_fallback():
    """on a machine without a CUDA device the engine refuses to start (it never computes on the CPU)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("this assertion is for GPU-less machines")
    from ai4e_b200 import Ai4eError
    from ai4e_b200.engine import Engine
    from tests import util
    with pytest.raises(Ai4eError) as ei:
        Engine(util.program_for("resnet18"), device=0, max_batch=2)
    assert ei.value.code in (-2, -1)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "aiforearth-api-platform_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                src = open(os.path.join(dirpath, f), encoding="utf-8", errors="replace").read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, re.M), f"{f} imports the oracle"
