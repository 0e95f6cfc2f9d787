"""CPU checks of the drop-in boundary: the C-ABI library loads, exports every symbol that
include/glia_b200.h declares, and fails loudly (no CPU fallback) without a usable GPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from glia_b200 import api

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "glia_b200.h")).read()
    return sorted(set(re.findall(r"GLIA_API[^;(]*?\b(glia_b200_\w+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    if not os.path.exists(api.LIB_PATH):
        api.build_library()
    names = _declared()
    assert len(names) >= 20
    L = C.CDLL(api.LIB_PATH)
    for n in names:
        assert hasattr(L, n), n
    assert set(api.EXPORTS) <= set(names)


def test_no_torch_types_in_header():
    src = open(os.path.join(ROOT, "include", "glia_b200.h")).read()
    assert "torch" not in src.lower() and "at::" not in src


def test_no_cpu_fallback_without_gpu():
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        pytest.skip("a GPU is present")
    pb = np.zeros((8, 8), np.float32)
    with pytest.raises(api.GliaError) as ei:
        api.MergeTreeBuilder((8, 8), api.default_config(pb))
    assert ei.value.code == 2  # GLIA_E_CUDA
    f = dict(treemap=np.array([[[2, 3], [0, 0], [0, 0]]], np.int32), nodestatus=np.array([[1, -1, -1]], np.int32),
             xbestsplit=np.array([[0.5, 0, 0]]), bestvar=np.array([[1, 0, 0]], np.int32),
             nodeclass=np.array([[0, 1, 2]], np.int32), ndbigtree=np.array([3], np.int32),
             orig_labels=np.array([-1, 1], np.int32))
    with pytest.raises(api.GliaError):
        api.Forest.from_arrays(f, 1)


def test_bad_arguments_are_status_codes_not_exits():
    L = api.lib()
    p = api.Params()
    p.ndim = 5
    h = C.c_void_p()
    assert L.glia_b200_create(C.byref(p), C.byref(h)) == 1  # GLIA_E_ARG
    assert L.glia_b200_rf_load(b"/nonexistent/model.bin", 0, C.byref(h)) == 5  # GLIA_E_IO


def test_oracle_is_not_linked_into_the_product():
    import subprocess
    out = subprocess.run(["nm", "-D", api.LIB_PATH], capture_output=True, text=True).stdout
    assert "go_build" not in out and "rf_oracle" not in out
    for fn in os.listdir(os.path.join(ROOT, "glia_b200")):
        if fn.endswith(".py"):
            assert "oracle" not in open(os.path.join(ROOT, "glia_b200", fn)).read().lower(), fn


def test_every_declared_symbol_is_documented():
    """INTEGRATION.md shows the reference-side binding (or the purpose) of every entry point the header declares."""
    doc = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    missing = [s for s in _declared() if s not in doc]
    assert not missing, missing
