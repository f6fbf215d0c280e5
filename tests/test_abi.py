"""The C-ABI library loads on a CPU-only box and exports every symbol include/femo_b200.h declares;
the product path fails loudly (no CPU fallback) when no GPU is present."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "femo_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(femo_[a-zA-Z0-9_]+)\s*\(", src)))


def test_header_symbols_exported(built_lib):
    lib = ctypes.CDLL(built_lib)
    names = _declared()
    assert len(names) >= 35
    for n in names:
        assert hasattr(lib, n), "libfemo_b200.so does not export " + n


def test_binding_covers_header(built_lib):
    from femo_alpha_b200 import _lib as L
    L.load()
    bound = set(L.SIGNATURES) | {"femo_last_error", "femo_version"}
    assert set(_declared()) == bound
    assert L.load().femo_version() >= 100


def test_no_cpu_fallback(built_lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from femo_alpha_b200.shell_device import ShellDevice
    nodes = np.array([[0, 0, 0], [1, 0, 0], [1, 1, 0], [0, 1, 0]], float)
    with pytest.raises(RuntimeError, match="no CPU fallback|CUDA"):
        ShellDevice(nodes, np.array([[0, 1, 2, 3]]))


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "femo_alpha_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dp, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M), f
                assert "oracle." not in txt or f.endswith(".md"), f
