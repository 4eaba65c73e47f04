"""CPU-side checks of the drop-in boundary: the C-ABI library loads and exports every
symbol include/b2g16.h declares; without a CUDA device it fails loudly (no fallback)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "b2g16.h")


def declared_symbols():
    txt = open(HEADER).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(b2g16_[a-z0-9_]+)\s*\(", txt)))


def test_header_symbols_exported():
    from b2g16 import lib
    so = lib.load()
    names = declared_symbols()
    assert len(names) >= 30
    for n in names:
        assert hasattr(so, n), "missing export: " + n


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    from b2g16 import lib
    with pytest.raises(lib.B2g16Error) as e:
        lib.Context(0)
    assert e.value.code == -4            # B2G16_E_CUDA


def test_product_path_never_imports_oracle():
    """The oracle is test infrastructure: nothing under the package may reference it."""
    pkg = os.path.join(ROOT, "crypto3-fil-proofs_b200")
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(d, f)).read()
                assert "coracle" not in txt and "oracle/" not in txt and "liboracle" not in txt, os.path.join(d, f)
