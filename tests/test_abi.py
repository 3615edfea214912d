"""CPU: the C-ABI library loads and exports every symbol include/fic.h declares; no compute calls."""
import ctypes
import os
import re

import pytest

from conftest import ROOT


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "fic.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(fic_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_expected_surface():
    syms = declared_symbols()
    for must in ("fic_bf_encode", "fic_bf_decode", "fic_qt_encode", "fic_hv_encode", "fic_tree_decode",
                 "fic_average_subsample", "fic_create", "fic_destroy", "fic_last_error"):
        assert must in syms


def test_library_exports_every_declared_symbol():
    import fic_b200
    L = ctypes.CDLL(fic_b200.LIB_PATH)
    for name in declared_symbols():
        assert hasattr(L, name), f"{name} declared in include/fic.h but not exported"


def test_binding_covers_every_declared_symbol():
    import fic_b200
    assert sorted(fic_b200.SIGNATURES) == declared_symbols()


def test_no_cpu_fallback_without_gpu():
    import fic_b200
    if fic_b200.lib().fic_device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(fic_b200.FicError, match="no CPU fallback"):
        fic_b200.Handle(0)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "fractal-image-compression_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src.lower() or f in (), f"{f} mentions the oracle"
