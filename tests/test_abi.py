"""CPU: the C-ABI library builds, loads and exports every symbol include/*.h declares."""
import ctypes
import os
import re

from tests.conftest import ROOT


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "nonmyopia_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(nmb_[a-z0-9_]+)\s*\(", src)))


def test_library_builds_and_exports_header_symbols():
    from nonmyopia_b200 import build
    path = build.build()
    lib = ctypes.CDLL(path)
    syms = _declared_symbols()
    assert len(syms) >= 5
    missing = [s for s in syms if not hasattr(lib, s)]
    assert not missing, missing
    lib.nmb_version.restype = ctypes.c_int
    assert lib.nmb_version() >= 100


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "nonmyopia_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                txt = open(os.path.join(dp, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M), f
