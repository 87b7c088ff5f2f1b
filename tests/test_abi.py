"""CPU-side checks of the drop-in boundary: the C-ABI library loads and exports every symbol
include/b200sv.h declares, and refuses to run without a CUDA device (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared(header):
    txt = open(os.path.join(ROOT, "include", header)).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(b200sv_[a-z0-9_]+|b200k_[a-z0-9_]+)\s*\(", txt)))


@pytest.fixture(scope="module")
def libpath():
    from catalyst_b200 import build
    return build.build_sv()


def test_library_exports_every_declared_symbol(libpath):
    L = C.CDLL(libpath)
    names = _declared("b200sv.h") + _declared("b200sv_kernels.h")
    assert len(names) > 50
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing


def test_python_binding_lists_the_header(libpath):
    from catalyst_b200 import capi
    declared = [n for n in _declared("b200sv.h") if n not in ("b200sv_rng_next_fn",)]
    assert sorted(capi.SYMBOLS) == sorted(declared)


def test_no_cpu_fallback(libpath):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from catalyst_b200 import capi
    with pytest.raises(capi.B200Error, match="no CUDA device"):
        capi.B200Device(2)


def test_product_never_imports_oracle():
    bad = []
    for dirpath, _, files in os.walk(os.path.join(ROOT, "catalyst_b200")):
        for f in files:
            if f.endswith((".py", ".cpp", ".cu", ".hpp", ".h")):
                src = open(os.path.join(dirpath, f), errors="ignore").read()
                if re.search(r"^\s*(from|import)\s+oracle|sv_oracle|#include\s+\"\.\./\.\./oracle", src, re.M):
                    bad.append(f)
    assert not bad, bad
