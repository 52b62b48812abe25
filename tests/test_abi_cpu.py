"""The C-ABI boundary without a GPU: libfsx.so loads, exports every symbol include/fsx.h declares,
the Python binding lists exactly those symbols, and the product fails loudly (no CPU fallback)."""
import ctypes
import os
import re

import pytest

import fsx
from conftest import ROOT


def _declared():
    text = open(os.path.join(ROOT, "include", "fsx.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(fsx_[a-z0-9_]+)\s*\(", text)))


def test_library_is_built_in_tree_and_loads():
    assert os.path.exists(fsx.LIB_PATH), "run __graft_entry__.build() / make -C hops-examples_b200"
    assert os.path.dirname(fsx.LIB_PATH).startswith(ROOT)
    lib = fsx.load()
    assert lib.fsx_abi_version() == 1


def test_every_declared_symbol_is_exported_and_bound():
    names = _declared()
    assert len(names) >= 35
    raw = ctypes.CDLL(fsx.LIB_PATH)
    for n in names:
        assert hasattr(raw, n), f"{n} declared in include/fsx.h but not exported by libfsx.so"
    assert sorted(fsx.SIGNATURES) == names, "fsx/_lib.py SIGNATURES must mirror include/fsx.h exactly"


def test_no_cpu_fallback_context_creation_fails_loudly_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the gpu tests")
    with pytest.raises(fsx.FsxError) as e:
        fsx.Context()
    assert "no CPU fallback" in str(e.value) or "CUDA" in str(e.value)


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "hops-examples_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h", ".scala", ".c")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                bad = re.search(r'#include\s*[<"][^>"]*oracle|libfsx_oracle|^\s*(from|import)\s+oracle|dlopen\([^)]*oracle',
                                text, flags=re.M)
                assert not bad, f"{os.path.join(dirpath, f)} links, includes or imports the oracle: {bad.group(0)!r}"
