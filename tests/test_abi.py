"""The C-ABI library loads here (no GPU, no libcuda) and exports every symbol that
include/b200dp.h declares; without a GPU it refuses to create a context instead of
falling back to anything."""
import ctypes
import os
import re

import pytest

import b200dp
from b200dp import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "b200dp.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b2dp_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_what_the_binding_binds():
    assert header_symbols() == sorted(_lib.SIGNATURES)


def test_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(_lib.SO_PATH)
    missing = [s for s in header_symbols() if not hasattr(lib, s)]
    assert not missing, missing


def test_abi_version():
    assert _lib.load().b2dp_abi_version() == 1


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(b200dp.B200dpError) as e:
        b200dp.Context.local([0])
    assert e.value.code == _lib.ERR_NO_DEVICE
    with pytest.raises(b200dp.B200dpError):
        b200dp.Context.rank(0, 2, 0)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "parallel-and-distributed-deep-learning_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in src and "from oracle" not in src and "b2dp_oracle" not in src, f


def test_adam_alpha_is_the_oracles():
    import oracle
    for t in (1, 2, 3, 10, 1000):
        assert b200dp.adam_alpha(1e-3, 0.9, 0.999, t) == float(oracle.adam_alpha(1e-3, 0.9, 0.999, t))
