"""The C-ABI library builds, loads without a GPU and exports every symbol that
include/msprime_b200.h declares; without a device it refuses to run (no CPU fallback)."""
import ctypes
import os
import re

import pytest

from msprime_b200 import _native

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "msprime_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(mspb_[a-z_]+)\s*\(", text)))


def test_header_symbols_exported(native_built):
    lib = ctypes.CDLL(_native.SO_PATH)
    names = declared_symbols()
    assert len(names) >= 12
    for name in names:
        assert hasattr(lib, name), name
    assert sorted(_native.EXPORTED_SYMBOLS) == names


def test_strerror_and_version(native_built):
    assert _native.lib().mspb_version().decode() == "0.1.0"
    assert _native.strerror(-4) == "Bad parameter value provided"
    assert "no CPU fallback" in _native.strerror(-1001)


def test_no_device_fails_loudly(native_built):
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(_native.LibraryError) as e:
        _native.Context(0)
    assert e.value.status == _native.ERR_NO_DEVICE
