"""CPU: the C-ABI library loads and exports every symbol include/iddiff_b200.h declares; the Python
binding covers all of them; and the product package never touches oracle/ or a CPU fallback."""
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "iddiff_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(idb_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    import __graft_entry__ as ge
    ge.build()
    from iddiff_b200 import _lib
    lib = _lib.load()
    names = _declared()
    assert len(names) >= 13
    for n in names:
        assert hasattr(lib, n), n
        assert n in _lib.SIGNATURES, f"{n} missing from the ctypes binding"
    assert sorted(_lib.SIGNATURES) == names
    assert lib.idb_abi_version() == 1


def test_argument_validation_without_a_gpu():
    """Rejected calls return before touching CUDA, so this runs on the CPU box."""
    from iddiff_b200 import _lib
    with pytest.raises(_lib.IdbError, match="NULL operand"):
        _lib.call("idb_fcn_output_layer", None, None, None, None, 0, 1.0, 128, 256, 64, 16, None)
    with pytest.raises(_lib.IdbError, match="n >= 3"):
        _lib.call("idb_id_from_spectrum", 16, 1, 2, 2, 16, None)
    import torch
    with pytest.raises(_lib.IdbError, match="CUDA tensors only"):
        _lib.ptr(torch.zeros(3))


def test_product_never_imports_oracle_or_falls_back():
    pkg = os.path.join(ROOT, "iddiff_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), f
                assert "/root/reference" not in text, f
