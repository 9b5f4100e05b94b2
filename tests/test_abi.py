"""CPU: the C-ABI library builds, loads and exports exactly the symbols include/sam2b.h declares; host-side error
behaviour without a GPU (no compute calls)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from sam2_opt_b200 import build, _lib
    build.build()
    return _lib.load()


def _declared():
    src = open(os.path.join(ROOT, "include", "sam2b.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(sam2b_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported_and_bound(lib):
    from sam2_opt_b200 import _lib
    names = _declared()
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/sam2b.h but not exported"
    assert sorted(_lib.SIGNATURES) == names, "ctypes SIGNATURES and include/sam2b.h disagree"


def test_no_torch_types_in_abi():
    src = open(os.path.join(ROOT, "include", "sam2b.h")).read()
    assert "torch" not in src.lower().replace("pytorch", "") and "at::" not in src and "#include <cuda" not in src


def test_version_and_error_strings(lib):
    assert b"sm_100a" in lib.sam2b_version()
    assert isinstance(lib.sam2b_last_error(), bytes)


def test_create_fails_loudly_without_gpu(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    h = ctypes.c_void_p()
    assert lib.sam2b_create(0, ctypes.byref(h)) != 0
    assert b"no CUDA device" in lib.sam2b_last_error()
    from sam2_opt_b200.engine import Engine
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        Engine({})


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "sam2_opt_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                txt = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M), f"{f} imports the oracle"
                assert "/root/reference" not in txt or f.endswith(".py") and "ref:" in txt, f
