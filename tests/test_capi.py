"""The C-ABI library loads and exports every symbol include/dv3_b200.h declares (no compute calls: no GPU needed)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "dv3_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(dv3_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from sheeprl_prey_b200 import build

    lib_path = build.build_library()
    lib = ctypes.CDLL(lib_path)
    names = declared_symbols()
    assert len(names) >= 30
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/dv3_b200.h but not exported"
    lib.dv3_version.restype = ctypes.c_int
    assert lib.dv3_version() >= 100


def test_python_binding_covers_header():
    from sheeprl_prey_b200._lib import SIGNATURES

    assert sorted(SIGNATURES) == declared_symbols()


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "sheeprl_prey_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(".py"):
                txt = open(os.path.join(dp, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M), os.path.join(dp, f)


def test_cuda_ops_fail_loudly_without_gpu():
    import pytest
    import torch

    from sheeprl_prey_b200._lib import CudaOps

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(RuntimeError):
        CudaOps("cpu")
