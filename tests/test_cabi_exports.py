"""CPU-side checks of the boundary: the shared library builds, loads without a GPU,
and exports every symbol include/hypatia_b200.h declares (no compute calls here)."""
import ctypes
import os
import re

import pytest

import hypatia_b200  # noqa: F401
from hypatia_b200 import lib as hlib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "hypatia_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(hyb_[a-zA-Z0-9_]+)\s*\(", text)))


@pytest.fixture(scope="module")
def built_lib():
    if not os.path.exists(hlib.LIB_PATH):
        import __graft_entry__ as ge

        ge.build()
    return ctypes.CDLL(hlib.LIB_PATH)


def test_header_declares_the_expected_surface():
    syms = header_symbols()
    for required in ("hyb_ctx_create", "hyb_model_load", "hyb_cones_check", "hyb_schur_update", "hyb_schur_solve",
                     "hyb_solve_subsystem3", "hyb_G_mul", "hyb_outer_prod", "hyb_posdef_solve"):
        assert required in syms


def test_library_exports_every_declared_symbol(built_lib):
    for name in header_symbols():
        assert hasattr(built_lib, name), f"{name} declared in include/hypatia_b200.h but not exported"


def test_python_binding_matches_header():
    assert sorted(hlib.SIGNATURES) == header_symbols()


def test_cone_desc_layout():
    # struct hyb_cone_desc: 2 x int32, 4 x int64, 2 x int32, 2 pointers
    assert ctypes.sizeof(hlib.ConeDesc) == 8 + 32 + 8 + 16


def test_no_cpu_fallback_without_gpu(built_lib):
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(hlib.HybError):
        hlib.Context(0)


def test_product_path_does_not_import_oracle():
    pkg = os.path.join(ROOT, "hypatia.jl_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f
