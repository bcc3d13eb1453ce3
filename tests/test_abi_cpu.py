"""CPU (-m "not gpu"): the C-ABI library builds for sm_100a, loads without a GPU and exports every
symbol include/gabril_b200.h declares; the product refuses to run without its CUDA extension."""
import ctypes
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "gabril_b200.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    src = re.sub(r"//[^\n]*", "", src)
    return sorted(set(re.findall(r"\b(gabril_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def lib():
    import gabril_carla_b200  # noqa: F401
    from gabril_carla_b200 import _lib
    if not os.path.exists(_lib.lib_path()):
        import __graft_entry__ as g
        g.build()
    return _lib.load()


def test_header_declares_the_path():
    syms = declared_symbols()
    for must in ("gabril_obs_patches", "gabril_gaze_splat", "gabril_conv_fwd", "gabril_conv_dgrad",
                 "gabril_conv_wgrad", "gabril_reg_mse", "gabril_gmd_mask", "gabril_adam", "gabril_init",
                 "gabril_last_error"):
        assert must in syms, must
    assert len(syms) >= 25


def test_library_exports_every_declared_symbol(lib):
    missing = [s for s in declared_symbols() if not hasattr(lib, s)]
    assert not missing, missing
    lib.gabril_version.restype = ctypes.c_int
    assert lib.gabril_version() >= 1


def test_no_torch_types_in_signatures():
    src = open(HEADER).read()
    assert 'extern "C"' in src
    code = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    code = re.sub(r"//[^\n]*", "", code)
    assert "at::" not in code and "torch" not in code.lower() and "Tensor" not in code


def test_library_is_sm100a_only(lib):
    from gabril_carla_b200 import _lib
    out = subprocess.run(["cuobjdump", "--list-elf", _lib.lib_path()], capture_output=True, text=True)
    if out.returncode != 0:
        pytest.skip("cuobjdump unavailable")
    archs = set(re.findall(r"sm_(\d+a?)", out.stdout))
    assert archs == {"100a"}, archs


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from gabril_carla_b200 import _lib
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        _lib.init()


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "gabril-carla_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                s = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", s, flags=re.M), f
