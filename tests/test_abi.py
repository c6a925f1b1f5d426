"""CPU tier: the C-ABI library loads and exports every symbol include/rt_b200.h declares (no compute calls)."""
import ctypes
import os
import re

from pyrt import lib as rtlib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "rt_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rt_[a-z0-9_]+)\s*\(", src)))


def test_header_and_binding_agree():
    assert _declared_symbols() == sorted(rtlib.ABI_SYMBOLS)


def test_library_exports_every_declared_symbol():
    assert os.path.exists(rtlib.LIB_PATH), "librt_b200.so is not built: run `python __graft_entry__.py build`"
    dll = ctypes.CDLL(rtlib.LIB_PATH)
    for name in _declared_symbols():
        assert hasattr(dll, name), name
    dll.rt_abi_version.restype = ctypes.c_int
    assert dll.rt_abi_version() == 2


def test_struct_layouts_match_header():
    from pyrt import types as T

    # rt_object: 13 material floats + 2 x 16 + type; rt_light 13 floats; rt_ray 8 floats (include/rt_b200.h)
    assert T.OBJECT_DTYPE.itemsize == 4 * (13 + 32 + 1)
    assert T.LIGHT_DTYPE.itemsize == 4 * 13
    assert T.RAY_DTYPE.itemsize == 32
    assert ctypes.sizeof(rtlib.Stats) == 10 * 8 + 2 * 4 + 4 * 4


def test_no_cpu_fallback_without_device():
    """On a box without a GPU rt_create must fail loudly (RT_ERR_NO_DEVICE), never fall back."""
    import pytest

    try:
        import torch

        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = os.path.exists("/dev/nvidia0")
    if has_gpu:
        pytest.skip("GPU present")
    with pytest.raises(rtlib.RtError) as e:
        rtlib.Context(0)
    assert e.value.code == -2


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "opencl-raytracer_b200")
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h")):
                txt = open(os.path.join(d, f), errors="replace").read()
                assert "import oracle" not in txt and "oracle/" not in txt and "liboracle" not in txt, os.path.join(d, f)
