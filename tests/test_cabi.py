"""The C-ABI library loads and exports exactly what include/rt.h declares; without a GPU it fails
loudly instead of falling back to a CPU path."""
import ctypes as C
import os
import re

import pytest

from real_time_ray_tracer_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    text = open(os.path.join(ROOT, "include", "rt.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rt_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported():
    names = declared_functions()
    assert len(names) >= 20
    lib = C.CDLL(_lib.LIB_PATH)
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    assert sorted(_lib.EXPORTS) == names
    assert lib.rt_abi_version() == 2


def test_struct_sizes_match_header():
    assert C.sizeof(_lib.rt_config) == 16 * 4
    assert C.sizeof(_lib.rt_stats) == 12 * 8 + 2 * 8
    assert C.sizeof(_lib.rt_halo_info) == 5 * 8


def test_no_cpu_fallback():
    lib = _lib.load()
    if lib.rt_device_count() > 0:
        pytest.skip("a CUDA device is present")
    cfg = _lib.rt_config(64, 48, 4, 20, 8, 0, 0, 0, 0, 0, 1)
    h = C.c_void_p()
    assert lib.rt_create(C.byref(cfg), C.byref(h)) == _lib.RT_ENODEVICE
    assert b"no CUDA device" in lib.rt_last_error(None)
    import real_time_ray_tracer_b200 as rtb
    with pytest.raises(rtb.RtError):
        rtb.RayTracer(64, 48)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "real_time_ray_tracer_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".hpp")):
                src = open(os.path.join(dp, f)).read()
                assert "import oracle" not in src and "from oracle" not in src, f
                assert "rt_oracle" not in src and "librt_oracle" not in src, f
