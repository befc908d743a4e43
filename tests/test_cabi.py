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


def test_header_is_plain_c(tmp_path):
    """include/rt.h is what a C or C++ host binds: it must compile as C99 on its own, with the documented constants"""
    import subprocess

    src = tmp_path / "hdr.c"
    src.write_text('#include "rt.h"\n'
                   "int main(void) { rt_config c = {0}; rt_uniforms u; rt_shape s; rt_stats st; (void)u; (void)s; (void)st;\n"
                   "  c.max_shapes = RT_MAX_SHAPES_LIMIT; c.flags = RT_FLAG_RECTANGLES | RT_FLAG_PIPELINE;\n"
                   "  return (RT_ABI_VERSION == 2 && RT_MAX_SHAPES == 10 && sizeof(rt_config) == 64 && sizeof(rt_shape) == 80 && c.flags == 12) ? 0 : 1; }\n")
    exe = tmp_path / "hdr"
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"),
                    str(src), "-o", str(exe)], check=True, capture_output=True)
    assert subprocess.run([str(exe)]).returncode == 0


def test_create_validates_before_touching_a_device():
    """bad arguments are RT_EINVAL with a message, with or without a GPU (the checks come first)"""
    lib = _lib.load()
    h = C.c_void_p()
    for kw, words in ((dict(flags=64), b"unknown flags"), (dict(max_shapes=257), b"max_shapes"), (dict(max_shapes=-1), b"max_shapes"),
                      (dict(flags=_lib.FLAG_PHONG_POSTPROCESS), b"RT_FLAG_PHONG_GBUFFER"), (dict(aa=0), b"aa"),
                      (dict(num_frames=17), b"num_frames")):
        cfg = _lib.rt_config(64, 48, 4, 20, 8, 0, 0, 0, 0, 0, 1)
        for k, v in kw.items():
            setattr(cfg, k, v)
        assert lib.rt_create(C.byref(cfg), C.byref(h)) == _lib.RT_EINVAL, kw
        assert words in lib.rt_last_error(None), (kw, lib.rt_last_error(None))
