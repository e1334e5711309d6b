"""CPU: the C-ABI library loads and exports every symbol include/b200pt.h declares; host-side
error behaviour without a GPU (no compute calls)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_header_symbols_exported(b200pt):
    lib = b200pt.load_library()
    hdr = open(os.path.join(ROOT, "include", "b200pt.h")).read()
    declared = set(re.findall(r"\b(b200pt_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    assert declared == set(b200pt.api.SYMBOLS)
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.b200pt_abi_version() == 1


def test_no_cpu_fallback_without_device(b200pt):
    lib = b200pt.load_library()
    if lib.b200pt_device_count() > 0:
        pytest.skip("GPU present")
    import numpy as np
    with pytest.raises(b200pt.B200ptError) as e:
        b200pt.HLBVH(np.zeros((3, 3), np.float32), np.array([[0, 1, 2]], np.int32))
    assert e.value.code == -3  # B200PT_ERR_NO_DEVICE: fails loudly, never computes on the CPU


def test_invalid_arguments(b200pt):
    lib = b200pt.load_library()
    assert lib.b200pt_scene_create(None, 0, None) == -1
    assert b"null" in lib.b200pt_last_error()
    assert lib.b200pt_bvh_build(None, None) == -1


def test_struct_layouts_match_header(b200pt):
    # sizes the C compiler gives the header's structs == ctypes mirrors
    import subprocess, tempfile, textwrap
    src = textwrap.dedent("""
        #include <stdio.h>
        #include "b200pt.h"
        int main(void) {
            printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu\\n", sizeof(b200pt_device_ray_set), sizeof(b200pt_texture), sizeof(b200pt_image),
                   sizeof(b200pt_medium), sizeof(b200pt_object), sizeof(b200pt_instance), sizeof(b200pt_ray_set), sizeof(b200pt_spectrum), sizeof(b200pt_material),
                   sizeof(b200pt_light), sizeof(b200pt_sphere), sizeof(b200pt_camera), sizeof(b200pt_filter),
                   sizeof(b200pt_film_desc), sizeof(b200pt_scene_desc), sizeof(b200pt_build_stats),
                   sizeof(b200pt_render_opts), sizeof(b200pt_render_stats));
            return 0;
        }""")
    with tempfile.TemporaryDirectory() as d:
        c = os.path.join(d, "t.c")
        open(c, "w").write(src)
        exe = os.path.join(d, "t")
        subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), c, "-o", exe], check=True)
        sizes = list(map(int, subprocess.run([exe], capture_output=True, text=True, check=True).stdout.split()))
    a = b200pt.api
    mirrors = [a.DeviceRaySet, a.Texture, a.Image, a.Medium, a.Object, a.Instance, a.RaySet, a.Spectrum, a.Material, a.Light, a.Sphere, a.Camera, a.Filter, a.FilmDesc, a.SceneDesc, a.BuildStats,
               a.RenderOpts, a.RenderStats]
    assert sizes == [C.sizeof(m) for m in mirrors]
