"""b200pt_film_read_srgb8_static: the sRGB8 conversion staged in a file-scope device array (no cudaMalloc in the call).

The entry point has no counterpart in the reference; its oracle is the existing
entry point b200pt_film_read_srgb8, which the other render tests pin against the reference: the two run the same arithmetic, so
the bytes must be EQUAL.  Checked on two calls in a row on films of different sizes and contents (the static array keeps the
last call's pixels), through the C-ABI.  This file is named to be collected last.
"""
import os
import re

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
STATIC_MAX_PIXELS = 1 << 21      # include/b200pt.h: b200pt_film_read_srgb8_static


def rendered_film(b200pt, name, res, spp):
    text = open(os.path.join(ROOT, "tests", "scenes", name)).read()
    text = re.sub(r'("integer xresolution"\s*\[\s*)\d+', r"\g<1>%d" % res[0], text)
    text = re.sub(r'("integer yresolution"\s*\[\s*)\d+', r"\g<1>%d" % res[1], text)
    sc = b200pt.pbrt_scene.load(text=text, spp=spp, scene_dir=os.path.join(ROOT, "tests", "scenes"))
    integ = b200pt.WavefrontPathIntegrator(sc, device=0)
    film = integ.create_film()
    if spp:
        integ.render(film)
    return integ, film


def test_static_srgb8_equals_the_allocating_entry_point_twice_in_a_row(b200pt):
    ia, fa = rendered_film(b200pt, "cornell-box-specular.pbrt", (96, 64), 16)
    ib, fb = rendered_film(b200pt, "materials.pbrt", (40, 56), 4)
    try:
        want_a, nan_a = fa.srgb8()
        want_b, nan_b = fb.srgb8()
        assert want_a.std() > 1.0 and want_b.std() > 1.0          # real images, not constants
        got_a, gnan_a = fa.srgb8_static()                          # first call
        got_b, gnan_b = fb.srgb8_static()                          # second call: smaller film, the array holds film a's pixels
        assert got_a.shape == (64, 96, 4) and got_b.shape == (56, 40, 4)
        assert np.array_equal(got_a, want_a) and gnan_a == nan_a
        assert np.array_equal(got_b, want_b) and gnan_b == nan_b
        again_a, _ = fa.srgb8_static()                             # and back: nothing of film b is left over
        assert np.array_equal(again_a, want_a)
        # a NaN pixel is counted per call (the counter lives in the shared array and is cleared by every call)
        raw = fa.raw()
        raw[3, 5, 0] = np.nan
        b200pt.api._check(fa.lib.b200pt_copy_to_device(fa.accel.handle, fa.device_ptr, raw.ctypes.data, raw.nbytes))
        w1, n1 = fa.srgb8()
        g1, m1 = fa.srgb8_static()
        g2, m2 = fa.srgb8_static()
        assert n1 == m1 == m2 == 1 and np.array_equal(g1, w1) and np.array_equal(g2, w1)
    finally:
        fa.close()
        fb.close()
        ia.close()
        ib.close()


def test_static_srgb8_refuses_films_beyond_its_array(b200pt):
    integ, film = rendered_film(b200pt, "cornell-box-specular.pbrt", (2048, 1025), 0)      # 2,099,200 pixels, never rendered
    try:
        assert film.width * film.height > STATIC_MAX_PIXELS
        with pytest.raises(b200pt.B200ptError) as e:
            film.srgb8_static()
        assert "static" in str(e.value) and "pixels" in str(e.value)
        out, nan = film.srgb8()                                     # the allocating entry point still takes it
        assert out.shape == (1025, 2048, 4) and nan == 0
    finally:
        film.close()
        integ.close()
