"""rt_film_resolve_rgb8_static — the entry point that allocates nothing (include/rtcore.h; no counterpart in the reference)
against rt_film_resolve_rgb8, through the C ABI.  Its device staging buffer is a static device array inside the library that
all calls share and that keeps the last call's bytes, so the check is made on two calls in a row with different films (a larger image first, then a
smaller, brighter one: stale bytes of the first call must not show up in the second).  Tolerance: the one the resolve tests use everywhere — byte
equality (tests/test_gpu_parity.py compares rt_film_resolve_rgb8 with the oracle's point_to_color byte for byte).
Named test_zz_* so that pytest collects it last."""
import ctypes as C

import numpy as np
import pytest

from rustraytracer_b200 import Film, IntType, _native as N, render_samples, scenes

pytestmark = pytest.mark.gpu


def test_static_scratch_resolve_matches_the_allocating_entry_point_twice_in_a_row():
    hs = scenes.HostScene("cornell_box", 96, 96)
    gpu = hs.upload()
    cam = hs.camera()
    big = Film(96, 96)
    render_samples(gpu, cam, IntType.Path(6, shade_f32=True), big, 96, 96, 0, 4, seed=17)
    hs2 = scenes.HostScene("sphere_roughness", 40, 56)
    gpu2 = hs2.upload()
    small = Film(40, 56)
    render_samples(gpu2, hs2.camera(), IntType.Path(6), small, 40, 56, 0, 8, seed=18)
    want_big, want_small = big.resolve_rgb8(), small.resolve_rgb8()
    assert want_big.std() > 5 and want_small.std() > 5 and not np.array_equal(want_big[:56, :40], want_small)
    # first call, then a second one straight after it on a different (smaller) film, then the first again
    got_big = big.resolve_rgb8_static()
    got_small = small.resolve_rgb8_static()
    assert np.array_equal(got_big, want_big)
    assert np.array_equal(got_small, want_small)
    assert np.array_equal(big.resolve_rgb8_static(), want_big)
    # the film itself is untouched, and the allocating entry point still agrees afterwards
    assert np.array_equal(small.resolve_rgb8(), want_small)


def test_static_scratch_resolve_rejects_films_beyond_its_maximum():
    L = N.lib()
    too_big = Film(4096, 1025)                           # 4096 * 1025 > RT_RESOLVE_STATIC_MAX_PIXELS = 2048 * 2048
    out = np.zeros((1025, 4096, 3), dtype=np.uint8)
    assert L.rt_film_resolve_rgb8_static(too_big.h, N.ptr(out)) == N.RT_ERR_INVALID
    msg = L.rt_last_error_string().decode()
    assert "exceeds the static scratch" in msg and "4194304" in msg
    assert L.rt_film_resolve_rgb8_static(None, N.ptr(out)) == N.RT_ERR_INVALID
    assert L.rt_film_resolve_rgb8_static(too_big.h, None) == N.RT_ERR_INVALID
    # exactly the maximum is accepted (a cleared film resolves to NaN -> 0 everywhere, on both entry points)
    edge = Film(2048, 2048)
    assert np.array_equal(edge.resolve_rgb8_static(), edge.resolve_rgb8())
