"""pt_resolve_colors_noalloc -- the entry point for callers that cannot call cudaMalloc (include/pt_api.h) -- against
pt_resolve_colors, through the C ABI.  Collected last on purpose; passes alone in a fresh
process and after the other tests."""
import ctypes as C

import numpy as np
import pytest

from pathtracer_b200 import render

pytestmark = pytest.mark.gpu


def _colours(n, seed):
    rng = np.random.default_rng(seed)
    rgb = (rng.gamma(0.6, 1.5, (n, 3)) * rng.choice([0.01, 1.0, 40.0], (n, 1))).astype(np.float32)
    rgb[: min(n, 8)] = [[0, 0, 0], [1, 1, 1], [np.nan, 0.5, 2], [np.inf, -1, 1e-9], [0.0031308, 0.04, 0.18], [-0.0, 5, 1e30], [255, 1e-30, 0.5], [3, 2, 1]][: min(n, 8)]
    return rgb


def test_noalloc_resolve_equals_the_allocating_entry_point_twice_in_a_row():
    rc = render.RenderCore(0)
    try:
        # two calls in a row on the shared scratch: a full-size one, then a SHORTER one with other data -- the second must not
        # see anything of the first (the scratch keeps the last call's contents)
        for n, seed in ((render.NOALLOC_MAX_COLORS, 1), (12345, 2)):
            rgb = _colours(n, seed)
            want = rc.resolve_colors(rgb)                    # the existing entry point (byte-exact vs the reference, test_gpu_parity)
            got = rc.resolve_colors_noalloc(rgb)
            assert got.shape == want.shape and np.array_equal(got, want)      # the tolerance of the resolve tests: byte-exact
        # and the same call repeated gives the same bytes
        assert np.array_equal(rc.resolve_colors_noalloc(rgb), want)
        # a second context on the same device shares the scratch address
        rc2 = render.RenderCore(0)
        assert np.array_equal(rc2.resolve_colors_noalloc(rgb[:777]), want[:777])
        rc2.close()
    finally:
        rc.close()


def test_noalloc_resolve_rejects_sizes_beyond_the_scratch():
    rc = render.RenderCore(0)
    try:
        L = render.lib()
        n = render.NOALLOC_MAX_COLORS + 1
        rgb = np.zeros((n, 3), np.float32)
        out = np.full(n, 0xDEADBEEF, np.uint32)
        code = L.pt_resolve_colors_noalloc(rc._h, n, rgb.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p))
        assert code == -7 and b"PT_NOALLOC_MAX_COLORS" in L.pt_last_error(rc._h)
        assert (out == 0xDEADBEEF).all()                     # nothing was written
        assert L.pt_resolve_colors_noalloc(rc._h, 0, rgb.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p)) == 0
        assert L.pt_resolve_colors_noalloc(rc._h, 4, None, out.ctypes.data_as(C.c_void_p)) == -1
    finally:
        rc.close()
