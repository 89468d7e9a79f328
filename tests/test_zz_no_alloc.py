"""gamer_b200_fbm_static — the allocation-free entry point (include/gamer_b200.h; no counterpart in the reference)
— against the existing gamer_b200_probe_fbm, through the C ABI.

Collected last on purpose. Works alone in a fresh process and after the other tests:
it makes its own context for the existing entry point and uses nothing another test leaves behind."""
import ctypes as C

import numpy as np
import pytest

from gamer_b200 import _lib

pytestmark = pytest.mark.gpu


def fbm_static(lib, pts, octaves, pers, scale):
    pts = np.ascontiguousarray(pts, dtype=np.float32)
    out = np.full(pts.shape[0], np.nan, dtype=np.float32)
    rc = lib.gamer_b200_fbm_static(pts.ctypes.data, pts.shape[0], octaves, pers, scale, out.ctypes.data)
    return rc, out


def test_fbm_static_equals_probe_fbm_on_two_calls_in_a_row():
    from gamer_b200.rasterizer import Context
    lib = _lib.load()
    limit = lib.gamer_b200_fbm_static_max_points()
    assert limit == 262144
    rng = np.random.default_rng(11)
    first = rng.uniform(-1, 1, (limit, 3)).astype(np.float32)          # the largest call the scratch holds
    second = rng.uniform(-40, 40, (70001, 3)).astype(np.float32)        # smaller: the array still holds the first call's tail
    ctx = Context([0])
    try:
        want1 = ctx.probe_fbm(first, 10, 1.0, 1.0)
        want2 = ctx.probe_fbm(second, 9, -0.8, 0.5)
        rc1, got1 = fbm_static(lib, first, 10, 1.0, 1.0)
        rc2, got2 = fbm_static(lib, second, 9, -0.8, 0.5)
        assert rc1 == 0 and rc2 == 0, lib.gamer_b200_last_error(None)
        # the same device function on the same inputs: the tests' tolerance for fBm is 4e-6 (test_gpu_parity.test_fbm);
        # these agree bit for bit
        assert np.max(np.abs(got1 - want1)) <= 4e-6 and np.array_equal(got1, want1)
        assert np.max(np.abs(got2 - want2)) <= 4e-6 and np.array_equal(got2, want2)
        # and once more, the first input again: nothing of the second call leaks into it
        rc3, got3 = fbm_static(lib, first[:1000], 10, 1.0, 1.0)
        assert rc3 == 0 and np.array_equal(got3, want1[:1000])
    finally:
        ctx.close()


def test_fbm_static_rejects_sizes_beyond_the_scratch():
    lib = _lib.load()
    limit = lib.gamer_b200_fbm_static_max_points()
    pts = np.zeros((limit + 1, 3), dtype=np.float32)
    rc, out = fbm_static(lib, pts, 4, 0.5, 1.0)
    assert rc == -1                                                      # GAMER_E_ARG
    msg = lib.gamer_b200_last_error(None).decode()
    assert str(limit) in msg and "exceed" in msg, msg
    assert np.isnan(out).all()                                           # nothing was written
    rc, _ = fbm_static(lib, pts[:0], 4, 0.5, 1.0)
    assert rc == -1
