"""GPU test of mib_sample_lanes_noalloc (include/mi_b200.h): the allocation-free variant of mib_debug_sample_lanes.

The entry point has no counterpart in the reference; it exists for callers that cannot call cudaMalloc.  Its per-lane
scratch is a static device array shared by every call, so the checks are:
  * it returns what the allocating entry point returns for the same lanes (same kernel, same lanes: bit-identical, which
    is inside the 1e-4 relative tolerance of the lane parity tests), on two calls in a row;
  * the second call is not polluted by what the first one left in the array (different scene and variant, fewer lanes);
  * a request beyond MIB_NOALLOC_MAX_FLOATS is refused with a clear error, and the library still works afterwards.

Collected last on purpose: the file name sorts after every other test file.
"""
import ctypes as C

import numpy as np
import pytest

from mitsuba3_nasa_b200 import scene as S, scenes, _lib

pytestmark = pytest.mark.gpu

RTOL = 1e-4          # the tolerance of the lane-by-lane parity tests (test_gpu_parity_forward.py)


def _close(a, b):
    a, b = a.astype(np.float64), b.astype(np.float64)
    scale = np.maximum(np.abs(b).max(axis=1, keepdims=True), 1e-6)
    return (np.abs(a - b) <= RTOL * scale).all()


def _cloud():
    grid = np.random.default_rng(3).uniform(0.0, 1.0, (12, 12, 12, 1)).astype(np.float32)
    d = scenes.c3_cloud(res=16, spp=8, max_depth=16, grid=grid, optical_depth=6.0, phase={'type': 'rayleigh'})
    return _lib.DeviceScene(S.load_dict(d, 'llvm_ad_mono_polarized'), 0)


def _cube():
    return _lib.DeviceScene(S.load_dict(scenes.c1_homogeneous_cube(res=16, spp=16), 'llvm_ad_rgb'), 0)


def test_no_alloc_matches_allocating_entry_point_twice_in_a_row():
    cloud, cube = _cloud(), _cube()
    n1 = 16 * 16 * 8                                     # polarized cloud: 4 floats per lane
    ref1 = cloud.sample_lanes(2, 8, 0, n1)
    got1 = cloud.sample_lanes_noalloc(2, 8, 0, n1)
    assert got1.shape == ref1.shape == (n1, 4)
    assert np.abs(ref1).max() > 0
    assert _close(got1, ref1)
    # second call straight after: another scene and variant, a ragged lane range inside the wavefront -- the
    # array still holds the first call's lanes and must not leak into this result
    b, n2 = 77, 1501
    ref2 = cube.sample_lanes(1, 16, b, n2)
    got2 = cube.sample_lanes_noalloc(1, 16, b, n2)
    assert got2.shape == ref2.shape == (n2, cube.raw_channels)
    assert _close(got2, ref2)
    # a third scene with another channel count (mono: I and weight per lane)
    mono = _lib.DeviceScene(S.load_dict(scenes.c3_cloud(res=16, spp=8, max_depth=16, grid=scenes.cloud_grid(12),
                                                        phase={'type': 'hg', 'g': 0.6}), 'llvm_ad_mono'), 0)
    ref_m = mono.sample_lanes(3, 8, 5, 999)
    got_m = mono.sample_lanes_noalloc(3, 8, 5, 999)
    assert got_m.shape == ref_m.shape == (999, 2)
    assert _close(got_m, ref_m)
    # and the first request again, from the no-alloc entry point twice in a row
    got3 = cloud.sample_lanes_noalloc(2, 8, 0, n1)
    got4 = cloud.sample_lanes_noalloc(2, 8, 0, n1)
    assert np.array_equal(got3, got1) and np.array_equal(got4, got1)
    # the scene's own lane range is untouched by the call (a later film render walks the whole wavefront)
    img = cloud.render(2, 8, develop=False)
    assert cloud.counters()['n_samples'] == n1
    assert np.isfinite(np.asarray(img)).all()


def test_no_alloc_rejects_sizes_beyond_the_static_array():
    cube = _cube()
    cap = _lib.DeviceScene.NOALLOC_MAX_FLOATS
    ch = cube.raw_channels                               # rgb: R, G, B, weight per lane
    too_many = cap // ch + 1
    out = np.empty((1, ch), np.float32)                  # never written: the call must fail before any launch
    rc = cube.l.mib_sample_lanes_noalloc(cube.h, C.c_uint32(1), C.c_uint32(16), C.c_uint64(0), C.c_uint64(too_many),
                                         out.ctypes.data_as(C.POINTER(C.c_float)))
    assert rc != 0
    msg = cube.l.mib_last_error().decode()
    assert 'MIB_NOALLOC_MAX_FLOATS' in msg and str(too_many) in msg
    rc = cube.l.mib_sample_lanes_noalloc(cube.h, C.c_uint32(1), C.c_uint32(16), C.c_uint64(0), C.c_uint64(0),
                                         out.ctypes.data_as(C.POINTER(C.c_float)))
    assert rc != 0
    # the library is still usable, and the largest request that fits goes through (lanes beyond the 16x16x16 wavefront
    # are not requested: the range stays inside it)
    n = 16 * 16 * 16
    assert _close(cube.sample_lanes_noalloc(1, 16, 0, n), cube.sample_lanes(1, 16, 0, n))


def test_no_alloc_exact_maximum_size():
    """Maximum size: 256 x 256 pixels x 4 spp x 4 Stokes floats = MIB_NOALLOC_MAX_FLOATS exactly."""
    grid = np.random.default_rng(4).uniform(0.0, 1.0, (8, 8, 8, 1)).astype(np.float32)
    d = scenes.c3_cloud(res=256, spp=4, max_depth=8, grid=grid, optical_depth=3.0, phase={'type': 'rayleigh'})
    dev = _lib.DeviceScene(S.load_dict(d, 'llvm_ad_mono_polarized'), 0)
    n = 256 * 256 * 4
    assert n * 4 == _lib.DeviceScene.NOALLOC_MAX_FLOATS
    ref = dev.sample_lanes(5, 4, 0, n)
    for _ in range(2):
        assert _close(dev.sample_lanes_noalloc(5, 4, 0, n), ref)
