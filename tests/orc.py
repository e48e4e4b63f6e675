"""Test-side ctypes wrapper of the parity oracle (oracle/liboracle.so) and of the
reference-header shim (oracle/_ref/libmiref.so).  TEST INFRASTRUCTURE ONLY: the
product package never imports this module."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")

from mitsuba3_nasa_b200 import _abi as A  # noqa: E402  (POD layout only)

_lib = None
_ref = None


def oracle_lib():
    global _lib
    if _lib is None:
        path = os.path.join(ORACLE_DIR, "liboracle.so")
        if not os.path.exists(path):
            subprocess.check_call(["make", "-C", ORACLE_DIR, "liboracle.so"])
        _lib = C.CDLL(path)
        _lib.orc_last_error.restype = C.c_char_p
    return _lib


def ref_lib():
    """Reference headers compiled in oracle/_ref (None if neither prebuilt nor buildable)."""
    global _ref
    if _ref is None:
        path = os.path.join(ORACLE_DIR, "_ref", "libmiref.so")
        if not os.path.exists(path) and os.path.isdir("/root/reference/include/mitsuba"):
            subprocess.call(["make", "-C", ORACLE_DIR, "ref"])
        if os.path.exists(path):
            _ref = C.CDLL(path)
    return _ref


def fp(a):
    return a.ctypes.data_as(C.c_void_p)


class OracleScene:
    def __init__(self, scene):
        """scene: mitsuba3_nasa_b200.scene.Scene (flattened description + buffers)."""
        self.lib = oracle_lib()
        self.h = C.c_void_p()
        self.spectrum = scene.desc.spectrum
        self.W, self.H = scene.desc.sensor.crop_width, scene.desc.sensor.crop_height
        self._check(self.lib.orc_scene_create(C.byref(scene.desc), C.byref(self.h)))
        for slot, arr in scene.buffers.items():
            self.set_buffer(slot, arr)
        self.commit()

    def _check(self, rc):
        if rc != 0:
            raise RuntimeError(self.lib.orc_last_error().decode())

    def set_buffer(self, slot, arr):
        arr = np.ascontiguousarray(arr, dtype=np.float32).reshape(-1)
        self._check(self.lib.orc_scene_set_buffer(self.h, C.c_int(slot), fp(arr), C.c_size_t(arr.size)))

    def set_desc(self, desc):
        self._check(self.lib.orc_scene_set_desc(self.h, C.byref(desc)))

    def commit(self):
        self._check(self.lib.orc_scene_commit(self.h))

    @property
    def raw_channels(self):
        return 2 if self.spectrum == A.SPEC_MONO else 4

    @property
    def dev_channels(self):
        return {A.SPEC_MONO: 1, A.SPEC_RGB: 3, A.SPEC_MONO_POLARIZED: 4}[self.spectrum]

    def render(self, seed, spp, lanes=(0, 0), nthreads=8):
        raw = np.zeros((self.H, self.W, self.raw_channels), np.float32)
        dev = np.zeros((self.H, self.W, self.dev_channels), np.float32)
        self._check(self.lib.orc_render(self.h, C.c_uint32(seed), C.c_uint32(spp), C.c_uint64(lanes[0]), C.c_uint64(lanes[1]),
                                        C.c_int(nthreads), fp(raw), fp(dev)))
        return raw, dev

    def sample_lanes(self, seed, spp, lane_begin, count, nthreads=8):
        out = np.zeros((count, self.raw_channels), np.float32)
        self._check(self.lib.orc_sample_lanes(self.h, C.c_uint32(seed), C.c_uint32(spp), C.c_uint64(lane_begin),
                                              C.c_uint64(count), C.c_int(nthreads), fp(out)))
        return out

    def sample_pixels_scalar(self, pixel_index, pixel_seed, spp):
        """Per-sample aovs under the seeding of the reference's scalar_* variants: [pixel][sample][C]."""
        pi = np.ascontiguousarray(pixel_index, np.uint32)
        ps = np.ascontiguousarray(pixel_seed, np.uint32)
        out = np.zeros((pi.size, spp, self.raw_channels), np.float32)
        self._check(self.lib.orc_sample_pixels_scalar(self.h, fp(pi), fp(ps), C.c_uint32(pi.size), C.c_uint32(spp), fp(out)))
        return out

    def render_1(self, seed, spp, lanes=(0, 0), nthreads=8, normalize=True):
        out = np.zeros(4, np.float64)
        self._check(self.lib.orc_render_1(self.h, C.c_uint32(seed), C.c_uint32(spp), C.c_uint64(lanes[0]), C.c_uint64(lanes[1]),
                                          C.c_int(nthreads), C.c_int(int(normalize)), fp(out)))
        return out

    # ---- adjoint (oracle/oracle_adjoint_api.inc) ----
    def render_backward(self, seed, spp, grad_in, lanes=(0, 0), nthreads=8):
        g = np.ascontiguousarray(np.asarray(grad_in, np.float32).reshape(self.H, self.W, self.dev_channels))
        self._check(self.lib.orc_render_backward(self.h, C.c_uint32(seed), C.c_uint32(spp), fp(g), C.c_uint64(lanes[0]),
                                                 C.c_uint64(lanes[1]), C.c_int(nthreads)))

    def render_1_backward(self, seed, spp, grad4, lanes=(0, 0), nthreads=8):
        g = np.zeros(4, np.float64)
        gi = np.asarray(grad4, np.float64).reshape(-1)
        g[:gi.size] = gi
        self._check(self.lib.orc_render_1_backward(self.h, C.c_uint32(seed), C.c_uint32(spp), fp(g), C.c_uint64(lanes[0]),
                                                   C.c_uint64(lanes[1]), C.c_int(nthreads)))

    def grad_get(self, slot, count):
        out = np.zeros(count, np.float64)
        self._check(self.lib.orc_grad_get(self.h, C.c_int(slot), fp(out), C.c_size_t(count)))
        return out

    def grad_get_scalar(self, medium=0):
        out = np.zeros(4, np.float64)
        self._check(self.lib.orc_grad_get_scalar(self.h, C.c_int(medium), fp(out)))
        return out

    def grad_zero(self):
        self._check(self.lib.orc_grad_zero(self.h))

    def counters(self):
        c = A.Counters()
        self.lib.orc_counters_get(self.h, C.byref(c))
        return c.as_dict()

    def __del__(self):
        try:
            self.lib.orc_scene_destroy(self.h)
        except Exception:
            pass
