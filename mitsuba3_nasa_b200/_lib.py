"""Loader of the C-ABI library libmi_b200.so (hand-written sm_100a kernels).

There is no CPU or PyTorch fallback: if the library is missing or no CUDA device
is present, rendering raises instead of silently computing something else.
"""
import ctypes as C
import os

import numpy as np

from . import _abi as A

_HERE = os.path.dirname(os.path.abspath(__file__))
# MIB_B200_LIB: alternative build of the same library (kernel A/B experiments)
LIB_PATH = os.environ.get("MIB_B200_LIB") or os.path.join(_HERE, "libmi_b200.so")
_lib = None


class NativeLibraryError(RuntimeError):
    pass


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise NativeLibraryError(
                "libmi_b200.so not found at %s: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(make -C mitsuba3_nasa_b200/csrc). There is no CPU fallback." % LIB_PATH)
        _lib = C.CDLL(LIB_PATH)
        _lib.mib_last_error.restype = C.c_char_p
        A.check_abi(_lib)
    return _lib


def _fp(a):
    return a.ctypes.data_as(C.c_void_p)


class DeviceScene:
    """Owns one mib_scene handle on one GPU."""

    def __init__(self, scene, device=0):
        self.l = lib()
        self.scene = scene
        self.h = C.c_void_p()
        self.device_index = int(device)
        self._ck(self.l.mib_scene_create(C.byref(scene.desc), C.c_int(device), C.byref(self.h)))
        for slot, arr in scene.buffers.items():
            self.set_buffer(slot, arr)
        self.commit()

    def _ck(self, rc):
        if rc != 0:
            raise RuntimeError(self.l.mib_last_error().decode())

    # -- parameters --
    def set_buffer(self, slot, arr):
        arr = np.ascontiguousarray(arr, dtype=np.float32).reshape(-1)
        self._ck(self.l.mib_scene_set_buffer(self.h, C.c_int(slot), _fp(arr), C.c_size_t(arr.size)))

    def get_buffer(self, slot, count):
        out = np.empty(count, np.float32)
        self._ck(self.l.mib_scene_get_buffer(self.h, C.c_int(slot), _fp(out), C.c_size_t(count)))
        return out

    def buffer_dev(self, slot):
        p, n = C.c_void_p(), C.c_size_t()
        self._ck(self.l.mib_scene_buffer_dev(self.h, C.c_int(slot), C.byref(p), C.byref(n)))
        return p.value, n.value

    def set_desc(self, desc):
        self._ck(self.l.mib_scene_set_desc(self.h, C.byref(desc)))

    def get_desc(self):
        d = A.SceneDesc()
        self._ck(self.l.mib_scene_get_desc(self.h, C.byref(d)))
        return d

    def commit(self):
        self._ck(self.l.mib_scene_commit(self.h))

    def set_lane_stripes(self, rank, world, chunk_lanes):
        """Rank `rank` of `world` renders the lane chunks rank, rank + world, ... (chunk_lanes lanes each); world <= 1: off."""
        self._ck(self.l.mib_scene_set_lane_stripes(self.h, C.c_uint32(rank), C.c_uint32(world), C.c_uint64(max(int(chunk_lanes), 1))))

    def set_lane_range(self, begin, end):
        self._ck(self.l.mib_scene_set_lane_range(self.h, C.c_uint64(begin), C.c_uint64(end)))

    # -- shapes --
    @property
    def size(self):
        s = self.scene.desc.sensor
        return s.crop_height, s.crop_width

    @property
    def raw_channels(self):
        return 2 if self.scene.desc.spectrum == A.SPEC_MONO else 4

    @property
    def dev_channels(self):
        return {A.SPEC_MONO: 1, A.SPEC_RGB: 3, A.SPEC_MONO_POLARIZED: 4}[self.scene.desc.spectrum]

    # -- forward --
    def render(self, seed, spp, develop=True, to_host=True):
        H, W = self.size
        out = None
        if to_host:
            out = np.empty((H, W, self.dev_channels if develop else self.raw_channels), np.float32)
        self._ck(self.l.mib_render(self.h, C.c_uint32(seed), C.c_uint32(spp), C.c_int(int(develop)),
                                   _fp(out) if to_host else None))
        return out

    def film_dev(self):
        rp, rn, dp, dn = C.c_void_p(), C.c_size_t(), C.c_void_p(), C.c_size_t()
        self._ck(self.l.mib_film_dev(self.h, C.byref(rp), C.byref(rn), C.byref(dp), C.byref(dn)))
        return (rp.value, rn.value), (dp.value, dn.value)

    def film_develop(self):
        self._ck(self.l.mib_film_develop(self.h))

    def film_read(self, developed=True):
        H, W = self.size
        out = np.empty((H, W, self.dev_channels if developed else self.raw_channels), np.float32)
        self._ck(self.l.mib_film_read(self.h, C.c_int(int(developed)), _fp(out)))
        return out

    def render_1(self, seed, spp, normalize=True):
        out = np.zeros(4, np.float64)
        f = self.l.mib_render_1 if normalize else self.l.mib_render_1_raw
        self._ck(f(self.h, C.c_uint32(seed), C.c_uint32(spp), _fp(out)))
        return out

    def forward_kernel(self):
        """'pool' or 'thread': the kernel family render() / render_1() launch for this scene (mib_scene_forward_kernel)."""
        k = C.c_int(0)
        self._ck(self.l.mib_scene_forward_kernel(self.h, C.byref(k)))
        return 'pool' if k.value else 'thread'

    def supergrid(self, medium=0):
        """Majorant supergrid of a heterogeneous medium (fast mode, integrator 'majorant_resolution' > 0): [M, M, M] (z, y, x)."""
        M = int(self.scene.desc.integrator.majorant_res)
        out = np.empty(M * M * M, dtype=np.float32)
        self._ck(self.l.mib_debug_supergrid(self.h, C.c_int(medium), _fp(out), C.c_size_t(out.size)))
        return out.reshape(M, M, M)

    def sample_lanes(self, seed, spp, lane_begin, count):
        out = np.empty((count, self.raw_channels), np.float32)
        self._ck(self.l.mib_debug_sample_lanes(self.h, C.c_uint32(seed), C.c_uint32(spp), C.c_uint64(lane_begin),
                                               C.c_uint64(count), _fp(out)))
        return out

    NOALLOC_MAX_FLOATS = 1 << 20

    def sample_lanes_noalloc(self, seed, spp, lane_begin, count):
        """mib_sample_lanes_noalloc: sample_lanes() with its scratch in the library's static device array (no cudaMalloc);
        count * raw_channels <= NOALLOC_MAX_FLOATS (mi_b200.h)."""
        out = np.empty((count, self.raw_channels), np.float32)
        self._ck(self.l.mib_sample_lanes_noalloc(self.h, C.c_uint32(seed), C.c_uint32(spp), C.c_uint64(lane_begin),
                                                 C.c_uint64(count), _fp(out)))
        return out

    # -- adjoint --
    def render_backward(self, seed, spp, grad_in):
        g = np.ascontiguousarray(grad_in, dtype=np.float32)
        self._ck(self.l.mib_render_backward(self.h, C.c_uint32(seed), C.c_uint32(spp), _fp(g)))

    def render_1_backward(self, seed, spp, grad_in4):
        g = np.zeros(4, np.float32)
        gi = np.asarray(grad_in4, dtype=np.float32).reshape(-1)
        g[:gi.size] = gi
        self._ck(self.l.mib_render_1_backward(self.h, C.c_uint32(seed), C.c_uint32(spp), _fp(g)))

    def grad_get(self, slot, count):
        out = np.empty(count, np.float32)
        self._ck(self.l.mib_grad_get(self.h, C.c_int(slot), _fp(out), C.c_size_t(count)))
        return out

    def grad_dev(self, slot):
        p, n = C.c_void_p(), C.c_size_t()
        self._ck(self.l.mib_grad_dev(self.h, C.c_int(slot), C.byref(p), C.byref(n)))
        return p.value, n.value

    def grad_get_scalar(self, medium):
        out = np.zeros(4, np.float64)
        self._ck(self.l.mib_grad_get_scalar(self.h, C.c_int(medium), _fp(out)))
        return out

    def grad_scalar_dev(self):
        p, n = C.c_void_p(), C.c_size_t()
        self._ck(self.l.mib_grad_scalar_dev(self.h, C.byref(p), C.byref(n)))
        return p.value, n.value

    def grad_zero(self):
        self._ck(self.l.mib_grad_zero(self.h))

    def counters(self):
        c = A.Counters()
        self._ck(self.l.mib_counters_get(self.h, C.byref(c)))
        return c.as_dict()

    def stream(self):
        p = C.c_void_p()
        self._ck(self.l.mib_scene_stream(self.h, C.byref(p)))
        return p.value

    def close(self):
        if self.h:
            self.l.mib_scene_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def debug_math(op, x):
    """Run a device math function on an array (op: 0 sincos, 1 exp, 2 log, 3 cbrt, 4 asin)."""
    l = lib()
    x = np.ascontiguousarray(x, np.float32)
    out = np.empty(2 * x.size if op == 0 else x.size, np.float32)
    rc = l.mib_debug_math(C.c_int(op), _fp(x), _fp(out), C.c_size_t(x.size))
    if rc != 0:
        raise RuntimeError(l.mib_last_error().decode())
    return out


def debug_fresnel(op, x, eta):
    """Run a Fresnel block of csrc/dfresnel.cuh on an array (op: 0 fresnel, 1 fresnel_polarized, 2 specular_reflection,
    3 specular_transmission, 4 refract with x = [n, 6] rows of (wi, m))."""
    l = lib()
    x = np.ascontiguousarray(x, np.float32)
    n = x.size // (6 if op == 4 else 1)
    out = np.empty((n, {0: 4, 1: 7, 2: 16, 3: 16, 4: 3}[op]), np.float32)
    rc = l.mib_debug_fresnel(C.c_int(op), _fp(x), C.c_float(eta), _fp(out), C.c_size_t(n))
    if rc != 0:
        raise RuntimeError(l.mib_last_error().decode())
    return out


def debug_microfacet(op, x, type_=0, alpha_u=0.1, alpha_v=0.1):
    """Run a block of csrc/dmicrofacet.cuh on an array (op: 0 eval/smith_g1/pdf with x = [n, 6] rows of (wi, m); 1 visible-normal
    sampling with x = [n, 5] rows of (wi, sample2); 2 erf, 3 erfinv, 4 exp2 with x = [n])."""
    l = lib()
    x = np.ascontiguousarray(x, np.float32)
    n = x.size // {0: 6, 1: 5}.get(op, 1)
    out = np.empty((n, {0: 3, 1: 4}.get(op, 1)), np.float32)
    rc = l.mib_debug_microfacet(C.c_int(op), C.c_int(type_), C.c_float(alpha_u), C.c_float(alpha_v), _fp(x), _fp(out), C.c_size_t(n))
    if rc != 0:
        raise RuntimeError(l.mib_last_error().decode())
    return out


def fp32_peak(device=0, reps=5):
    """Measured FP32 FMA roof of the device in TFLOP/s (mib_debug_fp32_peak)."""
    l = lib()
    out = C.c_double()
    rc = l.mib_debug_fp32_peak(C.c_int(device), C.c_int(reps), C.byref(out))
    if rc != 0:
        raise RuntimeError(l.mib_last_error().decode())
    return out.value


def l2_peak(device=0, reps=3):
    """Measured L2 streaming-read roof of the device in GB/s (mib_debug_l2_peak)."""
    l = lib()
    out = C.c_double()
    rc = l.mib_debug_l2_peak(C.c_int(device), C.c_int(reps), C.byref(out))
    if rc != 0:
        raise RuntimeError(l.mib_last_error().decode())
    return out.value
