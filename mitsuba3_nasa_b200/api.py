"""The reference-facing Python surface for the volumetric hot path.

Mirrors src/python/python/util.py (traverse :292-350, SceneParameters :20-290,
render :512-625, render_1 :627-670) and the integrator methods of
src/python/python/ad/integrators/common.py (render, render_1, render_backward,
render_1_backward).  Computation happens only in libmi_b200.so (CUDA).
"""
import numpy as np

from . import _abi as A
from . import scene as _scene
from ._lib import DeviceScene

_VARIANTS = ['scalar_rgb', 'llvm_ad_rgb', 'llvm_ad_mono', 'llvm_ad_mono_polarized',
             'b200_ad_rgb', 'b200_ad_mono', 'b200_ad_mono_polarized']
_variant = None


def variants():
    return list(_VARIANTS)


def set_variant(*names):
    """mi.set_variant: accepts the reference's variant names; results are defined to equal llvm_ad_*."""
    global _variant
    for n in names:
        if n in _VARIANTS:
            _variant = n
            return
    raise ImportError("Requested an unsupported variant \"%s\". The following variants are available: %s."
                      % (", ".join(names), ", ".join(_VARIANTS)))


def variant():
    return _variant


def sample_tea_32(v0, v1, rounds=4):
    """include/mitsuba/core/random.h:76-91 (host copy, used for seed_grad: util.py:617-619)."""
    v0, v1, s = int(v0) & 0xffffffff, int(v1) & 0xffffffff, 0
    for _ in range(rounds):
        s = (s + 0x9e3779b9) & 0xffffffff
        v0 = (v0 + ((((v1 << 4) & 0xffffffff) + 0xa341316c) ^ ((v1 + s) & 0xffffffff) ^ ((v1 >> 5) + 0xc8013ea4))) & 0xffffffff
        v1 = (v1 + ((((v0 << 4) & 0xffffffff) + 0xad90777d) ^ ((v0 + s) & 0xffffffff) ^ ((v0 >> 5) + 0x7e95761e))) & 0xffffffff
    return v0, v1


class Integrator:
    """Integrator object bound to a scene (volpath, volpathaos, prbvolpath, prb_volpathaos)."""

    def __init__(self, scene):
        self._scene = scene

    def render(self, scene=None, sensor=0, seed=0, spp=0, develop=True, evaluate=True):
        return (scene or self._scene)._render(seed, spp, develop)

    def render_1(self, scene=None, sensor=0, seed=0, spp=0, develop=True, evaluate=True, thread_count=0):
        return (scene or self._scene)._render_1(seed, spp)

    def render_backward(self, scene, params, grad_in, sensor=0, seed=0, spp=0):
        (scene or self._scene)._render_backward(seed, spp, grad_in)

    def render_1_backward(self, scene, params, grad_in, sensor=0, seed=0, spp=0):
        (scene or self._scene)._render_1_backward(seed, spp, grad_in)


class Film:
    """The part of mi.Film the render tests use (src/films/hdrfilm.cpp): size(), bitmap() = the last developed image,
    write() = an OpenEXR file in the film's component_format, PIZ-compressed as Bitmap::write does."""

    def __init__(self, scene):
        self._scene = scene

    def size(self):
        s = self._scene.flat.desc.sensor
        return (int(s.film_width), int(s.film_height))

    def crop_size(self):
        s = self._scene.flat.desc.sensor
        return (int(s.crop_width), int(s.crop_height))

    def bitmap(self, raw=False):
        img = self._scene._last_image
        if img is None:
            raise RuntimeError("No storage allocated, was prepare() called first?")       # hdrfilm.cpp develop() without a render
        return img

    def write(self, path):
        from .exr import write_exr
        fmt = self._scene.flat.film_component_format
        if fmt == 'uint32':
            raise RuntimeError("Film.write: component_format 'uint32' is not supported by this build")
        write_exr(path, np.asarray(self.bitmap()), compression="piz", component_format=fmt)


class Sensor:
    """scene.sensors()[i]: film() plus read access to the flattened sensor record (crop_width, to_world, ...)."""

    def __init__(self, scene):
        self._scene = scene
        self._film = Film(scene)

    def film(self):
        return self._film

    def __getattr__(self, name):
        return getattr(self._scene.flat.desc.sensor, name)


class Scene:
    """mi.Scene stand-in: the flattened description + its device replica."""

    def __init__(self, flat, device=0):
        self.flat = flat
        self.device = DeviceScene(flat, device)
        self._integrator = Integrator(self)
        self._last_image = None
        self._sensor = Sensor(self)

    def integrator(self):
        return self._integrator

    def sensors(self):
        return [self._sensor]

    def _spp(self, spp):
        return int(spp) if spp else int(self.flat.sample_count)

    def _render(self, seed, spp, develop=True):
        img = self.device.render(int(seed), self._spp(spp), develop=develop)
        if develop:
            self._last_image = img
        return img

    def _render_1(self, seed, spp):
        out = self.device.render_1(int(seed), self._spp(spp))
        if self.flat.desc.spectrum == A.SPEC_MONO_POLARIZED:
            m = np.zeros((4, 4), np.float64)     # integrator.cpp:707-713: Stokes vector in column 0
            m[:, 0] = out
            return m
        return float(out[0])

    def _render_backward(self, seed, spp, grad_in):
        self.device.render_backward(int(seed), self._spp(spp), grad_in)

    def _render_1_backward(self, seed, spp, grad_in):
        g = np.asarray(grad_in, dtype=np.float64)
        if g.ndim == 2:
            g = g[:, 0]
        self.device.render_1_backward(int(seed), self._spp(spp), g)

    def counters(self):
        return self.device.counters()


def load_dict(d, parallel=True, device=0):
    """mi.load_dict (src/core/python/xml_v.cpp:112-146)."""
    if _variant is None:
        raise ImportError("set_variant() must be called before load_dict()")
    return Scene(_scene.load_dict(d, _variant), device)


def load_string(string, parallel=True, device=0, **kwargs):
    """mi.load_string (src/core/python/xml_v.cpp:88-109): an XML scene description; keyword arguments fill `$name` parameters."""
    if _variant is None:
        raise ImportError("set_variant() must be called before load_string()")
    from . import xmlscene
    return Scene(_scene.load_dict(xmlscene.xml_to_dict(string, None, **kwargs), _variant), device)


def load_file(path, update_scene=False, parallel=True, device=0, **kwargs):
    """mi.load_file (src/core/python/xml_v.cpp:64-85).  Relative `filename` properties and <include> tags are resolved against
    the directory of the scene file, as the reference's file resolver does."""
    if _variant is None:
        raise ImportError("set_variant() must be called before load_file()")
    import os
    from . import xmlscene
    if not os.path.exists(path):
        raise FileNotFoundError('"%s": file does not exist!' % path)
    with open(path, 'r') as f:
        d = xmlscene.xml_to_dict(f.read(), os.path.dirname(os.path.abspath(path)), **kwargs)
    return Scene(_scene.load_dict(d, _variant), device)


class ParamArray(np.ndarray):
    """What params[key] returns for array-valued parameters: a numpy array that remembers where it came from, so that
    dr.enable_grad(params[key]) / dr.grad(params[key]) (dr.py) can find the parameter, as the Dr.Jit arrays of the
    reference carry their own AD index."""
    _mi_params = None
    _mi_key = None

    def __array_finalize__(self, obj):
        # results of arithmetic are plain data: only the object handed out by params[key] names a parameter
        self._mi_params = None
        self._mi_key = None


class ParamScalar(float):
    """Scalar-valued parameter with the same back-reference."""
    _mi_params = None
    _mi_key = None


def _tag(value, params, key):
    if isinstance(value, np.ndarray):
        out = value.view(ParamArray)
    else:
        out = ParamScalar(value)
    out._mi_params, out._mi_key = params, key
    return out


class SceneParameters:
    """Dictionary-like parameter map (util.py:20-290): params[key] = value; params.update()."""

    def __init__(self, scene):
        self.scene = scene
        self._dirty = set()
        self._torch = {}           # key -> torch tensor assigned by the user (leaf tensors make mi.render differentiable)

    def keys(self):
        return sorted(self.scene.flat.params.keys())

    def __contains__(self, k):
        return k in self.scene.flat.params

    def __iter__(self):
        return iter(self.keys())

    def __len__(self):
        return len(self.scene.flat.params)

    def items(self):
        return [(k, self[k]) for k in self.keys()]

    def __getitem__(self, key):
        return _tag(self._value(key), self, key)

    def _value(self, key):
        if key not in self.scene.flat.params:
            raise KeyError(key)
        acc = self.scene.flat.params[key]
        f = self.scene.flat
        kind = acc['kind']
        if kind == 'buffer':
            return f.buffers[acc['slot']].reshape(f.buffer_shapes[acc['slot']])
        if kind == 'volume_const':
            v = acc['vol']
            return np.array(v.value[:]) if f.spectrum == A.SPEC_RGB else float(v.value[0])
        if kind == 'phase_g':
            return float(acc['phase'].g)
        if kind == 'medium_scale':
            return float(acc['medium'].scale)
        if kind == 'bsdf_reflectance':
            b = acc['bsdf']
            return np.array(b.reflectance[:]) if f.spectrum == A.SPEC_RGB else float(b.reflectance[0])
        if kind == 'emitter_radiance':
            e = acc['emitter']
            return np.array(e.radiance[:]) if f.spectrum == A.SPEC_RGB else float(e.radiance[0])
        if kind == 'bsdf_eta':
            return float(acc['bsdf'].eta)
        if kind == 'bsdf_alpha':
            return float(acc['bsdf'].alpha_u)
        raise KeyError(key)

    def __setitem__(self, key, value):
        if key not in self.scene.flat.params:
            raise KeyError(key)
        if hasattr(value, 'detach') and hasattr(value, 'requires_grad'):     # torch tensor
            self._torch[key] = value
            value = value.detach().cpu().numpy()
        else:
            self._torch.pop(key, None)
        self._assign(key, value)

    def enable_grad(self, key):
        """dr.enable_grad(params[key]): returns the leaf tensor whose .grad receives d(loss)/d(params[key])."""
        import torch
        from . import ad
        if ad.grad_source(self.scene, key) is None:
            raise RuntimeError("parameter '%s' is not differentiable in this backend (sigma_t / albedo of media are)" % key)
        t = torch.tensor(np.array(self._value(key), dtype=np.float32), requires_grad=True)
        self._torch[key] = t
        self._assign(key, t.detach().numpy())
        return t

    def attached(self):
        return {k: t for k, t in self._torch.items() if getattr(t, 'requires_grad', False)}

    def _assign(self, key, value):
        acc = self.scene.flat.params[key]
        f = self.scene.flat
        kind = acc['kind']
        if kind == 'buffer':
            arr = np.ascontiguousarray(np.asarray(value, dtype=np.float32))
            slot = acc['slot']
            if arr.size != f.buffers[slot].size:
                raise ValueError("set: size mismatch for '%s' (%d vs %d); resizing grids is not supported" %
                                 (key, arr.size, f.buffers[slot].size))
            f.buffers[slot] = arr.reshape(-1).copy()
        elif kind in ('volume_const', 'bsdf_reflectance', 'emitter_radiance'):
            tgt = {'volume_const': 'vol', 'bsdf_reflectance': 'bsdf', 'emitter_radiance': 'emitter'}[kind]
            field = {'volume_const': 'value', 'bsdf_reflectance': 'reflectance', 'emitter_radiance': 'radiance'}[kind]
            col = np.broadcast_to(np.asarray(value, dtype=np.float64), (3,))
            getattr(acc[tgt], field)[:] = [float(x) for x in col]
            if kind == 'volume_const':
                # (an attached RGB constant keeps three channels so that its gradient comes back per channel)
                acc['vol'].channels = 3 if (f.spectrum == A.SPEC_RGB and (len(set(col.tolist())) > 1 or key in self._torch)) else 1
        elif kind == 'bsdf_eta':
            eta = np.float32(value)
            acc['bsdf'].eta = float(eta)
            acc['bsdf'].inv_eta = float(np.float32(1.0) / eta)                   # roughdielectric.cpp:245
        elif kind == 'bsdf_alpha':
            acc['bsdf'].alpha_u = acc['bsdf'].alpha_v = float(np.float32(value))
        elif kind == 'phase_g':
            acc['phase'].g = float(value)
        elif kind == 'medium_scale':
            acc['medium'].scale = float(value)
        self._dirty.add(key)

    def update(self, values=None):
        """Push modified parameters to the device and recompute derived data (parameters_changed)."""
        if values:
            for k, v in values.items():
                self[k] = v
        f, dev = self.scene.flat, self.scene.device
        for key in self._dirty:
            acc = f.params[key]
            if acc['kind'] == 'buffer':
                dev.set_buffer(acc['slot'], f.buffers[acc['slot']])
        dev.set_desc(f.desc)
        dev.commit()
        done = sorted(self._dirty)
        self._dirty.clear()
        return done


def traverse(scene):
    """mi.traverse (util.py:292-350)."""
    return SceneParameters(scene)


def _differentiable(scene, params, seed, seed_grad, spp, spp_grad, one):
    from . import ad
    it = scene.flat.desc.integrator.type
    if it not in (A.INTEG_PRBVOLPATH, A.INTEG_PRBVOLPATHAOS):
        raise RuntimeError("differentiable rendering needs a prbvolpath / prb_volpathaos integrator")
    att = params.attached()
    keys = sorted(att)
    if seed_grad == 0:
        seed_grad = sample_tea_32(seed, 1)[0]                         # util.py:617-619
    elif seed_grad == seed:                                           # util.py:620-622
        raise Exception('The primal and differential seed should be different to ensure unbiased gradient computation!')
    return ad.render_op().apply(scene, params, keys, int(seed), int(seed_grad), int(spp), int(spp_grad), one, *[att[k] for k in keys])


def render(scene, params=None, sensor=0, integrator=None, seed=0, seed_grad=0, spp=0, spp_grad=0):
    """mi.render (util.py:512-625).  With leaf tensors attached to `params` (SceneParameters.enable_grad, or
    params[key] = tensor) the result is a torch tensor wired to them (ad.py); otherwise a numpy array."""
    if params is not None and isinstance(params, SceneParameters) and params.attached():
        return _differentiable(scene, params, seed, seed_grad, spp, spp_grad, False)
    return scene.integrator().render(scene, sensor, seed, spp, develop=True, evaluate=False)


def render_1(scene, params=None, sensor=0, integrator=None, seed=0, seed_grad=0, spp=0, spp_grad=0):
    """mi.render_1 (util.py:627-670): one Spectrum / Mueller matrix (Stokes vector in column 0)."""
    if params is not None and isinstance(params, SceneParameters) and params.attached():
        return _differentiable(scene, params, seed, seed_grad, spp, spp_grad, True)
    return scene.integrator().render_1(scene, sensor, seed, spp)
