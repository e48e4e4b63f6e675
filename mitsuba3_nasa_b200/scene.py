"""Host-side mirror of the reference's plugin tree for the volumetric hot path.

`load_dict` builds a tree of `Plugin` nodes from nested dicts exactly as
`mi.load_dict` does (src/core/python/xml_v.cpp:112-146): `'type'` keys select
plugins, `{'type': 'ref', 'id': ...}` reuses an object, floats / rgb dicts are
auto-wrapped into `uniform` / `srgb` textures and `constvolume`s
(include/mitsuba/core/properties.h:341-386).  Plugin names, parameter names and
defaults are those of the reference (citations next to each plugin below).
`Scene.flatten()` turns the tree into the POD `mib_scene_desc` + float buffers
consumed by the C ABI (include/mi_b200.h).
"""
import os

import numpy as np

from . import _abi as A, plastic
from .transforms import ScalarTransform4f, coordinate_system

RAY_EPSILON = 1500.0 * 2.0 ** -24   # include/mitsuba/core/math.h:17-23


# include/mitsuba/render/ior.h:22-48 (named indices of refraction)
_IOR = {"vacuum": 1.0, "helium": 1.000036, "hydrogen": 1.000132, "air": 1.000277, "carbon dioxide": 1.00045, "water": 1.3330,
        "acetone": 1.36, "ethanol": 1.361, "carbon tetrachloride": 1.461, "glycerol": 1.4729, "benzene": 1.501,
        "silicone oil": 1.52045, "bromine": 1.661, "water ice": 1.31, "fused quartz": 1.458, "pyrex": 1.470,
        "acrylic glass": 1.49, "polypropylene": 1.49, "bk7": 1.5046, "sodium chloride": 1.544, "amber": 1.55, "pet": 1.5750,
        "diamond": 2.419}


def _lookup_ior(v):
    if isinstance(v, str):
        if v.lower() not in _IOR:
            raise PluginError("Unable to find an IOR value for \"%s\"! Valid choices are: %s" % (v, ", ".join(_IOR)))
        return _IOR[v.lower()]
    return float(v)


class PluginError(RuntimeError):
    pass


def _mean32(values):
    # dr::mean of a Color in float32 (sum left to right, times 1 / n)
    acc = np.float32(values[0])
    for v in values[1:]:
        acc = np.float32(acc + np.float32(v))
    return np.float32(acc * (np.float32(1.0) / np.float32(len(values))))


def _luminance(rgb):
    # include/mitsuba/core/spectrum.h:431-434
    return float(rgb[0] * 0.212671 + rgb[1] * 0.715160 + rgb[2] * 0.072169)


# directories searched for relative file names of `obj` shapes (the role of the reference's FileResolver; the XML front end
# resolves file names itself, xmlscene.py)
SEARCH_PATHS = []


def _parse_fov(pr, aspect):
    """parse_fov (src/render/sensor.cpp:149-203): the horizontal field of view in degrees from 'fov' + 'fov_axis' or from
    'focal_length' (default "50mm": a 35 mm film's DIAGONAL field of view)."""
    import math
    if pr.has('fov') and pr.has('focal_length'):
        raise PluginError("Please specify either a focal length ('focal_length') or a field of view ('fov')!")
    if pr.has('fov'):
        fov = float(pr.get('fov'))
        axis = str(pr.get('fov_axis', 'x')).lower()
        if axis == 'smaller':
            axis = 'y' if aspect > 1 else 'x'
        elif axis == 'larger':
            axis = 'x' if aspect > 1 else 'y'
    else:
        f = str(pr.get('focal_length', '50mm'))
        if f.endswith('mm'):
            f = f[:-2]
        try:
            value = float(f)
        except ValueError:
            raise PluginError("Could not parse the focal length (must be of the form <x>mm, where <x> is a positive integer)!")
        fov = 2.0 * math.degrees(math.atan(math.sqrt(36.0 * 36 + 24 * 24) / (2.0 * value)))
        axis = 'diagonal'
    if axis == 'x':
        result = fov
    elif axis == 'y':
        result = math.degrees(2.0 * math.atan(math.tan(0.5 * math.radians(fov)) * aspect))
    elif axis == 'diagonal':
        diagonal = 2.0 * math.tan(0.5 * math.radians(fov))
        width = diagonal / math.sqrt(1.0 + 1.0 / (aspect * aspect))
        result = math.degrees(2.0 * math.atan(width * 0.5))
    else:
        raise PluginError("The 'fov_axis' parameter must be set to one of 'smaller', 'larger', 'diagonal', 'x', or 'y'!")
    if result <= 0.0 or result >= 180.0:
        raise PluginError("The horizontal field of view must be in the range [0, 180]!")
    return result


class Plugin:
    """One instantiated plugin.  `props` holds the parameters (defaults applied)."""
    CATEGORY = {
        'volpath': 'integrator', 'volpathaos': 'integrator', 'prbvolpath': 'integrator', 'prb_volpathaos': 'integrator',
        'moment': 'integrator', 'volpathmis': 'integrator',
        'homogeneous': 'medium', 'heterogeneous': 'medium',
        'gridvolume': 'volume', 'constvolume': 'volume',
        'hg': 'phase', 'rayleigh': 'phase', 'isotropic': 'phase', 'tabphase': 'phase', 'tabphase_polarized': 'phase',
        'mie': 'phase',
        'lognormal': 'sizedistr', 'gamma': 'sizedistr', 'powerlaw': 'sizedistr', 'monodisperse': 'sizedistr',
        'independent': 'sampler', 'hdrfilm': 'film',
        'box': 'rfilter', 'gaussian': 'rfilter', 'tent': 'rfilter',
        'perspective': 'sensor', 'distant': 'sensor', 'radiancemeter': 'sensor',
        'constant': 'emitter', 'directional': 'emitter', 'area': 'emitter', 'envmap': 'emitter', 'spot': 'emitter',
        'cube': 'shape', 'rectangle': 'shape', 'sphere': 'shape', 'obj': 'shape',
        'null': 'bsdf', 'diffuse': 'bsdf', 'dielectric': 'bsdf', 'roughdielectric': 'bsdf', 'thindielectric': 'bsdf', 'plastic': 'bsdf', 'roughplastic': 'bsdf',
        'uniform': 'texture', 'srgb': 'texture', 'rgb': 'texture', 'd65': 'texture', 'checkerboard': 'texture',
        'scene': 'scene',
    }

    def __init__(self, type_, props, children, id_=None):
        self.type = type_
        self.category = Plugin.CATEGORY[type_]
        self.props = props          # plain values
        self.children = children    # name -> Plugin
        self.id = id_

    def __repr__(self):
        return "%s[%s]" % (self.type, ", ".join("%s=%r" % kv for kv in self.props.items()))


def _properties_less(a, b):
    """Properties' SortKey (src/core/properties.cpp:48-71): the order in which a plugin sees its child objects -- byte order, except
    that a trailing run of digits after the common prefix compares as a number ("_arg_2" < "_arg_10")."""
    i = 0
    while i < len(a) and i < len(b) and a[i] == b[i]:
        i += 1
    while i > 0 and a[i - 1].isdigit():
        i -= 1
    ta, tb = a[i:], b[i:]
    if ta[:1].isdigit() and tb[:1].isdigit() and ta.isdigit() and tb.isdigit():
        return int(ta) < int(tb)
    return ta < tb


def properties_order(names):
    import functools
    return sorted(names, key=functools.cmp_to_key(lambda a, b: -1 if _properties_less(a, b) else (1 if _properties_less(b, a) else 0)))


class _Loader:
    def __init__(self, variant):
        self.variant = variant
        self.by_id = {}
        self.root_dict = None

    def is_rgb(self):
        return 'rgb' in self.variant

    # -- generic dict -> Plugin --
    def load(self, d, name=None):
        if not isinstance(d, dict) or 'type' not in d:
            raise PluginError("load_dict(): expected a dict with a 'type' key for %r" % (name,))
        t = d['type']
        if t == 'ref':
            if d['id'] not in self.by_id:
                # forward reference to a top-level object declared later in the dict
                if self.root_dict is not None and isinstance(self.root_dict.get(d['id']), dict):
                    self.by_id[d['id']] = self.load(self.root_dict[d['id']], d['id'])
                else:
                    raise PluginError("load_dict(): reference to unknown id '%s'" % d['id'])
            return self.by_id[d['id']]
        if t not in Plugin.CATEGORY:
            raise PluginError("Plugin \"%s\" not found! (only the plugins of the volumetric hot path exist "
                              "in this build, see DESIGN.md)" % t)
        props, children = {}, {}
        if t == 'scene' and self.root_dict is None:
            self.root_dict = d
        for k, v in d.items():
            if k in ('type', 'id', '_document_order'):
                continue
            if isinstance(v, dict):
                if t == 'scene' and k in self.by_id:
                    children[k] = self.by_id[k]     # already instantiated through a forward reference
                else:
                    children[k] = self.load(v, k)
            else:
                props[k] = v
        p = Plugin(t, props, children, d.get('id', name))
        p.document_order = bool(d.get('_document_order', False))     # xmlscene.py: children already in the parser's `_arg_N` order
        if p.id is not None and p.id not in self.by_id:
            self.by_id[p.id] = p
        return p


# --------------------------------------------------------------------------
# Property access with reference semantics: defaults + "unqueried property" errors
# --------------------------------------------------------------------------
class _Props:
    def __init__(self, plugin):
        self.p = plugin
        self.queried = set()

    def get(self, name, default=None, required=False):
        self.queried.add(name)
        if name in self.p.props:
            return self.p.props[name]
        if required:
            raise PluginError("Property \"%s\" has not been specified! (plugin %s)" % (name, self.p.type))
        return default

    def has(self, name):
        return name in self.p.props or name in self.p.children

    def child(self, name):
        self.queried.add(name)
        return self.p.children.get(name)

    def children_of(self, category):
        out = []
        for k, c in self.p.children.items():
            if c.category == category:
                self.queried.add(k)
                out.append((k, c))
        return out

    def finish(self):
        extra = (set(self.p.props) | set(self.p.children)) - self.queried
        if extra:
            raise PluginError("Error while loading \"%s\": unreferenced propert%s %s" %
                              (self.p.type, "y" if len(extra) == 1 else "ies", sorted(extra)))


def _to_transform(v):
    if v is None:
        return ScalarTransform4f()
    return v if isinstance(v, ScalarTransform4f) else ScalarTransform4f(v)


class Scene:
    """Flattened scene + parameter map.  Mirrors the parts of mi.Scene the hot path needs."""

    def __init__(self, root, variant):
        self.root = root
        self.variant = variant
        self.spectrum = A.SPEC_MONO_POLARIZED if 'polarized' in variant else (A.SPEC_RGB if 'rgb' in variant else A.SPEC_MONO)
        self.desc = A.SceneDesc()
        self.buffers = {}      # slot -> np.float32 array (flat)
        self.buffer_shapes = {}
        self.params = {}       # dotted key -> accessor dict
        self._n_buffers = 0
        self._media_index = {}
        self._phase_index = {}
        self._bsdf_index = {}
        self.film_component_format = 'float16'
        self._flatten()

    # ------------------------------------------------------------------
    def _color(self, v, what):
        """float | [r,g,b] | texture Plugin -> 3 floats for the current variant."""
        if isinstance(v, Plugin):
            pr = _Props(v)
            if v.type in ('rgb', 'srgb'):
                val = pr.get('value', required=True)
                pr.finish()
                val = np.broadcast_to(np.asarray(val, dtype=np.float64), (3,))
                if self.spectrum == A.SPEC_RGB:
                    return [float(x) for x in val]
                lum = _luminance(val)
                return [lum] * 3
            if v.type == 'uniform':
                val = float(pr.get('value', 1.0))
                pr.finish()
                return [val] * 3
            if v.type == 'd65':
                # mono / rgb variants: D65 collapses to its scale factor (texture_d65, properties.h)
                val = float(pr.get('scale', 1.0))
                pr.finish()
                return [val] * 3
            raise PluginError("%s: unsupported texture plugin '%s'" % (what, v.type))
        if isinstance(v, (int, float)):
            return [float(v)] * 3
        val = np.asarray(v, dtype=np.float64)
        if self.spectrum == A.SPEC_RGB:
            return [float(x) for x in np.broadcast_to(val, (3,))]
        return [_luminance(np.broadcast_to(val, (3,)))] * 3

    def _new_buffer(self, arr, shape=None):
        slot = self._n_buffers
        if slot >= A.MIB_MAX_BUFFERS:
            raise PluginError("too many data buffers (max %d)" % A.MIB_MAX_BUFFERS)
        self._n_buffers += 1
        a = np.ascontiguousarray(arr, dtype=np.float32)
        self.buffers[slot] = a.reshape(-1).copy()
        self.buffer_shapes[slot] = tuple(a.shape if shape is None else shape)
        return slot

    # -- volumes: src/volumes/const.cpp:61-84, src/volumes/grid.cpp:156-299 --
    def _volume(self, out, v, default, key):
        out.buffer = -1
        out.max_value = 0.0
        ident = np.eye(4, dtype=np.float32)
        out.to_local[:] = ident.reshape(-1).tolist()
        if v is None:
            v = default
        if isinstance(v, Plugin) and v.type == 'gridvolume':
            pr = _Props(v)
            data = pr.get('data')
            grid = pr.get('grid')
            fname = pr.get('filename')
            file_bbox = None
            if data is None and grid is not None:                 # a VolumeGrid object (grid.cpp:170-176)
                data = np.asarray(grid)
                if hasattr(grid, 'bbox'):
                    file_bbox = grid.bbox()
            if data is None and fname is not None:
                from .volumegrid import read_vol
                data, file_bbox = read_vol(fname)
            if data is None:
                raise PluginError("gridvolume: one of 'filename', 'grid' or 'data' must be specified")
            data = np.asarray(data, dtype=np.float32)
            if data.ndim == 3:
                data = data[..., None]
            if data.ndim != 4 or data.shape[3] not in (1, 3):
                raise PluginError("gridvolume: data must be [z, y, x, c] with c in {1, 3}")
            ftype = pr.get('filter_type', 'trilinear')
            wmode = pr.get('wrap_mode', 'clamp')
            pr.get('raw', False); pr.get('accel', True)
            use_grid_bbox = bool(pr.get('use_grid_bbox', False))
            fixed_max = pr.get('max_value')
            to_world = _to_transform(pr.get('to_world'))
            if use_grid_bbox:                                     # grid.cpp:288-293, volumegrid.h:26-30
                if file_bbox is None:
                    raise PluginError("use_grid_bbox is unsupported with tensor input and requires a volume grid")
                bmin, bmax = np.asarray(file_bbox[0], np.float64), np.asarray(file_bbox[1], np.float64)
                to_world = to_world @ ScalarTransform4f.translate(bmin.tolist()) @ ScalarTransform4f.scale((bmax - bmin).tolist())
            pr.finish()
            if ftype not in ('nearest', 'trilinear'):
                raise PluginError("Invalid filter type \"%s\", must be one of: \"nearest\" or \"trilinear\"!" % ftype)
            if wmode not in ('repeat', 'mirror', 'clamp'):
                raise PluginError("Invalid wrap mode \"%s\", must be one of: \"repeat\", \"mirror\", or \"clamp\"!" % wmode)
            out.type = A.VOLUME_GRID
            out.channels = data.shape[3]
            out.res[:] = [data.shape[2], data.shape[1], data.shape[0]]
            out.filter = A.FILTER_NEAREST if ftype == 'nearest' else A.FILTER_TRILINEAR
            out.wrap = {'repeat': A.WRAP_REPEAT, 'mirror': A.WRAP_MIRROR, 'clamp': A.WRAP_CLAMP}[wmode]
            out.buffer = self._new_buffer(data)
            to_local = to_world.inverse()
            out.to_local[:] = to_local.f32().reshape(-1).tolist()
            # update_bbox (include/mitsuba/render/volume.h:99-110)
            corners = np.array([to_world @ [x, y, z] for x in (0, 1) for y in (0, 1) for z in (0, 1)])
            out.bbox_min[:] = corners.min(0).astype(np.float32).tolist()
            out.bbox_max[:] = corners.max(0).astype(np.float32).tolist()
            out.fixed_max = 0
            if fixed_max is not None:                             # grid.cpp:296-297: m_fixed_max, used as the majorant as it is
                out.max_value = float(fixed_max)
                out.fixed_max = 1
            self.params[key + '.data'] = {'kind': 'buffer', 'slot': out.buffer}
            return
        # constvolume (explicit plugin, bare float or texture)
        if isinstance(v, Plugin) and v.type == 'constvolume':
            pr = _Props(v)
            val = pr.child('value') or pr.get('value', 1.0)
            pr.finish()
            v = val
        col = self._color(v, key)
        out.type = A.VOLUME_CONST
        out.channels = 3 if (self.spectrum == A.SPEC_RGB and len(set(col)) > 1) else 1
        out.value[:] = col
        out.res[:] = [1, 1, 1]
        self.params[key + '.value.value'] = {'kind': 'volume_const', 'vol': out}

    # -- phase functions --
    def _phase(self, p, key):
        if id(p) in self._phase_index:
            return self._phase_index[id(p)]
        i = self.desc.n_phases
        if i >= A.MIB_MAX_PHASES:
            raise PluginError("too many phase functions")
        self.desc.n_phases += 1
        self._phase_index[id(p)] = i
        out = self.desc.phases[i]
        out.buf_nodes = out.buf_m11 = out.buf_m12 = out.buf_m33 = out.buf_m34 = -1
        pr = _Props(p)
        if p.type == 'isotropic':                      # src/phase/isotropic.cpp
            out.type = A.PHASE_ISOTROPIC
        elif p.type == 'hg':                           # src/phase/hg.cpp:50-62
            g = float(pr.get('g', 0.8))
            if g >= 1.0 or g <= -1.0:
                raise PluginError("The asymmetry parameter must lie in the interval (-1, 1)!")
            out.type = A.PHASE_HG
            out.g = g
            self.params[key + '.g'] = {'kind': 'phase_g', 'phase': out}
        elif p.type == 'rayleigh':                     # src/phase/rayleigh.cpp:41-48
            rho = float(pr.get('depolarization', 0.0))
            if rho < 0.0 or rho >= 1.0:
                raise PluginError("Depolarization factor must be between 0 and 1!")
            out.type = A.PHASE_RAYLEIGH
            out.depolarization = rho
        elif p.type in ('tabphase', 'tabphase_polarized'):   # src/phase/tabphase.cpp:44-79, tabphase_polarized.cpp:48-112
            def parse(name):
                s = pr.get(name, required=True)
                if isinstance(s, str):
                    return np.array([float(t) for t in s.replace(',', ' ').split()], dtype=np.float32)
                return np.asarray(s, dtype=np.float32)
            cost = parse('cost')
            names = ['values'] if p.type == 'tabphase' else ['m11', 'm12', 'm33', 'm34']
            tabs = [parse(n) for n in names]
            for n, t in zip(names, tabs):
                if t.size != cost.size:
                    raise PluginError("%s: 'cost' and '%s' parameters must have the same size!" % (p.type, n))
            out.type = A.PHASE_TABPHASE if p.type == 'tabphase' else A.PHASE_TABPHASE_POLARIZED
            out.n_nodes = cost.size
            out.buf_nodes = self._new_buffer(cost)
            out.buf_m11 = self._new_buffer(tabs[0])
            self.params[key + ('.values' if p.type == 'tabphase' else '.m11')] = {'kind': 'buffer', 'slot': out.buf_m11}
            if p.type == 'tabphase_polarized':
                out.buf_m12 = self._new_buffer(tabs[1])
                out.buf_m33 = self._new_buffer(tabs[2])
                out.buf_m34 = self._new_buffer(tabs[3])
                for n, s in zip(names[1:], (out.buf_m12, out.buf_m33, out.buf_m34)):
                    self.params[key + '.' + n] = {'kind': 'buffer', 'slot': s}
        elif p.type == 'mie':
            # src/phase/mie.cpp:251-257: mie is tabulated before rendering
            if self.spectrum == A.SPEC_RGB:
                raise PluginError("Mie phase function may only be used in monochromatic or spectral mode!")   # mie.cpp:67-68
            from .mie import tabulate_mie_plugin
            cost, m11, m12, m33, m34 = tabulate_mie_plugin(p, pr)
            out.type = A.PHASE_TABPHASE_POLARIZED
            out.n_nodes = cost.size
            out.buf_nodes = self._new_buffer(cost)
            out.buf_m11 = self._new_buffer(m11)
            out.buf_m12 = self._new_buffer(m12)
            out.buf_m33 = self._new_buffer(m33)
            out.buf_m34 = self._new_buffer(m34)
        else:
            raise PluginError("unsupported phase function '%s'" % p.type)
        pr.finish()
        return i

    # -- media: src/media/homogeneous.cpp:137-147, heterogeneous.cpp:155-167, src/render/medium.cpp:12-32 --
    def _medium(self, m, key):
        if id(m) in self._media_index:
            return self._media_index[id(m)]
        i = self.desc.n_media
        if i >= A.MIB_MAX_MEDIA:
            raise PluginError("too many media")
        self.desc.n_media += 1
        self._media_index[id(m)] = i
        if m.id is not None and m.id in self.root.children and self.root.children[m.id] is m:
            key = m.id
        out = self.desc.media[i]
        pr = _Props(m)
        out.type = A.MEDIUM_HOMOGENEOUS if m.type == 'homogeneous' else A.MEDIUM_HETEROGENEOUS
        phases = pr.children_of('phase')
        if len(phases) > 1:
            raise PluginError("Only a single phase function can be specified per medium")
        if phases:
            out.phase = self._phase(phases[0][1], key + '.phase_function')
        else:
            out.phase = self._phase(Plugin('isotropic', {}, {}), key + '.phase_function')
        out.sample_emitters = int(bool(pr.get('sample_emitters', True)))
        out.has_spectral_extinction = int(bool(pr.get('has_spectral_extinction', True)))
        out.scale = float(pr.get('scale', 1.0))
        self._volume(out.albedo, pr.child('albedo') or pr.get('albedo'), 0.75, key + '.albedo')
        self._volume(out.sigma_t, pr.child('sigma_t') or pr.get('sigma_t'), 1.0, key + '.sigma_t')
        if out.type == A.MEDIUM_HOMOGENEOUS and out.sigma_t.type != A.VOLUME_CONST:
            raise PluginError("homogeneous medium: sigma_t must be constant in this build")
        self.params[key + '.scale'] = {'kind': 'medium_scale', 'medium': out}
        pr.finish()
        return i

    # -- bsdfs: src/bsdfs/null.cpp, diffuse.cpp --
    def _bsdf(self, b, key, emitter=False):
        if b is not None and id(b) in self._bsdf_index:
            return self._bsdf_index[id(b)]
        i = self.desc.n_bsdfs
        if i >= A.MIB_MAX_BSDFS:
            raise PluginError("too many BSDFs")
        self.desc.n_bsdfs += 1
        out = self.desc.bsdfs[i]
        if b is None:                     # default BSDF of a shape: diffuse, black on a shape that emits (shape.cpp:64-70)
            out.type = A.BSDF_DIFFUSE
            out.reflectance[:] = [0.0 if emitter else 0.5] * 3
            return i
        self._bsdf_index[id(b)] = i
        pr = _Props(b)
        if b.type == 'null':
            out.type = A.BSDF_NULL
        elif b.type == 'diffuse':
            out.type = A.BSDF_DIFFUSE
            refl = pr.child('reflectance')
            if refl is not None and refl.type == 'checkerboard':
                # src/textures/checkerboard.cpp:58-84
                tp = _Props(refl)
                out.texture = A.TEXTURE_CHECKERBOARD
                out.color0[:] = self._color(tp.child('color0') or tp.get('color0', 0.4), key)
                out.color1[:] = self._color(tp.child('color1') or tp.get('color1', 0.2), key)
                to_uv = tp.get('to_uv')
                m = np.eye(3) if to_uv is None else np.asarray(getattr(to_uv, 'matrix', to_uv), dtype=np.float64)
                if m.shape == (4, 4):                  # the XML parser reads every <transform> as 4 x 4: keep the 2D affine part
                    m = np.array([[m[0, 0], m[0, 1], m[0, 3]], [m[1, 0], m[1, 1], m[1, 3]], [0, 0, 1.0]])
                out.to_uv[:] = m[:2, :].astype(np.float32).reshape(-1).tolist()
                out.reflectance[:] = [0.5 * (a + b) for a, b in zip(out.color0[:], out.color1[:])]      # mean()
                tp.finish()
            else:
                out.reflectance[:] = self._color(refl or pr.get('reflectance', 0.5), key)
                self.params[key + '.reflectance.value'] = {'kind': 'bsdf_reflectance', 'bsdf': out}
        elif b.type == 'thindielectric':
            # src/bsdfs/thindielectric.cpp:100-130
            int_ior = _lookup_ior(pr.get('int_ior', 'bk7'))
            ext_ior = _lookup_ior(pr.get('ext_ior', 'air'))
            if int_ior < 0 or ext_ior < 0:
                raise PluginError("The interior and exterior indices of refraction must be positive!")
            out.type = A.BSDF_THINDIELECTRIC
            out.eta = float(np.float32(int_ior) / np.float32(ext_ior))
            for name in ('specular_reflectance', 'specular_transmittance'):
                v = pr.child(name) or pr.get(name)
                if v is not None and any(abs(c - 1.0) > 0 for c in self._color(v, key)):
                    raise PluginError("%s: only %s = 1 is supported" % (b.type, name))
        elif b.type in ('dielectric', 'roughdielectric'):
            # src/bsdfs/dielectric.cpp:210-222, roughdielectric.cpp:166-218 (the ocean interface of the AOS scenes)
            int_ior = _lookup_ior(pr.get('int_ior', 'bk7'))
            ext_ior = _lookup_ior(pr.get('ext_ior', 'air'))
            if int_ior < 0 or ext_ior < 0 or (b.type == 'roughdielectric' and int_ior == ext_ior):
                raise PluginError("The interior and exterior indices of refraction must be positive and differ!")
            out.eta = float(np.float32(int_ior) / np.float32(ext_ior))
            out.inv_eta = float(np.float32(ext_ior) / np.float32(int_ior)) if b.type == 'dielectric' else float(np.float32(1.0) / np.float32(out.eta))
            for name in ('specular_reflectance', 'specular_transmittance'):
                v = pr.child(name) or pr.get(name)
                if v is not None and any(abs(c - 1.0) > 0 for c in self._color(v, key)):
                    raise PluginError("%s: only %s = 1 is supported" % (b.type, name))
            if b.type == 'dielectric':
                out.type = A.BSDF_DIELECTRIC
            else:
                out.type = A.BSDF_ROUGHDIELECTRIC
                distr = str(pr.get('distribution', 'beckmann')).lower()
                if distr not in ('beckmann', 'ggx'):
                    raise PluginError("Specified an invalid distribution \"%s\", must be \"beckmann\" or \"ggx\"!" % distr)
                out.distribution = A.MICROFACET_BECKMANN if distr == 'beckmann' else A.MICROFACET_GGX
                out.sample_visible = int(bool(pr.get('sample_visible', True)))
                if pr.has('alpha_sq'):
                    if pr.has('alpha'):
                        raise PluginError("Microfacet model: please specify either 'alpha' or 'alpha_sq'.")
                    au = av = float(np.sqrt(np.float32(pr.get('alpha_sq'))))
                elif pr.has('alpha_u') or pr.has('alpha_v'):
                    if not (pr.has('alpha_u') and pr.has('alpha_v')):
                        raise PluginError("Microfacet model: both 'alpha_u' and 'alpha_v' must be specified.")
                    if pr.has('alpha'):
                        raise PluginError("Microfacet model: please specifyeither 'alpha' or 'alpha_u'/'alpha_v'.")
                    au, av = float(pr.get('alpha_u')), float(pr.get('alpha_v'))
                else:
                    au = av = float(pr.get('alpha', 0.1))
                if not out.sample_visible and au != av:
                    raise PluginError("roughdielectric: sample_visible = false with an anisotropic roughness is not implemented")
                out.alpha_u, out.alpha_v = au, av
                # roughdielectric.cpp:224-236 traverse(): `eta`, and `alpha` (a uniform texture: key `alpha.value`) when isotropic.  Not
                # differentiable here (the adjoint of the interface is not built); params.update() re-commits the scene, which is what
                # the finite-difference side of the reference's adjoint test does (test_ad_volpathaos.py:504-529)
                self.params[key + '.eta'] = {'kind': 'bsdf_eta', 'bsdf': out}
                if au == av:
                    self.params[key + '.alpha.value'] = {'kind': 'bsdf_alpha', 'bsdf': out}
        elif b.type in ('plastic', 'roughplastic'):
            # src/bsdfs/plastic.cpp:151-200, src/bsdfs/roughplastic.cpp:169-253
            int_ior = _lookup_ior(pr.get('int_ior', 'polypropylene'))
            ext_ior = _lookup_ior(pr.get('ext_ior', 'air'))
            rough = b.type == 'roughplastic'
            if int_ior < 0 or ext_ior < 0 or (rough and int_ior == ext_ior):
                raise PluginError("The interior and exterior indices of refraction must be positive%s!" % (" and differ" if rough else ""))
            f32 = np.float32
            eta = f32(int_ior) / f32(ext_ior)
            out.type = A.BSDF_ROUGHPLASTIC if rough else A.BSDF_PLASTIC
            out.eta = float(eta)
            out.inv_eta_2 = float(f32(1.0) / (eta * eta))
            for name in ('diffuse_reflectance', 'specular_reflectance'):
                c = pr.child(name)
                if c is not None and c.type not in ('rgb', 'srgb', 'uniform', 'd65'):
                    raise PluginError("%s: textured %s is not supported" % (b.type, name))
            diff = self._color(pr.child('diffuse_reflectance') or pr.get('diffuse_reflectance', 0.5), key)
            out.reflectance[:] = diff
            d_mean = _mean32(diff if self.spectrum == A.SPEC_RGB else diff[:1])
            s_mean = f32(1.0)
            spec = pr.child('specular_reflectance') or pr.get('specular_reflectance')
            out.specular_reflectance[:] = [1.0] * 3
            if spec is not None:
                sr = self._color(spec, key)
                out.has_specular_reflectance = 1
                out.specular_reflectance[:] = sr
                s_mean = _mean32(sr if self.spectrum == A.SPEC_RGB else sr[:1])
            out.specular_sampling_weight = float(s_mean / (d_mean + s_mean))
            out.nonlinear = int(bool(pr.get('nonlinear', False)))
            if rough:
                distr = str(pr.get('distribution', 'beckmann')).lower()
                if distr not in ('beckmann', 'ggx'):
                    raise PluginError("Specified an invalid distribution \"%s\", must be \"beckmann\" or \"ggx\"!" % distr)
                out.distribution = A.MICROFACET_BECKMANN if distr == 'beckmann' else A.MICROFACET_GGX
                out.sample_visible = int(bool(pr.get('sample_visible', True)))
                if not out.sample_visible:
                    raise PluginError("eval_transmittance(): requires visible normal sampling!")       # microfacet.h:520-521
                if pr.has('alpha_u') or pr.has('alpha_v'):
                    raise PluginError("The 'roughplastic' plugin currently does not support anisotropic microfacet distributions!")
                alpha = float(np.sqrt(f32(pr.get('alpha_sq')))) if pr.has('alpha_sq') else float(pr.get('alpha', 0.1))
                out.alpha_u = out.alpha_v = alpha
                ext, internal = plastic.rough_tables(out.distribution, alpha, out.eta)
                out.ext_transmittance[:] = ext.tolist()
                out.internal_reflectance = float(internal)
            else:
                out.internal_reflectance = float(plastic.fresnel_diffuse_reflectance(f32(1.0) / eta))
        else:
            raise PluginError("unsupported BSDF '%s'" % b.type)
        pr.finish()
        return i

    # -- shapes: src/shapes/cube.cpp, rectangle.cpp:86-125, sphere.cpp --
    def _shape(self, s, key):
        i = self.desc.n_shapes
        if i >= A.MIB_MAX_SHAPES:
            raise PluginError("too many shapes (max %d)" % A.MIB_MAX_SHAPES)
        self.desc.n_shapes += 1
        out = self.desc.shapes[i]
        pr = _Props(s)
        to_world = _to_transform(pr.get('to_world'))
        if pr.get('flip_normals', False):
            to_world = to_world @ ScalarTransform4f.scale([1, 1, -1])
        out.type = {'cube': A.SHAPE_CUBE, 'rectangle': A.SHAPE_RECTANGLE, 'sphere': A.SHAPE_SPHERE, 'obj': A.SHAPE_MESH}[s.type]
        out.mesh_buffer = -1
        mesh = None
        if s.type == 'obj':
            # src/shapes/obj.cpp: world-space triangle mesh + our BVH in one buffer (mesh.py)
            from . import mesh as M
            fname = pr.get('filename', required=True)
            if not os.path.isabs(fname) and not os.path.exists(fname):
                for d in self.search_paths:
                    if os.path.exists(os.path.join(d, fname)):
                        fname = os.path.join(d, fname)
                        break
            try:
                mesh = M.load_obj(fname, to_world.matrix, face_normals=bool(pr.get('face_normals', False)),
                                  flip_tex_coords=bool(pr.get('flip_tex_coords', True)))
            except (FileNotFoundError, ValueError) as e:
                raise PluginError(str(e))
            is_emitter = bool(pr.children_of('emitter'))
            try:
                slot = self._new_buffer(mesh.pack(with_area_table=is_emitter))
                if is_emitter:
                    out.inv_surface_area = mesh.area_table()[2]
            except ValueError as e:
                raise PluginError(str(e))
            out.mesh_buffer, out.mesh_vertices, out.mesh_faces = slot, len(mesh.positions), len(mesh.faces)
            out.mesh_nodes, out.mesh_flags = len(mesh.nodes), mesh.flags | (A.MESH_HAS_AREA_TABLE if is_emitter else 0)
        if s.type == 'sphere':
            c = np.asarray(pr.get('center', [0, 0, 0]), dtype=np.float64)
            r = float(pr.get('radius', 1.0))
            to_world = to_world @ ScalarTransform4f.translate(c) @ ScalarTransform4f.scale(r)
            out.center[:] = (to_world @ [0, 0, 0]).astype(np.float32).tolist()
            out.radius = float(np.linalg.norm(to_world.transform_vector([1, 0, 0])))
        out.to_world[:] = to_world.f32().reshape(-1).tolist()
        out.to_object[:] = to_world.inverse().f32().reshape(-1).tolist()
        if s.type == 'rectangle':
            dp_du = to_world.transform_vector([2, 0, 0])
            dp_dv = to_world.transform_vector([0, 2, 0])
            n = np.linalg.inv(to_world.matrix[:3, :3]).T @ np.array([0, 0, 1.0])
            n = n / np.linalg.norm(n)
            out.frame_s[:] = dp_du.astype(np.float32).tolist()
            out.frame_t[:] = dp_dv.astype(np.float32).tolist()
            out.frame_n[:] = n.astype(np.float32).tolist()
            out.inv_surface_area = float(1.0 / np.linalg.norm(np.cross(dp_du, dp_dv)))
        bsdfs = pr.children_of('bsdf')
        out.bsdf = self._bsdf(bsdfs[0][1] if bsdfs else None, key + '.bsdf', emitter=bool(pr.children_of('emitter')))
        interior, exterior = pr.child('interior'), pr.child('exterior')
        out.interior_medium = self._medium(interior, key + '.interior_medium') if interior else -1
        out.exterior_medium = self._medium(exterior, key + '.exterior_medium') if exterior else -1
        out.emitter = -1
        ems = pr.children_of('emitter')
        if ems:
            out.emitter = self._emitter(ems[0][1], key + '.emitter', shape=i)
        pr.finish()
        # world-space bbox corners for the scene bounding sphere
        if s.type == 'obj':
            lo, hi = mesh.bbox()
            pts = [lo.astype(np.float64), hi.astype(np.float64)]
        elif s.type == 'rectangle':
            pts = [to_world @ [x, y, 0] for x in (-1, 1) for y in (-1, 1)]
        elif s.type == 'cube':
            pts = [to_world @ [x, y, z] for x in (-1, 1) for y in (-1, 1) for z in (-1, 1)]
        else:
            c, r = np.array(out.center[:]), out.radius
            pts = [c - r, c + r]
        self._bbox_pts.extend(pts)
        return i

    # -- emitters: src/emitters/constant.cpp:55-92, directional.cpp:60-108, area.cpp --
    def _emitter(self, e, key, shape=-1):
        i = self.desc.n_emitters
        if i >= A.MIB_MAX_EMITTERS:
            raise PluginError("too many emitters")
        self.desc.n_emitters += 1
        out = self.desc.emitters[i]
        out.shape = shape
        out.env_buffer = -1
        pr = _Props(e)
        if e.type == 'envmap':
            # src/emitters/envmap.cpp:108-232: texels + luminance warp in one buffer (envmap.py)
            from . import envmap as EM, exr
            out.type = A.EMITTER_ENVMAP
            if self.desc.environment >= 0:
                raise PluginError("Only one environment emitter can be specified per scene.")
            self.desc.environment = i
            if pr.has('data'):
                rgb = np.asarray(pr.get('data'), np.float32)
            else:
                fname = pr.get('filename', required=True)
                if not os.path.isabs(fname) and not os.path.exists(fname):
                    for d_ in self.search_paths:
                        if os.path.exists(os.path.join(d_, fname)):
                            fname = os.path.join(d_, fname)
                            break
                if not os.path.exists(fname):
                    raise PluginError('envmap: file "%s" not found' % fname)
                img, names = exr.read_exr(fname)
                idx = [names.index(c) for c in ('R', 'G', 'B')] if all(c in names for c in 'RGB') else [0] * 3
                rgb = np.ascontiguousarray(img[..., idx], np.float32)
            if rgb.ndim != 3 or rgb.shape[2] != 3:
                raise PluginError("envmap: expected an RGB image")
            try:
                buf, W, H, offs, widths = EM.build(rgb, mono=self.spectrum != A.SPEC_RGB)
            except ValueError as ex:
                raise PluginError("envmap: %s" % ex)
            out.env_buffer = self._new_buffer(buf)
            out.env_res[:] = [W, H]
            out.env_scale = float(pr.get('scale', 1.0))
            tw = _to_transform(pr.get('to_world'))
            out.env_to_world[:] = tw.f32().reshape(-1).tolist()
            out.env_to_local[:] = tw.inverse().f32().reshape(-1).tolist()
            out.env_levels = len(offs)
            for k, (o_, w_) in enumerate(zip(offs, widths)):
                out.env_level_offset[k], out.env_level_width[k] = int(o_), int(w_)
            pr.get('mis_compensation', False)
        elif e.type == 'spot':
            # src/emitters/spot.cpp:89-116
            out.type = A.EMITTER_SPOT
            out.radiance[:] = self._color(pr.child('intensity') or pr.get('intensity', 1.0), key)
            if pr.child('texture') is not None or pr.has('texture'):
                raise PluginError("spot: projection textures are not supported")
            cutoff = np.float32(pr.get('cutoff_angle', 20.0))
            beam = np.float32(pr.get('beam_width', float(cutoff * np.float32(3.0) / np.float32(4.0))))
            cutoff_r, beam_r = np.float32(np.deg2rad(np.float64(cutoff))), np.float32(np.deg2rad(np.float64(beam)))
            if cutoff_r < beam_r:
                raise PluginError("spot: cutoff_angle must not be smaller than beam_width")
            out.spot_cutoff_angle = float(cutoff_r)
            out.spot_inv_transition = float(np.float32(1.0) / (cutoff_r - beam_r))
            out.spot_cos_cutoff, out.spot_cos_beam = float(np.float32(np.cos(np.float64(cutoff_r)))), float(np.float32(np.cos(np.float64(beam_r))))
            tw = _to_transform(pr.get('to_world'))
            out.env_to_world[:] = tw.f32().reshape(-1).tolist()
            out.env_to_local[:] = tw.inverse().f32().reshape(-1).tolist()
            self.params[key + '.intensity.value'] = {'kind': 'emitter_radiance', 'emitter': out}
        elif e.type == 'constant':
            out.type = A.EMITTER_CONSTANT
            out.radiance[:] = self._color(pr.child('radiance') or pr.get('radiance', 1.0), key)
            if self.desc.environment >= 0:
                raise PluginError("Only one environment emitter can be specified per scene.")
            self.desc.environment = i
            self.params[key + '.radiance.value'] = {'kind': 'emitter_radiance', 'emitter': out}
        elif e.type == 'directional':
            out.type = A.EMITTER_DIRECTIONAL
            out.radiance[:] = self._color(pr.child('irradiance') or pr.get('irradiance', 1.0), key)
            if pr.has('direction'):
                if pr.has('to_world'):
                    raise PluginError("Only one of the parameters 'direction' and 'to_world' can be specified at the same time!'")
                d = np.asarray(pr.get('direction'), dtype=np.float64)
                d = d / np.linalg.norm(d)
                up, _ = coordinate_system(d)
                tw = ScalarTransform4f.look_at([0, 0, 0], d, up)
            else:
                tw = _to_transform(pr.get('to_world'))
            out.direction[:] = tw.transform_vector([0, 0, 1]).astype(np.float32).tolist()
            self.params[key + '.irradiance.value'] = {'kind': 'emitter_radiance', 'emitter': out}
        elif e.type == 'area':
            out.type = A.EMITTER_AREA
            out.radiance[:] = self._color(pr.child('radiance') or pr.get('radiance', 1.0), key)
            if shape < 0:
                raise PluginError("area emitter must be attached to a shape")
            if self.desc.shapes[shape].type not in (A.SHAPE_RECTANGLE, A.SHAPE_MESH):
                raise PluginError("area emitters are supported on rectangle and obj shapes only (emitter sampling of cube / sphere is not implemented)")
            self.params[key + '.radiance.value'] = {'kind': 'emitter_radiance', 'emitter': out}
        else:
            raise PluginError("unsupported emitter '%s'" % e.type)
        pr.finish()
        return i

    # -- sensor + film + sampler --
    def _sensor(self, s):
        out = self.desc.sensor
        pr = _Props(s)
        out.medium = -1
        med = pr.children_of('medium')
        if med:
            out.medium = self._medium(med[0][1], 'sensor.medium')
        # film: src/films/hdrfilm.cpp:122-243, src/render/film.cpp:10-51
        film = None
        for k, c in pr.children_of('film'):
            film = c
        fw, fh, rf, rad, std = 768, 576, A.RFILTER_GAUSSIAN, 2.0, 0.5
        crop = None
        sample_border = False
        if film is not None:
            fp = _Props(film)
            fw, fh = int(fp.get('width', 768)), int(fp.get('height', 576))
            crop = (int(fp.get('crop_offset_x', 0)), int(fp.get('crop_offset_y', 0)),
                    int(fp.get('crop_width', fw)), int(fp.get('crop_height', fh)))
            sample_border = bool(fp.get('sample_border', False))
            for n in ('pixel_format', 'file_format', 'compensate'):
                fp.get(n)
            # hdrfilm.cpp:60-64: float16 by default; only used when the film is written to disk (api.Film.write)
            self.film_component_format = str(fp.get('component_format', 'float16'))
            if self.film_component_format not in ('float16', 'float32', 'uint32'):
                raise PluginError("The \"component_format\" parameter must either be equal to \"float16\", \"float32\", or \"uint32\"."
                                  " Found %s instead." % self.film_component_format)
            for k, c in fp.children_of('rfilter'):
                rp = _Props(c)
                if c.type == 'box':
                    rf, rad = A.RFILTER_BOX, 0.5
                elif c.type == 'gaussian':          # src/rfilters/gaussian.cpp:50-53
                    std = float(rp.get('stddev', 0.5))
                    rf, rad = A.RFILTER_GAUSSIAN, 4.0 * std
                elif c.type == 'tent':
                    rad = float(rp.get('radius', 1.0))
                    rf = A.RFILTER_TENT
                rp.finish()
            fp.finish()
        if crop is None:
            crop = (0, 0, fw, fh)
        out.film_width, out.film_height = fw, fh
        out.crop_offset_x, out.crop_offset_y, out.crop_width, out.crop_height = crop
        out.rfilter, out.rfilter_radius, out.rfilter_stddev = rf, rad, std
        out.sample_border = int(sample_border)
        # sampler: src/render/sampler.cpp:12-20
        self.sample_count = 4
        self.desc.sampler_base_seed = 0
        for k, c in pr.children_of('sampler'):
            sp = _Props(c)
            if c.type != 'independent':
                raise PluginError("only the 'independent' sampler exists in this build")
            self.sample_count = int(sp.get('sample_count', 4))
            self.desc.sampler_base_seed = int(sp.get('seed', 0))
            sp.finish()
        ident = np.eye(4, dtype=np.float32).reshape(-1).tolist()
        out.sample_to_camera[:] = ident
        if s.type == 'perspective':               # src/sensors/perspective.cpp:120-199
            out.type = A.SENSOR_PERSPECTIVE
            tw = _to_transform(pr.get('to_world'))
            fov = _parse_fov(pr, fw / fh)
            near, far = float(pr.get('near_clip', 1e-2)), float(pr.get('far_clip', 1e4))
            ppx = float(pr.get('principal_point_offset_x', 0.0))
            ppy = float(pr.get('principal_point_offset_y', 0.0))
            out.near_clip, out.far_clip = near, far
            out.principal_point_offset[:] = [ppx, ppy]
            # perspective_projection (include/mitsuba/render/sensor.h:228-263)
            aspect = fw / fh
            rel_size = (crop[2] / fw, crop[3] / fh)
            rel_off = (crop[0] / fw, crop[1] / fh)
            T = ScalarTransform4f
            c2s = (T.scale([1.0 / rel_size[0], 1.0 / rel_size[1], 1.0]) @
                   T.translate([-rel_off[0], -rel_off[1], 0.0]) @
                   T.scale([-0.5, -0.5 * aspect, 1.0]) @
                   T.translate([-1.0, -1.0 / aspect, 0.0]) @
                   T.perspective(fov, near, far))
            out.sample_to_camera[:] = c2s.inverse().f32().reshape(-1).tolist()
            out.to_world[:] = tw.f32().reshape(-1).tolist()
            out.needs_aperture_sample = 0
        elif s.type == 'distant':                 # src/sensors/distant.cpp:140-178
            out.type = A.SENSOR_DISTANT
            if (fw, fh) != (1, 1):
                raise PluginError("This sensor only supports films of size 1x1 Pixels!")
            if pr.has('direction'):
                if pr.has('to_world'):
                    raise PluginError("Only one of the parameters 'direction' and 'to_world'can be specified at the same time!'")
                d = np.asarray(pr.get('direction'), dtype=np.float64)
                d = d / np.linalg.norm(d)
                _, up = coordinate_system(d)
                tw = ScalarTransform4f.look_at([0, 0, 0], d, up)
            else:
                tw = _to_transform(pr.get('to_world'))
            out.to_world[:] = tw.f32().reshape(-1).tolist()
            tgt = pr.get('target')
            if tgt is not None:
                out.target_type = 1
                out.target[:] = [float(x) for x in tgt]
            else:
                out.target_type = 0
            out.needs_aperture_sample = 1          # include/mitsuba/render/endpoint.h:401 default
        elif s.type == 'radiancemeter':           # src/sensors/radiancemeter.cpp:55-95
            out.type = A.SENSOR_RADIANCEMETER
            if pr.has('origin') or pr.has('direction'):
                o = np.asarray(pr.get('origin', required=True), dtype=np.float64)
                d = np.asarray(pr.get('direction', required=True), dtype=np.float64)
                d = d / np.linalg.norm(d)
                _, up = coordinate_system(d)
                tw = ScalarTransform4f.look_at(o, o + d, up)
            else:
                tw = _to_transform(pr.get('to_world'))
            pr.get('wavelength', -1.0)
            out.to_world[:] = tw.f32().reshape(-1).tolist()
            out.needs_aperture_sample = 0
        pr.finish()

    # -- integrators: src/render/integrator.cpp:24-27,1281-1298; common.py:32-41 --
    def _integrator(self, it):
        out = self.desc.integrator
        if it.type == 'moment':
            # src/integrators/moment.cpp wraps integrators and adds their second moments as AOVs (the fork's render tests use
            # it to get variance images); the radiance channels are those of the first nested integrator, which is what runs here
            nested = [c for c in it.children.values() if c.category == 'integrator']
            if len(nested) != 1:
                raise PluginError("moment: exactly one nested integrator is supported")
            return self._integrator(nested[0])
        out.type = {'volpath': A.INTEG_VOLPATH, 'volpathaos': A.INTEG_VOLPATHAOS, 'volpathmis': A.INTEG_VOLPATHMIS,
                    'prbvolpath': A.INTEG_PRBVOLPATH, 'prb_volpathaos': A.INTEG_PRBVOLPATHAOS}[it.type]
        pr = _Props(it)
        is_prb = it.type.startswith('prb')
        max_depth = int(pr.get('max_depth', 6 if is_prb else -1))
        if max_depth < 0 and max_depth != -1:
            raise PluginError("\"max_depth\" must be set to -1 (infinite) or a value >= 0")
        rr_depth = int(pr.get('rr_depth', 5))
        if rr_depth <= 0:
            raise PluginError("\"rr_depth\" must be set to a value greater than zero!")
        out.max_depth, out.rr_depth = max_depth, rr_depth
        out.hide_emitters = int(bool(pr.get('hide_emitters', False)))
        for n in ('timeout', 'block_size', 'samples_per_pass'):
            pr.get(n)
        if it.type == 'volpathmis':
            if self.spectrum == A.SPEC_MONO_POLARIZED:
                raise PluginError("This integrator currently does not support polarized mode!")      # volpathmis.cpp:134-136
            out.spectral_mis = int(bool(pr.get('use_spectral_mis', True)))                          # volpathmis.cpp:93
        out.rng_arg_order = A.RNG_ARGS_LEFT_TO_RIGHT
        out.double_phase_quirk = 1 if it.type == 'volpath' else 0
        # backend extension (not in the fork): M^3 supergrid of local majorants, 0 = the reference's global majorant
        out.majorant_res = int(pr.get('majorant_resolution', 0))
        if it.type == 'prb_volpathaos':
            # extension, off by default: the gradient that agrees with finite differences of volpathaos (mi_b200.h unbiased_adjoint)
            out.unbiased_adjoint = int(bool(pr.get('unbiased', False)))
        if out.majorant_res < 0 or out.majorant_res > 128:
            raise PluginError("\"majorant_resolution\" must be in [0, 128]")
        if it.type == 'prb_volpathaos' and self.spectrum != A.SPEC_MONO_POLARIZED:
            # prb_volpathaos.py:72-78
            raise PluginError("prb_volpathaos requires a monochromatic polarized variant")
        pr.finish()

    def _flatten(self):
        d = self.desc
        d.abi_version = A.MIB_ABI_VERSION
        d.spectrum = self.spectrum
        d.environment = -1
        self._bbox_pts = []
        self.search_paths = list(SEARCH_PATHS)
        root = self.root
        integrator = sensor = None
        # Scene::Scene walks props.objects() (scene.cpp:27-66), i.e. the children sorted by their property names (load_dict: the
        # dictionary keys; XML: `_arg_N` in document order); emitters are numbered in that order -- an area light when its shape
        # comes up -- and Scene::sample_emitter_direction picks among them by index
        names = list(root.children) if getattr(root, 'document_order', False) else properties_order(root.children)
        for name in names:
            c = root.children[name]
            if c.category == 'integrator':
                integrator = c
            elif c.category == 'sensor':
                sensor = c
            elif c.category == 'shape':
                self._shape(c, name)
            elif c.category == 'emitter':
                self._emitter(c, name)
            elif c.category in ('medium', 'bsdf', 'phase', 'texture', 'volume'):
                pass  # only instantiated when referenced
            else:
                raise PluginError("unexpected top-level object '%s' of type '%s'" % (name, c.type))
        if integrator is None:
            integrator = Plugin('volpath', {}, {})
        if sensor is None:
            raise PluginError("the scene needs a sensor")
        self._sensor(sensor)
        self._integrator(integrator)
        # scene bounding sphere (include/mitsuba/core/bbox.h:330-333) -> emitters / distant sensor set_scene
        if self._bbox_pts:
            pts = np.array(self._bbox_pts, dtype=np.float32)
            bmin, bmax = pts.min(0), pts.max(0)
            center = ((bmin + bmax) * np.float32(0.5)).astype(np.float32)
            radius = np.float32(np.linalg.norm((center - bmax).astype(np.float32)))
            radius = max(np.float32(RAY_EPSILON), np.float32(radius * np.float32(1.0 + RAY_EPSILON)))
        else:
            center, radius = np.zeros(3, np.float32), np.float32(RAY_EPSILON)
        for i in range(d.n_emitters):
            d.emitters[i].bsphere_center[:] = center.tolist()
            d.emitters[i].bsphere_radius = float(radius)
        d.sensor.bsphere_center[:] = center.tolist()
        d.sensor.bsphere_radius = float(radius)


def load_dict(d, variant):
    loader = _Loader(variant)
    root = loader.load(d, None)
    if root.type != 'scene':
        raise PluginError("load_dict(): the top-level object must be of type 'scene' in this build")
    return Scene(root, variant)
