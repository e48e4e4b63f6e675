"""Mitsuba XML scene description -> load_dict() dictionary (the data format in front of the path).

Mirrors the parser of the reference, src/core/xml.cpp:
  * tags (:70-107): plugin objects (scene, integrator, sensor, film, sampler, rfilter, shape, bsdf, medium, phase, volume,
    emitter, texture, sizedistr ...), properties float / integer / boolean / string / point / vector / rgb / spectrum /
    transform (translate, rotate, scale, lookat, matrix), ref, default, include, path;
  * `$name` parameter substitution from keyword arguments and <default> tags (:588-607, :470-497);
  * transform operations are applied in document order, each one multiplied from the LEFT (:846-925);
  * unnamed child objects get the argument names `_arg_0`, `_arg_1`, ... (:419-423).
Only plugins of the volumetric hot path can be instantiated afterwards (scene.py refuses the others, as load_dict does).
The module is host-side set-up; nothing here touches the device.
"""
import math
import os
import re
import xml.etree.ElementTree as ET

import numpy as np

from .scene import PluginError, Plugin
from .transforms import ScalarTransform4f

_PROPERTY_TAGS = ('float', 'integer', 'boolean', 'string', 'point', 'vector', 'rgb', 'spectrum', 'transform')
_TRANSFORM_OPS = ('translate', 'rotate', 'scale', 'lookat', 'matrix')
# XML tag -> plugin categories it may instantiate (xml.cpp:109-117 registers the class alias as the tag; "spectrum"
# doubles as a tag for texture plugins)
_OBJECT_TAGS = ('scene', 'integrator', 'sensor', 'film', 'sampler', 'rfilter', 'shape', 'bsdf', 'medium', 'phase', 'volume',
                'emitter', 'texture', 'sizedistr')


class XMLError(PluginError):
    pass


def _tokens(s):
    return [t for t in re.split(r'[\s,]+', s.strip()) if t]


def _float(s, what):
    try:
        return float(s)
    except ValueError:
        raise XMLError('could not parse floating point value "%s" (%s)' % (s, what))


def _coordinate_system_s(n):
    # include/mitsuba/core/vector.h:116-136 (first tangent)
    sign = math.copysign(1.0, n[2])
    a = -1.0 / (sign + n[2])
    b = n[0] * n[1] * a
    return np.array([math.copysign(1.0, n[2]) * n[0] * n[0] * a + 1.0, math.copysign(1.0, n[2]) * b,
                     -math.copysign(1.0, n[2]) * n[0]])


class _Parser:
    def __init__(self, params, base_dir):
        self.params = dict(params)          # name -> string
        self.used = set()
        self.base_dir = base_dir
        self.search = [base_dir] if base_dir else []      # file resolver paths: the scene's directory, then every <path>

    # -- `$name` substitution (xml.cpp:470-497): longest names first, unknown -> error --
    def subst(self, node):
        for k, v in list(node.attrib.items()):
            if '$' not in v:
                continue
            for name in sorted(self.params, key=len, reverse=True):
                if '$' + name in v:
                    v = v.replace('$' + name, self.params[name])
                    self.used.add(name)
            if '$' in v:
                raise XMLError('undefined parameter(s) in "%s" of <%s>' % (v, node.tag))
            node.attrib[k] = v

    def resolve(self, path):
        """FileResolver::resolve: the first search directory that holds the file, else the path as given."""
        if not os.path.isabs(path):
            for d in self.search:
                if os.path.exists(os.path.join(d, path)):
                    return os.path.join(d, path)
        return path

    def check(self, node, allowed, required=None):
        for k in node.attrib:
            if k not in allowed:
                raise XMLError('unexpected attribute "%s" in <%s>' % (k, node.tag))
        for k in (allowed if required is None else required):
            if k not in node.attrib:
                raise XMLError('missing attribute "%s" in <%s>' % (k, node.tag))

    def xyz(self, node, default=0.0):
        """expand_value_to_xyz + parse_vector (xml.cpp:230-262): value="v" | value="x y z" | x= y= z=."""
        if 'value' in node.attrib:
            t = _tokens(node.attrib['value'])
            if len(t) == 1:
                t = t * 3
            if len(t) != 3:
                raise XMLError('"value" attribute of <%s> must have exactly 1 or 3 elements' % node.tag)
            return [_float(x, node.tag) for x in t]
        return [_float(node.attrib.get(a, repr(default)), node.tag) for a in ('x', 'y', 'z')]

    def named_vector(self, node, attr):
        t = _tokens(node.attrib[attr])
        if len(t) != 3:
            raise XMLError('<%s>: "%s" must have three components' % (node.tag, attr))
        return np.array([_float(x, attr) for x in t])

    def transform(self, node):
        self.check(node, ('name',))
        m = ScalarTransform4f()
        for op in node:
            self.subst(op)
            if op.tag == 'translate':
                m = ScalarTransform4f.translate(self.xyz(op)) @ m
            elif op.tag == 'scale':
                m = ScalarTransform4f.scale(self.xyz(op, 1.0)) @ m
            elif op.tag == 'rotate':
                if 'angle' not in op.attrib:
                    raise XMLError('missing attribute "angle" in <rotate>')
                angle = _float(op.attrib.pop('angle'), 'rotate')
                m = ScalarTransform4f.rotate(self.xyz(op), angle) @ m
            elif op.tag == 'lookat':
                self.check(op, ('origin', 'target', 'up'), ('origin', 'target'))
                o, t = self.named_vector(op, 'origin'), self.named_vector(op, 'target')
                up = self.named_vector(op, 'up') if 'up' in op.attrib else np.zeros(3)
                if not np.any(up):
                    d = (t - o) / np.linalg.norm(t - o)
                    up = _coordinate_system_s(d)
                r = ScalarTransform4f.look_at(o, t, up)
                if np.isnan(r.matrix).any():
                    raise XMLError('invalid lookat transformation')
                m = r @ m
            elif op.tag == 'matrix':
                self.check(op, ('value',))
                t = [_float(x, 'matrix') for x in _tokens(op.attrib['value'])]
                if len(t) == 16:
                    mat = np.array(t).reshape(4, 4)
                elif len(t) == 9:
                    mat = np.eye(4)
                    mat[:3, :3] = np.array(t).reshape(3, 3)
                else:
                    raise XMLError('matrix: expected 16 or 9 values')
                m = ScalarTransform4f(mat) @ m
            else:
                raise XMLError('transform nodes can only contain transform operations (got <%s>)' % op.tag)
        return m

    def prop(self, node, within_spectrum=False):
        tag = node.tag
        if tag == 'transform':
            return self.transform(node)
        if tag in ('point', 'vector'):
            return self.xyz(node)
        if tag == 'spectrum' and 'type' in node.attrib:
            return self.obj(node)            # <spectrum type="..."> instantiates a texture plugin
        if tag == 'spectrum':
            has_value, has_file = 'value' in node.attrib, 'filename' in node.attrib
            if has_value == has_file:
                raise XMLError("'spectrum' tag requires one of \"value\" or \"filename\" attributes")
            t = _tokens(node.attrib['value']) if has_value else []
            if len(t) == 1:
                return {'type': 'uniform', 'value': _float(t[0], 'spectrum')}
            if has_file:
                raise XMLError("'spectrum' tag: spectrum files are not supported by this build")
            # wavelength:value pairs (xml.cpp:806-826).  The reference turns them into a texture through its spectral
            # upsampling; that is outside the mono / rgb hot path, so the list is handed on as the `irregular` spectrum
            # plugin and load_dict refuses it like any other plugin that does not exist in this build.
            wl, val = [], []
            for tok in t:
                pair = tok.split(':')
                if len(pair) != 2:
                    raise XMLError('invalid spectrum (expected wavelength:value pairs)')
                wl.append(_float(pair[0], 'spectrum'))
                val.append(_float(pair[1], 'spectrum'))
            return {'type': 'irregular', 'wavelengths': wl, 'values': val}
        self.check(node, ('name', 'value'))
        v = node.attrib['value']
        if tag == 'float':
            return _float(v, node.attrib['name'])
        if tag == 'integer':
            try:
                return int(v)
            except ValueError:
                raise XMLError('could not parse integer value "%s"' % v)
        if tag == 'boolean':
            if v.lower() not in ('true', 'false'):
                raise XMLError('could not parse boolean value "%s" -- must be "true" or "false"' % v)
            return v.lower() == 'true'
        if tag == 'string':
            if node.attrib['name'] == 'filename':
                return self.resolve(v)
            return v
        if tag == 'rgb':
            t = _tokens(v)
            if len(t) == 1:
                t = t * 3
            if len(t) != 3:
                raise XMLError("'rgb' tag requires one or three values (got \"%s\")" % v)
            return {'type': 'rgb', 'value': [_float(x, 'rgb') for x in t]}
        raise XMLError('unsupported tag <%s>' % tag)

    def obj(self, node):
        self.check(node, ('type', 'id', 'name', 'version'), ('type',) if node.tag != 'scene' else ())
        d = {'type': node.attrib.get('type', 'scene') if node.tag != 'scene' else 'scene'}
        if node.tag == 'spectrum' and Plugin.CATEGORY.get(d['type']) not in (None, 'texture'):
            raise XMLError('<spectrum type="%s">: not a texture plugin' % d['type'])
        if 'id' in node.attrib:
            if node.attrib['id'].startswith('_'):
                raise XMLError('invalid id "%s" in element "%s": leading underscores are reserved for internal identifiers.'
                               % (node.attrib['id'], node.tag))
            d['id'] = node.attrib['id']
        arg = 0
        for ch in node:
            self.subst(ch)
            tag = ch.tag
            if tag == 'default':
                self.check(ch, ('name', 'value'))
                if not ch.attrib['name']:
                    raise XMLError('<default>: name must by nonempty')
                self.params.setdefault(ch.attrib['name'], ch.attrib['value'])
                continue
            if tag == 'include':
                self.check(ch, ('filename',))
                path = self.resolve(ch.attrib['filename'])
                if not os.path.exists(path):
                    raise XMLError('included file "%s" not found' % ch.attrib['filename'])
                sub = ET.parse(path).getroot()
                if sub.tag == 'scene':
                    inner = self.obj(sub)
                    inner.pop('type')
                    for k, v in inner.items():
                        d[k if not k.startswith('_arg_') else self._arg(d, arg)] = v
                        arg += k.startswith('_arg_')
                else:
                    d[self._key(sub, d, arg)] = self.obj(sub)
                    arg += 1
                continue
            if tag == 'path':
                # xml.cpp:609-627: add a directory to the file resolver (relative to the scene file)
                self.check(ch, ('value',))
                path = ch.attrib['value']
                if not os.path.isabs(path) and self.base_dir:
                    path = os.path.join(self.base_dir, path)
                if not os.path.exists(path):
                    raise XMLError('<path>: folder "%s" not found' % ch.attrib['value'])
                self.search.append(os.path.normpath(path))
                continue
            if tag == 'alias':
                raise XMLError('<alias> is not supported by this build')
            if tag == 'ref':
                self.check(ch, ('name', 'id'), ('id',))
                name = ch.attrib.get('name') or self._arg(d, arg)
                arg += 'name' not in ch.attrib
                value = {'type': 'ref', 'id': ch.attrib['id']}
            elif tag in _PROPERTY_TAGS and not (tag == 'spectrum' and 'type' in ch.attrib):
                if 'name' not in ch.attrib:
                    raise XMLError('missing attribute "name" in <%s>' % tag)
                name, value = ch.attrib['name'], self.prop(ch)
            elif tag in _OBJECT_TAGS or tag == 'spectrum':
                name = self._key(ch, d, arg)
                arg += 'name' not in ch.attrib
                value = self.obj(ch)
            elif tag in _TRANSFORM_OPS:
                raise XMLError('transform operations can only occur in a transform node')
            else:
                raise XMLError('unexpected tag <%s> inside <%s>' % (tag, node.tag))
            if name in d:
                raise XMLError('Property "%s" was specified multiple times!' % name)
            d[name] = value
        return d

    @staticmethod
    def _arg(d, arg):
        return '_arg_%d' % arg

    def _key(self, ch, d, arg):
        # a named child keeps its name; an unnamed one with an id is stored under the id (so <ref id=...> finds it, also
        # ahead of its declaration); otherwise it gets the positional argument name of the reference parser
        if 'name' in ch.attrib:
            return ch.attrib['name']
        if 'id' in ch.attrib:
            return ch.attrib['id']
        return self._arg(d, arg)


def xml_to_dict(string, base_dir=None, **kwargs):
    """Parse a Mitsuba XML document and return the equivalent load_dict() dictionary."""
    try:
        root = ET.fromstring(string)
    except ET.ParseError as e:
        raise XMLError('Error while loading XML: %s' % e)
    if root.tag not in _OBJECT_TAGS:
        raise XMLError('root element "%s" must be an object' % root.tag)
    if 'version' not in root.attrib:
        raise XMLError('missing version attribute in root element "%s"' % root.tag)
    p = _Parser({k: str(v) for k, v in kwargs.items()}, base_dir)
    p.subst(root)
    d = p.obj(root)
    for k in kwargs:
        if k not in p.used:
            raise XMLError('Unused parameter "%s"!' % k)        # xml.cpp:976-981
    if d.get('type') == 'scene':
        d['_document_order'] = True          # scene.py: the children are in the parser's `_arg_N` order, not to be re-sorted by key
    return d


# ----------------------------------------------------------------------------------------------------------------
# The other direction: a load_dict() dictionary as an XML document the reference can load (the role of
# mi.xml.dict_to_xml, src/python/python/xml.py).  Grids held in memory are written as .vol files next to the scene.
# ----------------------------------------------------------------------------------------------------------------
def _fmt(v):
    return repr(float(v))


def _emit(lines, indent, key, v, out_dir, counter):
    pad = "    " * indent
    name = '' if key is None or key.startswith('_arg_') else ' name="%s"' % key
    if isinstance(v, dict):
        t = v.get('type')
        if t == 'ref':
            lines.append('%s<ref id="%s"%s/>' % (pad, v['id'], name))
            return
        if t == 'rgb':
            lines.append('%s<rgb%s value="%s"/>' % (pad, name, ", ".join(_fmt(x) for x in np.broadcast_to(np.asarray(v['value'], float), (3,)))))
            return
        if t == 'scene':
            tag, attrs = 'scene', ' version="3.0.0"'
        else:
            cat = Plugin.CATEGORY.get(t)
            if cat is None:
                raise XMLError('dict_to_xml: unknown plugin "%s"' % t)
            tag = 'spectrum' if cat == 'texture' and t != 'checkerboard' else cat
            attrs = ' type="%s"' % t
            if 'id' in v:
                attrs += ' id="%s"' % v['id']
            elif indent == 1 and key is not None and not key.startswith('_arg_'):
                attrs += ' id="%s"' % key               # top-level objects: the dictionary key is what <ref> points at
            elif name:
                attrs += name
        body = []
        keys = list(v)
        if t == 'scene' and not v.get('_document_order', False):
            # the reference's load_dict hands the scene its children sorted by key (properties.cpp:48-71); its XML parser in
            # document order: write them sorted, so that both front ends build the same scene (emitter indices)
            from .scene import properties_order
            keys = properties_order(keys)
        for k in keys:
            if k in ('type', 'id', '_document_order'):
                continue
            _emit(body, indent + 1, k, v[k], out_dir, counter)
        if body:
            lines.append('%s<%s%s>' % (pad, tag, attrs))
            lines.extend(body)
            lines.append('%s</%s>' % (pad, tag))
        else:
            lines.append('%s<%s%s/>' % (pad, tag, attrs))
    elif isinstance(v, ScalarTransform4f):
        lines.append('%s<transform%s>' % (pad, name))
        lines.append('%s    <matrix value="%s"/>' % (pad, " ".join(_fmt(x) for x in v.matrix.reshape(-1))))
        lines.append('%s</transform>' % pad)
    elif isinstance(v, bool):
        lines.append('%s<boolean%s value="%s"/>' % (pad, name, 'true' if v else 'false'))
    elif isinstance(v, (int, np.integer)):
        lines.append('%s<integer%s value="%d"/>' % (pad, name, int(v)))
    elif isinstance(v, (float, np.floating)):
        lines.append('%s<float%s value="%s"/>' % (pad, name, _fmt(v)))
    elif isinstance(v, str):
        lines.append('%s<string%s value="%s"/>' % (pad, name, v))
    elif key in ('grid', 'data') and (hasattr(v, 'channel_count') or (isinstance(v, np.ndarray) and v.ndim in (3, 4))):
        # gridvolume data held in memory (a VolumeGrid or a [z, y, x, c] array): the reference takes it as a .vol file
        if out_dir is None:
            raise XMLError('dict_to_xml: the scene holds a grid in memory; pass out_dir so that it can be written as a .vol file')
        from .volumegrid import VolumeGrid
        counter[0] += 1
        fname = 'grid_%d.vol' % counter[0]
        (v if hasattr(v, 'channel_count') else VolumeGrid(np.asarray(v, np.float32))).write(os.path.join(out_dir, fname))
        lines.append('%s<string name="filename" value="%s"/>' % (pad, fname))
    elif key in ('cost', 'values', 'm11', 'm12', 'm33', 'm34') and isinstance(v, (list, tuple, np.ndarray)):
        # phase tables given as arrays (an extension of this package): the reference takes comma-separated strings
        # (tabphase.cpp:44-60); 9 significant digits round-trip a float32
        lines.append('%s<string%s value="%s"/>' % (pad, name, ", ".join('%.9g' % float(x) for x in np.asarray(v, np.float32))))
    elif isinstance(v, (list, tuple, np.ndarray)) and np.asarray(v).shape == (3,):
        lines.append('%s<vector%s value="%s"/>' % (pad, name, ", ".join(_fmt(x) for x in v)))
    else:
        raise XMLError('dict_to_xml: cannot express property "%s" (%s) in XML' % (key, type(v).__name__))


def dict_to_xml(d, out_dir=None):
    """Returns the XML text of a load_dict() dictionary.  xml_to_dict(dict_to_xml(d)) flattens to the same scene."""
    lines = []
    _emit(lines, 0, None, d, out_dir, [0])
    return "\n".join(lines) + "\n"
