"""Triangle meshes of the `obj` shape plugin (src/shapes/obj.cpp, include/mitsuba/render/mesh.h), host side.

    load_obj(path, to_world)  ->  Mesh  (world-space float32 arrays, as OBJMesh builds them)
    Mesh.pack()               ->  one float32 buffer in the layout `mib_shape` documents (mi_b200.h), BVH included

The parser follows obj.cpp:212-395: `v` lines are transformed to world space as they are read, `vn` lines are transformed as
normals and normalised, `vt` lines are flipped vertically (flip_tex_coords, default true), faces are triangulated as a fan
and vertices are the distinct (v, vt, vn) index triples in order of first use.  A file without `vn` lines gets angle-weighted
vertex normals (mesh.cpp:321-372) unless `face_normals` is set.

The BVH is ours (the reference hands meshes to Embree or to its own kd-tree; neither is part of the path): a binary tree
over triangle bounding boxes, median split along the widest axis of the centroid bounds, at most LEAF triangles per leaf,
boxes padded by 1e-5 of the mesh extent so that the box test never rejects a ray the triangle test would accept.  The CUDA
kernels and the CPU oracle walk the same nodes in the same order, so the two pick the same triangle bit for bit."""
import os

import numpy as np

LEAF = int(os.environ.get('MIB_BVH_LEAF', '4'))       # triangles per leaf (MIB_BVH_LEAF: A/B runs, profiles/coverage_bench.py)
INNER = 0x80000000


class Mesh:
    def __init__(self, positions, faces, normals=None, texcoords=None, name="mesh"):
        self.positions = np.ascontiguousarray(positions, np.float32).reshape(-1, 3)
        self.faces = np.ascontiguousarray(faces, np.uint32).reshape(-1, 3)
        self.normals = None if normals is None else np.ascontiguousarray(normals, np.float32).reshape(-1, 3)
        self.texcoords = None if texcoords is None else np.ascontiguousarray(texcoords, np.float32).reshape(-1, 2)
        self.name = name
        if len(self.faces) == 0:
            raise ValueError("mesh '%s' has no faces" % name)
        if self.faces.max() >= len(self.positions):
            raise ValueError("mesh '%s': face references an invalid vertex" % name)
        self.nodes, self.order = build_bvh(self.positions, self.faces)

    @property
    def flags(self):
        return (1 if self.normals is not None else 0) | (2 if self.texcoords is not None else 0)

    def bbox(self):
        return self.positions.min(axis=0), self.positions.max(axis=0)

    def area_table(self):
        """Mesh::build_pmf (mesh.cpp:430-462) + DiscreteDistribution::compute_cdf_scalar (distr_1d.h:243-262): float32 triangle
        areas, their running sum accumulated in double and stored as float32, and the normalisation 1 / total."""
        p = self.positions
        p0, p1, p2 = p[self.faces[:, 0]], p[self.faces[:, 1]], p[self.faces[:, 2]]
        e0, e1 = (p1 - p0).astype(np.float32), (p2 - p0).astype(np.float32)
        c = np.stack([_fms32(e0[:, 1], e1[:, 2], e0[:, 2] * e1[:, 1]), _fms32(e0[:, 2], e1[:, 0], e0[:, 0] * e1[:, 2]),
                      _fms32(e0[:, 0], e1[:, 1], e0[:, 1] * e1[:, 0])], axis=1)
        area = (np.float32(0.5) * np.sqrt((c.astype(np.float32) ** 2).sum(axis=1, dtype=np.float32), dtype=np.float32)).astype(np.float32)
        cdf = np.cumsum(area.astype(np.float64)).astype(np.float32)
        if not cdf[-1] > 0:
            raise ValueError("DiscreteDistribution: no probability mass found!")
        return area, cdf, float(np.float32(1.0) / cdf[-1])

    def pack(self, with_area_table=False):
        parts = [self.positions.reshape(-1)]
        if self.normals is not None:
            parts.append(self.normals.reshape(-1))
        if self.texcoords is not None:
            parts.append(self.texcoords.reshape(-1))
        parts.append(self.faces.reshape(-1).view(np.float32))
        parts.append(self.nodes.reshape(-1))
        parts.append(self.order.view(np.float32))
        if with_area_table:
            area, cdf, _ = self.area_table()
            parts += [area, cdf]
        return np.ascontiguousarray(np.concatenate(parts), np.float32)


def _fma32(a, b, c):
    """float32 fused multiply-add of arrays (a * b is exact in float64)."""
    return (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(np.float32)


def _fms32(a, b, c):
    """float32 fused multiply-subtract a * b - c."""
    return (a.astype(np.float64) * b.astype(np.float64) - c.astype(np.float64)).astype(np.float32)


def _transform_point(M, p):
    # Transform::transform_affine(Point): column 3, then fmadd of columns 0..2 (transform.h)
    M = M.astype(np.float32)
    r = np.broadcast_to(M[:3, 3], p.shape).astype(np.float32)
    for i in range(3):
        r = _fma32(np.broadcast_to(M[:3, i], p.shape), p[:, i:i + 1].repeat(3, axis=1), r)
    return r


def _transform_normal(Minv, n):
    # Transform::transform_affine(Normal): the inverse transpose applied to the normal, no translation
    T = Minv.astype(np.float32).T
    r = (np.broadcast_to(T[:3, 0], n.shape) * n[:, 0:1]).astype(np.float32)
    for i in (1, 2):
        r = _fma32(np.broadcast_to(T[:3, i], n.shape), n[:, i:i + 1].repeat(3, axis=1), r)
    return r


def _normalize32(v):
    l2 = (v.astype(np.float32) ** 2).sum(axis=1, dtype=np.float32)
    with np.errstate(divide='ignore', invalid='ignore'):
        return (v * (np.float32(1) / np.sqrt(l2, dtype=np.float32))[:, None]).astype(np.float32)


def load_obj(path, to_world=None, face_normals=False, flip_tex_coords=True):
    """obj.cpp:146-408.  to_world: 4 x 4 matrix (float64) or None."""
    if not os.path.exists(path):
        raise FileNotFoundError('Error while loading OBJ file "%s": file not found' % os.path.basename(path))
    v, vn, vt = [], [], []
    keys, tris = {}, []
    with open(path, 'r') as f:
        for line in f:
            cur = line.strip(" \t\r\n")
            if len(cur) < 2:
                continue
            if cur[0] == 'v' and cur[1] in ' \t':
                v.append([float(x) for x in cur[2:].split()[:3]])
            elif cur.startswith('vn') and len(cur) > 2 and cur[2] in ' \t':
                if not face_normals:
                    vn.append([float(x) for x in cur[3:].split()[:3]])
            elif cur.startswith('vt') and len(cur) > 2 and cur[2] in ' \t':
                vt.append([float(x) for x in cur[3:].split()[:2]])
            elif cur[0] == 'f' and cur[1] in ' \t':
                ids = []
                for tok in cur[2:].split():
                    parts = tok.split('/')
                    key = (int(parts[0]), int(parts[1]) if len(parts) > 1 and parts[1] else 0,
                           int(parts[2]) if len(parts) > 2 and parts[2] else 0)
                    if key[0] - 1 >= len(v) or key[0] < 1:
                        raise ValueError('Error while loading OBJ file "%s": reference to invalid vertex %d!' % (os.path.basename(path), key[0]))
                    if key not in keys:
                        keys[key] = len(keys)
                    ids.append(keys[key])
                for k in range(2, len(ids)):            # fan: (0, k-1, k)
                    tris.append((ids[0], ids[k - 1], ids[k]))
    if not tris:
        raise ValueError('Error while loading OBJ file "%s": no faces' % os.path.basename(path))
    M = np.eye(4) if to_world is None else np.asarray(to_world, np.float64)
    v = np.asarray(v, np.float32).reshape(-1, 3)
    vw = _transform_point(M, v)
    if not np.all(np.isfinite(vw)):
        raise ValueError("mesh contains invalid vertex position data")
    key_list = sorted(keys, key=keys.get)
    kv = np.array([k[0] - 1 for k in key_list])
    positions = vw[kv]
    texcoords = None
    if vt:
        vt = np.asarray(vt, np.float32).reshape(-1, 2)
        if flip_tex_coords:
            vt[:, 1] = np.float32(1.0) - vt[:, 1]
        kt = np.array([k[1] - 1 for k in key_list])
        if kt.max() >= len(vt):
            raise ValueError("reference to invalid texture coordinate")
        texcoords = np.where((kt >= 0)[:, None], vt[np.maximum(kt, 0)], np.float32(0))
    faces = np.asarray(tris, np.uint32)
    normals = None
    if not face_normals:
        if vn:
            nw = _normalize32(_transform_normal(np.linalg.inv(M), np.asarray(vn, np.float32).reshape(-1, 3)))
            kn = np.array([k[2] - 1 for k in key_list])
            if kn.max() >= len(nw):
                raise ValueError("reference to invalid normal")
            normals = np.where((kn >= 0)[:, None], nw[np.maximum(kn, 0)], np.float32(0))
        else:
            normals = vertex_normals(positions, faces)
    return Mesh(positions, faces, normals, texcoords, os.path.basename(path))


def vertex_normals(positions, faces):
    """Angle-weighted vertex normals (mesh.cpp:321-372, Thuermer & Wuethrich)."""
    p = positions.astype(np.float64)
    v0, v1, v2 = p[faces[:, 0]], p[faces[:, 1]], p[faces[:, 2]]
    n = np.cross(v1 - v0, v2 - v0)
    l = np.linalg.norm(n, axis=1)
    ok = l > 0
    n[ok] /= l[ok][:, None]
    out = np.zeros_like(p)

    def angle(a, b):
        a = a / np.maximum(np.linalg.norm(a, axis=1), 1e-300)[:, None]
        b = b / np.maximum(np.linalg.norm(b, axis=1), 1e-300)[:, None]
        return np.arccos(np.clip((a * b).sum(axis=1), -1.0, 1.0))
    ang = [angle(v1 - v0, v2 - v0), angle(v2 - v1, v0 - v1), angle(v0 - v2, v1 - v2)]
    for j in range(3):
        np.add.at(out, faces[ok, j], n[ok] * ang[j][ok][:, None])
    l = np.linalg.norm(out, axis=1)
    bad = l == 0
    out[~bad] /= l[~bad][:, None]
    out[bad] = [1.0, 0.0, 0.0]
    return out.astype(np.float32)


def build_bvh(positions, faces):
    """-> (nodes float32 [N, 8], order uint32 [F]).  Node: bbox min (3), bbox max (3), then two uint32 bit patterns: inner
    node (left child, right child | 0x80000000), leaf (first entry of `order`, count)."""
    tri = positions[faces.astype(np.int64)]                    # [F, 3, 3]
    lo, hi = tri.min(axis=1), tri.max(axis=1)
    cen = 0.5 * (lo.astype(np.float64) + hi.astype(np.float64))
    pad = np.float32(1e-5) * np.float32(max(float((hi.max(axis=0) - lo.min(axis=0)).max()), 1e-30))
    nodes, order = [], []

    def rec(idx):
        me = len(nodes)
        nodes.append(None)
        bmin = lo[idx].min(axis=0) - pad
        bmax = hi[idx].max(axis=0) + pad
        if len(idx) <= LEAF:
            first = len(order)
            order.extend(int(i) for i in idx)
            nodes[me] = (bmin, bmax, first, len(idx))
            return me
        c = cen[idx]
        ext = c.max(axis=0) - c.min(axis=0)
        ax = int(np.argmax(ext))
        if ext[ax] <= 0:                                        # coincident centroids: split by index
            srt = idx
        else:
            srt = idx[np.argsort(c[:, ax], kind='stable')]
        mid = len(srt) // 2
        left = rec(srt[:mid])
        right = rec(srt[mid:])
        nodes[me] = (bmin, bmax, left, right | INNER)
        return me

    import sys
    old = sys.getrecursionlimit()
    sys.setrecursionlimit(max(old, 10000))
    try:
        rec(np.arange(len(faces)))
    finally:
        sys.setrecursionlimit(old)
    out = np.zeros((len(nodes), 8), np.float32)
    ab = np.zeros((len(nodes), 2), np.uint32)
    for i, (bmin, bmax, a, b) in enumerate(nodes):
        out[i, 0:3], out[i, 3:6] = bmin, bmax
        ab[i] = (a, b)
    out[:, 6:8] = ab.view(np.float32)
    return out, np.asarray(order, np.uint32)


def mesh_offsets(n_vertices, n_faces, n_nodes, flags):
    """Float offsets of the parts of a packed mesh buffer (mirrors the device / oracle code)."""
    off = {'positions': 0}
    o = 3 * n_vertices
    if flags & 1:
        off['normals'] = o
        o += 3 * n_vertices
    if flags & 2:
        off['texcoords'] = o
        o += 2 * n_vertices
    off['faces'] = o
    o += 3 * n_faces
    off['nodes'] = o
    o += 8 * n_nodes
    off['order'] = o
    off['total'] = o + n_faces
    return off
