"""Host side of the `obj` shape (mitsuba3_nasa_b200/mesh.py; reference: src/shapes/obj.cpp:146-408, src/render/mesh.cpp:321-372)
and the oracle's mesh intersection (oracle/omesh.h; include/mitsuba/render/mesh.h:431-453, src/render/mesh.cpp:1368-1537).
The integrator-level pin of these against the reference's own scalar build is tests/test_oracle_vs_scalar_ref.py (mesh_*)."""
import os

import numpy as np
import pytest

from mitsuba3_nasa_b200 import mesh as M, scene as S, scenes
from mitsuba3_nasa_b200.transforms import ScalarTransform4f as T
from orc import OracleScene

MESHES = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "meshes")


def test_fan_triangulation_and_vertex_splitting():
    m = M.load_obj(os.path.join(MESHES, "box.obj"))
    # 6 quads -> 12 triangles; every corner is used by three faces with three different normals -> 24 vertices
    assert m.faces.shape == (12, 3) and m.positions.shape == (24, 3) and m.normals.shape == (24, 3) and m.texcoords is None
    # fan: the two triangles of a quad share their first vertex and the diagonal
    assert m.faces[0][0] == m.faces[1][0] and m.faces[0][2] == m.faces[1][1]
    # per-face normals as written in the file
    for f in m.faces:
        n = np.cross(m.positions[f[1]] - m.positions[f[0]], m.positions[f[2]] - m.positions[f[0]])
        n /= np.linalg.norm(n)
        for v in f:
            np.testing.assert_allclose(m.normals[v], n, atol=1e-6)


def test_texcoords_are_flipped_and_normals_transformed():
    m = M.load_obj(os.path.join(MESHES, "ground.obj"), T.rotate([1, 0, 0], 90.0).matrix)
    np.testing.assert_allclose(sorted(m.texcoords[:, 1].tolist()), [0, 0, 1, 1], atol=1e-7)      # v -> 1 - v
    assert m.texcoords[0].tolist() == [0.0, 1.0]
    np.testing.assert_allclose(m.normals, np.tile([0, -1, 0], (4, 1)), atol=1e-6)                # (0,0,1) rotated about x


def test_vertex_normals_are_computed_when_the_file_has_none():
    m = M.load_obj(os.path.join(MESHES, "box_no_normals.obj"))
    assert m.positions.shape == (8, 3) and m.normals.shape == (8, 3)
    np.testing.assert_allclose(m.normals, m.positions / np.sqrt(3.0), atol=1e-5)                  # cube corners: diagonal
    f = M.load_obj(os.path.join(MESHES, "box_no_normals.obj"), face_normals=True)
    assert f.normals is None and f.flags == 0


def test_bvh_holds_every_triangle_once_and_bounds_it():
    m = M.load_obj(os.path.join(MESHES, "sphere.obj"))
    assert sorted(m.order.tolist()) == list(range(len(m.faces)))
    ab = m.nodes[:, 6:8].copy().view(np.uint32)
    seen = np.zeros(len(m.faces), int)
    for i, (a, b) in enumerate(ab):
        lo, hi = m.nodes[i, 0:3], m.nodes[i, 3:6]
        if b & M.INNER:
            for c in (a, b & ~np.uint32(M.INNER)):
                assert np.all(m.nodes[c, 0:3] >= lo - 1e-6) and np.all(m.nodes[c, 3:6] <= hi + 1e-6)
        else:
            assert 1 <= b <= M.LEAF
            for f in m.order[a:a + b]:
                seen[f] += 1
                tri = m.positions[m.faces[f]]
                assert np.all(tri >= lo) and np.all(tri <= hi)
    assert np.all(seen == 1)
    off = M.mesh_offsets(len(m.positions), len(m.faces), len(m.nodes), m.flags)
    assert m.pack().size == off['total']


def test_errors_read_like_the_reference():
    with pytest.raises(FileNotFoundError, match="file not found"):
        M.load_obj(os.path.join(MESHES, "nope.obj"))
    d = scenes.c1_homogeneous_cube(res=4, spp=1)
    d['object'] = {'type': 'obj', 'filename': os.path.join(MESHES, "nope.obj")}
    with pytest.raises(S.PluginError, match="file not found"):
        S.load_dict(d, 'llvm_ad_rgb')


def test_mesh_cube_equals_the_analytic_cube_in_the_oracle():
    """The same scene with the cube as 12 triangles and as the analytic slab test: every lane agrees to float accuracy."""
    d = scenes.c1_homogeneous_cube(res=12, spp=8)
    a = OracleScene(S.load_dict(d, 'llvm_ad_rgb')).sample_lanes(0, 8, 0, 12 * 12 * 8)
    d['object'] = {'type': 'obj', 'filename': os.path.join(MESHES, "box.obj"), 'bsdf': {'type': 'null'},
                   'interior': {'type': 'ref', 'id': 'medium1'}}
    b = OracleScene(S.load_dict(d, 'llvm_ad_rgb')).sample_lanes(0, 8, 0, 12 * 12 * 8)
    assert np.abs(a - b).max() <= 2e-6 * np.abs(a).max()


def test_bvh_agrees_with_brute_force():
    """Nearest hit through the BVH (the traversal of oracle/omesh.h restated in numpy) = nearest hit over all triangles."""
    m = M.load_obj(os.path.join(MESHES, "sphere.obj"))
    rng = np.random.default_rng(0)
    tri = m.positions[m.faces.astype(np.int64)].astype(np.float64)
    # geometric check of the tree itself: random rays, brute force over the triangles
    for _ in range(50):
        org = rng.normal(size=3) * 3.0
        dirn = -org / np.linalg.norm(org) + rng.normal(size=3) * 0.2
        dirn /= np.linalg.norm(dirn)
        e1, e2 = tri[:, 1] - tri[:, 0], tri[:, 2] - tri[:, 0]
        p = np.cross(dirn, e2)
        det = (e1 * p).sum(axis=1)
        with np.errstate(divide='ignore', invalid='ignore'):
            inv = 1.0 / det
            tv = org - tri[:, 0]
            u = (tv * p).sum(axis=1) * inv
            q = np.cross(tv, e1)
            v = (q * dirn).sum(axis=1) * inv
            t = (e2 * q).sum(axis=1) * inv
        hit = (u >= 0) & (u <= 1) & (v >= 0) & (u + v <= 1) & (t >= 0)
        t_brute = t[hit].min() if hit.any() else np.inf
        t_bvh = _bvh_nearest(m, org, dirn)
        assert (np.isinf(t_brute) and np.isinf(t_bvh)) or abs(t_brute - t_bvh) <= 1e-5 * max(1.0, t_brute)


def _bvh_nearest(m, org, dirn):
    """The traversal of oracle/omesh.h restated in numpy float64 (box test + leaf triangles)."""
    ab = m.nodes[:, 6:8].copy().view(np.uint32)
    best, stack = np.inf, [0]
    with np.errstate(divide='ignore', invalid='ignore'):
        inv = 1.0 / dirn
    while stack:
        i = stack.pop()
        t1, t2 = (m.nodes[i, 0:3] - org) * inv, (m.nodes[i, 3:6] - org) * inv
        tmin, tmax = max(0.0, np.fmin(t1, t2).max()), min(best, np.fmax(t1, t2).min())
        if not tmin <= tmax:
            continue
        a, b = int(ab[i, 0]), int(ab[i, 1])
        if b & M.INNER:
            stack += [b & 0x7fffffff, a]
        else:
            for f in m.order[a:a + b]:
                p0, p1, p2 = m.positions[m.faces[f]].astype(np.float64)
                e1, e2 = p1 - p0, p2 - p0
                p = np.cross(dirn, e2)
                det = e1 @ p
                if det == 0:
                    continue
                tv = org - p0
                u = (tv @ p) / det
                q = np.cross(tv, e1)
                v = (dirn @ q) / det
                t = (e2 @ q) / det
                if 0 <= u <= 1 and v >= 0 and u + v <= 1 and 0 <= t < best:
                    best = t
    return best
