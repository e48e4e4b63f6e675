"""Validates the adjoint ORACLE itself (oracle/oadjoint.h) before the CUDA kernels are compared with it.

The reference has no fixed-seed gradient golden; its own methodology is finite differences
(src/integrators/tests/test_ad_volpathaos.py:476-529).  Path replay with detached sampling makes the albedo
gradient the exact derivative of the same-seed primal estimator (albedo only enters through weights) as long
as Russian roulette is off, so that check is tight; sigma_t changes the sampling densities, so that check is
statistical.  prb_volpathaos is biased by construction: its restatement is checked literally
(ratio = albedo) and with the test-only unbiased switch.
"""
import os
import subprocess
import sys

import numpy as np
import pytest

from mitsuba3_nasa_b200 import scene as S, scenes
from orc import OracleScene

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _homog(variant, integ, albedo, sigma_t, spp, phase=None):
    d = scenes.c1_homogeneous_cube(res=4, spp=spp, max_depth=12)
    d['integrator'] = {'type': integ, 'max_depth': 12, 'rr_depth': 1000}
    d['medium1']['sigma_t'] = sigma_t
    d['medium1']['albedo'] = albedo
    if phase:
        d['medium1']['phase'] = phase
    d['sensor']['film']['pixel_format'] = 'luminance'
    if integ == 'prb_volpathaos':
        d['emitter'] = {'type': 'directional', 'direction': [0.3, -0.2, -1.0], 'irradiance': 1.0}
    return S.load_dict(d, variant)


def _fd(variant, integ, key, x0, h, spp, phase, seed):
    vals = []
    for sgn in (+1, -1):
        kw = dict(albedo=0.8, sigma_t=0.6)
        kw[key] = x0 + sgn * h
        vals.append(OracleScene(_homog(variant, integ, spp=spp, phase=phase, **kw)).render_1(seed, spp, nthreads=8)[0])
    return (vals[0] - vals[1]) / (2 * h)


def test_prbvolpath_albedo_gradient_is_the_same_seed_derivative():
    spp = 2048
    o = OracleScene(_homog('llvm_ad_mono', 'prbvolpath', 0.8, 0.6, spp))
    o.render_1_backward(7, spp, [1.0, 0, 0, 0], nthreads=8)
    g = o.grad_get_scalar(0)
    fd = _fd('llvm_ad_mono', 'prbvolpath', 'albedo', 0.8, 1e-2, spp, None, 7)
    assert abs(g[0] / fd - 1) < 2e-4, (g[0], fd)


def test_prbvolpath_sigma_t_gradient_statistical():
    spp = 1 << 15
    o = OracleScene(_homog('llvm_ad_mono', 'prbvolpath', 0.8, 0.6, spp))
    o.render_1_backward(7, spp, [1.0, 0, 0, 0], nthreads=8)
    g = o.grad_get_scalar(0)
    fd = _fd('llvm_ad_mono', 'prbvolpath', 'sigma_t', 0.6, 3e-2, spp, None, 11)
    assert abs(g[3] / fd - 1) < 0.05, (g[3], fd)      # the reference's own tolerance (test_ad_volpathaos.py:599-618)


def test_prb_volpathaos_literal_bias_is_the_albedo():
    """I = replace_grad(1, weight) without the division: each real scatter contributes albedo x the true term, and
    d/d albedo sums one term per scatter, so the literal gradient is exactly albedo x the FD derivative."""
    spp = 2048
    ph = {'type': 'rayleigh'}
    o = OracleScene(_homog('llvm_ad_mono_polarized', 'prb_volpathaos', 0.8, 0.6, spp, ph))
    o.render_1_backward(7, spp, [1.0, 0, 0, 0], nthreads=8)
    g = o.grad_get_scalar(0)
    fd = _fd('llvm_ad_mono_polarized', 'prb_volpathaos', 'albedo', 0.8, 1e-2, spp, ph, 7)
    assert abs(g[0] / fd - 0.8) < 1e-3, (g[0], fd)


def test_prb_volpathaos_machinery_with_division_matches_fd():
    code = r'''
import sys
sys.path.insert(0, %r); sys.path.insert(0, %r)
import test_oracle_adjoint as t
from orc import OracleScene
spp = 2048
ph = {'type': 'rayleigh'}
o = OracleScene(t._homog('llvm_ad_mono_polarized', 'prb_volpathaos', 0.8, 0.6, spp, ph))
o.render_1_backward(7, spp, [1.0, 0, 0, 0], nthreads=8)
g = o.grad_get_scalar(0)
fd = t._fd('llvm_ad_mono_polarized', 'prb_volpathaos', 'albedo', 0.8, 1e-2, spp, ph, 7)
print("RATIO", g[0] / fd)
''' % (ROOT, os.path.join(ROOT, "tests"))
    env = dict(os.environ, ORC_PRB_AOS_UNBIASED="1")
    out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr
    ratio = float(out.stdout.split("RATIO")[1])
    assert abs(ratio - 1) < 2e-4, ratio


def test_grid_gradient_directional_derivative():
    rng = np.random.default_rng(3)
    n = 5
    grid = rng.uniform(0.2, 0.8, size=(n, n, n, 1)).astype(np.float32)
    grid[0, 0, 0, 0] = 1.0                                   # pins the global majorant
    v = rng.uniform(0.0, 1.0, size=grid.shape).astype(np.float32)
    v[0, 0, 0, 0] = 0

    def mk(g, spp):
        d = scenes.c3_cloud(res=4, spp=spp, max_depth=12, integrator='prbvolpath', grid=g, optical_depth=4.0, albedo=0.9,
                            phase={'type': 'hg', 'g': 0.3})
        d['integrator']['rr_depth'] = 1000
        d['object']['interior']['scale'] = 2.0
        d['sensor']['film']['pixel_format'] = 'luminance'
        return S.load_dict(d, 'llvm_ad_mono')
    spp = 1 << 15
    f = mk(grid, spp)
    o = OracleScene(f)
    slot = f.desc.media[0].sigma_t.buffer
    o.render_1_backward(5, spp, [1, 0, 0, 0], nthreads=8)
    g = o.grad_get(slot, grid.size)
    h = 0.05
    vals = [OracleScene(mk(grid + s * h * v, spp)).render_1(9, spp, nthreads=8)[0] for s in (1, -1)]
    fd = (vals[0] - vals[1]) / (2 * h)
    assert abs(float(g @ v.reshape(-1)) / fd - 1) < 0.06


def test_three_channel_albedo_grid_in_a_mono_variant():
    """A 3-channel grid in a 1-channel variant is evaluated as its luminance (grid.cpp:545-575): the gradient reaches all
    three channels with the luminance weights.  The albedo only enters through weights, so with Russian roulette off the
    adjoint is the exact derivative of the same-seed estimator (round 1 sent everything to channel 0 with weight 1)."""
    rng = np.random.default_rng(8)
    alb = rng.uniform(0.4, 0.9, size=(3, 3, 3, 3)).astype(np.float32)
    v = rng.uniform(-1.0, 1.0, size=alb.shape).astype(np.float32)
    from mitsuba3_nasa_b200.transforms import ScalarTransform4f as T

    def mk(a, spp):
        d = scenes.c1_homogeneous_cube(res=4, spp=spp, max_depth=10)
        d['integrator'] = {'type': 'prbvolpath', 'max_depth': 10, 'rr_depth': 1000}
        d['medium1']['albedo'] = {'type': 'gridvolume', 'data': a, 'to_world': T.translate([-1, -1, -1]) @ T.scale(2.0)}
        d['sensor']['film']['pixel_format'] = 'luminance'
        return S.load_dict(d, 'llvm_ad_mono')
    spp = 4096
    f = mk(alb, spp)
    o = OracleScene(f)
    slot = f.desc.media[0].albedo.buffer
    o.render_1_backward(3, spp, [1, 0, 0, 0], nthreads=8)
    g = o.grad_get(slot, alb.size).reshape(alb.shape)
    assert np.all(np.abs(g).reshape(-1, 3).sum(axis=0) > 0)          # every channel receives gradient
    h = 0.02
    vals = [OracleScene(mk(alb + s * h * v, spp)).render_1(3, spp, nthreads=8)[0] for s in (1, -1)]
    fd = (vals[0] - vals[1]) / (2 * h)
    assert abs(float((g * v).sum()) / fd - 1) < 2e-3, (float((g * v).sum()), fd)


def test_prb_volpathaos_albedo_gradient_against_finite_differences_of_the_reference():
    """The reference's own methodology (test_ad_volpathaos.py:504-529, :599-618) with the reference's own numbers: central differences of
    render_1 of the UNMODIFIED volpathaos (scalar_mono_polarized build, 2e5 samples, one random stream at albedo 0.79 and 0.81;
    tests/golden/scalar_ref.npz fd_ref_albedo_*, generated by tests/golden/make_scalar_golden.py) against the adjoint oracle's gradient
    of Stokes I w.r.t. the albedo.  prb_volpathaos' literal gradient is albedo x the true derivative (previous tests),
    so the quotient must be 0.8; with the test-only unbiased switch it is 1.  Tolerance: the reference's 5 %."""
    import scalar_cases
    gold = np.load(os.path.join(ROOT, "tests", "golden", "scalar_ref.npz"))
    lo, hi = (gold['fd_ref_albedo_' + t].reshape(4, 4)[:, 0].astype(np.float64) for t in ('minus', 'plus'))
    fd = (hi - lo) / 0.02
    spp = 1 << 17
    o = OracleScene(S.load_dict(scalar_cases.fd_scene(0.8, integrator='prb_volpathaos', spp=64), 'llvm_ad_mono_polarized'))
    o.render_1_backward(3, spp, [1.0, 0, 0, 0], nthreads=8)
    gs = o.grad_get_scalar(0)
    g = gs[0]
    print("adjoint %.6g, reference FD %.6g, quotient %.4f" % (g, fd[0], g / fd[0]))
    assert abs(g / fd[0] - 0.8) < 0.05 * 0.8
    # the same for sigma_t (reference render_1 at sigma_t 0.57 / 0.63): the literal gradient mixes unbiased transmittance terms with
    # scattering terms that carry the albedo factor (measured quotient 0.919); the unbiased switch must give 1
    lo_s, hi_s = (gold['fd_ref_sigma_t_' + t].reshape(4, 4)[:, 0].astype(np.float64) for t in ('minus', 'plus'))
    fd_s = (hi_s - lo_s) / 0.06
    print("sigma_t: adjoint %.6g, reference FD %.6g, quotient %.4f" % (gs[3], fd_s[0], gs[3] / fd_s[0]))
    assert 0.8 < gs[3] / fd_s[0] < 1.0
    code = r'''
import sys
sys.path.insert(0, %r); sys.path.insert(0, %r)
import scalar_cases
from mitsuba3_nasa_b200 import scene as S
from orc import OracleScene
o = OracleScene(S.load_dict(scalar_cases.fd_scene(0.8, integrator='prb_volpathaos', spp=64), 'llvm_ad_mono_polarized'))
o.render_1_backward(3, 1 << 17, [1.0, 0, 0, 0], nthreads=8)
g = o.grad_get_scalar(0)
print("GRAD", g[0], g[3])
''' % (ROOT, os.path.join(ROOT, "tests"))
    out = subprocess.run([sys.executable, "-c", code], env=dict(os.environ, ORC_PRB_AOS_UNBIASED="1"), capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr
    g1, g3 = (float(x) for x in out.stdout.split("GRAD")[1].split())
    print("unbiased switch: albedo quotient %.4f, sigma_t quotient %.4f" % (g1 / fd[0], g3 / fd_s[0]))
    assert abs(g1 / fd[0] - 1.0) < 0.05 and abs(g3 / fd_s[0] - 1.0) < 0.05


def test_grid_gradient_against_finite_differences_of_the_reference():
    """The C4 kind of gradient -- w.r.t. the cells of a density grid -- against the reference's own numbers: render_1 of the unmodified
    volpathaos (scalar_mono_polarized build, 4e5 samples, one random stream) on the grid +/- 0.05 v, v a fixed random direction that
    leaves the majorant cell alone (tests/scalar_cases.py fd_grid; golden fd_ref_grid_*), against <gradient, v> of the adjoint oracle.
    With the test-only unbiased switch the quotient must be 1 within the reference's 5 % (measured 1.013).  The literal
    prb_volpathaos is 13 % above the true derivative here (its scattering terms carry the albedo, 0.9, and its null-collision terms
    the reciprocal weight) -- outside the reference's own tolerance, which may be why the reference keeps its medium
    configurations switched off (test_ad_volpathaos.py:452-459); pinned as measured."""
    import scalar_cases
    gold = np.load(os.path.join(ROOT, "tests", "golden", "scalar_ref.npz"))
    lo, hi = (gold['fd_ref_grid_' + t].reshape(4, 4)[:, 0].astype(np.float64) for t in ('minus', 'plus'))
    fd = (hi[0] - lo[0]) / 0.1
    _, v = scalar_cases.fd_grid()
    code = r'''
import sys
sys.path.insert(0, %r); sys.path.insert(0, %r)
import scalar_cases
from mitsuba3_nasa_b200 import scene as S
from orc import OracleScene
f = S.load_dict(scalar_cases.fd_grid_scene(0, integrator='prb_volpathaos'), 'llvm_ad_mono_polarized')
o = OracleScene(f)
o.render_1_backward(3, 1 << 18, [1.0, 0, 0, 0], nthreads=8)
g = o.grad_get(f.desc.media[0].sigma_t.buffer, 125)
print("DOT", float(g @ scalar_cases.fd_grid()[1].reshape(-1)))
''' % (ROOT, os.path.join(ROOT, "tests"))
    q = {}
    for tag, env in (('literal', {}), ('unbiased', {'ORC_PRB_AOS_UNBIASED': '1'})):
        out = subprocess.run([sys.executable, "-c", code], env=dict(os.environ, **env), capture_output=True, text=True, timeout=600)
        assert out.returncode == 0, out.stderr
        q[tag] = float(out.stdout.split("DOT")[1]) / fd
    print("reference FD %.6g; quotient literal %.4f, unbiased %.4f" % (fd, q['literal'], q['unbiased']))
    assert abs(q['unbiased'] - 1.0) < 0.05
    assert abs(q['literal'] - 1.13) < 0.05
