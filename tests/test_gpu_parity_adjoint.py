"""GPU parity tests of the path-replay adjoint (prbvolpath, prb_volpathaos) through the C ABI.

The CUDA kernels scatter float gradients with atomics in an unspecified order while the
oracle accumulates in double, so gradients are compared within 1e-3 relative to the
largest gradient magnitude of the buffer (north_star: gradients within 1e-3); scalar
gradients (double accumulators on both sides) within 1e-3 relative.  Event counters of
both passes have to agree as for the forward walk.
"""
import numpy as np
import pytest

from mitsuba3_nasa_b200 import scene as S, scenes, _lib
from orc import OracleScene

pytestmark = pytest.mark.gpu

GRAD_TOL = 1e-3


def _homog(variant, integ, res=16, spp=16, sigma_t=1.5, albedo=0.8, phase=None, rr_depth=5):
    d = scenes.c1_homogeneous_cube(res=res, spp=spp, max_depth=16)
    d['integrator'] = {'type': integ, 'max_depth': 16, 'rr_depth': rr_depth}
    d['medium1']['sigma_t'] = sigma_t
    d['medium1']['albedo'] = albedo
    if phase:
        d['medium1']['phase'] = phase
    if 'mono' in variant:
        d['sensor']['film']['pixel_format'] = 'luminance'
    if integ == 'prb_volpathaos':
        d['emitter'] = {'type': 'directional', 'direction': [0.3, -0.2, -1.0], 'irradiance': 1.0}
    return S.load_dict(d, variant)


def _cloud(variant, integ, n=12, res=16, spp=16, albedo=0.9, albedo_grid=False, seed=3):
    rng = np.random.default_rng(seed)
    grid = rng.uniform(0.05, 1.0, size=(n, n, n, 1)).astype(np.float32)
    pol = 'polarized' in variant
    d = scenes.c3_cloud(res=res, spp=spp, max_depth=16, integrator=integ, grid=grid, optical_depth=6.0, albedo=albedo,
                        phase={'type': 'rayleigh'} if pol else {'type': 'hg', 'g': 0.5})
    if albedo_grid:
        agrid = rng.uniform(0.5, 0.95, size=(n, n, n, 3 if 'rgb' in variant else 1)).astype(np.float32)
        d['object']['interior']['albedo'] = {'type': 'gridvolume', 'data': agrid,
                                             'to_world': d['object']['interior']['sigma_t']['to_world']}
    if pol:
        d['emitter'] = {'type': 'directional', 'direction': [0.3, -0.2, -1.0], 'irradiance': 1.0}
    if 'mono' in variant:
        d['sensor']['film']['pixel_format'] = 'luminance'
    return S.load_dict(d, variant)


def _cmp_counters(gc, oc):
    for k in ('n_real', 'n_null', 'n_nee_steps', 'n_surface', 'n_lookups', 'n_samples'):
        assert abs(gc[k] - oc[k]) <= 2e-3 * max(oc[k], 1), (k, gc[k], oc[k])


def _cmp_grid(g, o, what):
    scale = np.abs(o).max()
    assert scale > 0, "oracle gradient of %s is identically zero" % what
    err = np.abs(g.astype(np.float64) - o).max() / scale
    print("%s: |grad|max %.4g, max err / max %.3g, relative L2 %.3g" %
          (what, scale, err, np.linalg.norm(g - o) / np.linalg.norm(o)))
    assert err <= GRAD_TOL, (what, err)


def _primal_compare(sc, spp, count):
    dev = _lib.DeviceScene(sc, 0)
    orc = OracleScene(sc)
    g = dev.sample_lanes(0, spp, 0, count).astype(np.float64)
    o = orc.sample_lanes(0, spp, 0, count).astype(np.float64)
    scale = np.maximum(np.abs(o).max(axis=1, keepdims=True), 1e-6)
    rel = np.abs(g - o) / scale
    bad = (rel > 1e-4).any(axis=1)
    print("primal lanes %d: max rel %.3g, over tol %d" % (count, rel.max(), bad.sum()))
    assert bad.mean() <= 2e-3
    _cmp_counters(dev.counters(), orc.counters())


@pytest.mark.parametrize("variant,integ", [("llvm_ad_mono", "prbvolpath"), ("llvm_ad_rgb", "prbvolpath"),
                                           ("llvm_ad_mono_polarized", "prb_volpathaos")])
def test_prb_primal_lanes_homogeneous(variant, integ):
    sc = _homog(variant, integ, phase={'type': 'rayleigh'} if 'polarized' in variant else None)
    _primal_compare(sc, 16, 16 * 16 * 16)


@pytest.mark.parametrize("variant,integ", [("llvm_ad_mono", "prbvolpath"), ("llvm_ad_rgb", "prbvolpath"),
                                           ("llvm_ad_mono_polarized", "prb_volpathaos")])
def test_prb_primal_lanes_cloud(variant, integ):
    _primal_compare(_cloud(variant, integ), 16, 16 * 16 * 16)


@pytest.mark.parametrize("variant,integ,grad4", [("llvm_ad_mono", "prbvolpath", [1, 0, 0, 0]),
                                                 ("llvm_ad_mono_polarized", "prb_volpathaos", [1, 0, 0, 0]),
                                                 ("llvm_ad_mono_polarized", "prb_volpathaos", [0.3, -1.0, 0.5, 0.45])])
def test_render_1_backward_homogeneous_scalars(variant, integ, grad4):
    sc = _homog(variant, integ, phase={'type': 'rayleigh'} if 'polarized' in variant else None)
    dev, orc = _lib.DeviceScene(sc, 0), OracleScene(sc)
    dev.grad_zero()
    dev.render_1_backward(3, 64, grad4)
    orc.render_1_backward(3, 64, grad4)
    g, o = dev.grad_get_scalar(0), orc.grad_get_scalar(0)
    print("scalar grads gpu", g, "oracle", o)
    assert abs(o[0]) > 0 and abs(o[3]) > 0
    assert np.allclose(g, o, rtol=GRAD_TOL, atol=GRAD_TOL * np.abs(o).max())
    _cmp_counters(dev.counters(), orc.counters())


def test_render_backward_rgb_homogeneous_scalars():
    sc = _homog("llvm_ad_rgb", "prbvolpath")
    dev, orc = _lib.DeviceScene(sc, 0), OracleScene(sc)
    gi = np.random.default_rng(5).uniform(-1, 1, size=(16, 16, 3)).astype(np.float32)
    dev.grad_zero()
    dev.render_backward(1, 32, gi)
    orc.render_backward(1, 32, gi)
    g, o = dev.grad_get_scalar(0), orc.grad_get_scalar(0)
    print("scalar grads gpu", g, "oracle", o)
    assert np.allclose(g, o, rtol=GRAD_TOL, atol=GRAD_TOL * np.abs(o).max())
    _cmp_counters(dev.counters(), orc.counters())


@pytest.mark.parametrize("variant,integ", [("llvm_ad_mono", "prbvolpath"), ("llvm_ad_rgb", "prbvolpath"),
                                           ("llvm_ad_mono_polarized", "prb_volpathaos")])
@pytest.mark.parametrize("albedo_grid", [False, True])
def test_render_backward_cloud_grid(variant, integ, albedo_grid):
    sc = _cloud(variant, integ, albedo_grid=albedo_grid)
    dev, orc = _lib.DeviceScene(sc, 0), OracleScene(sc)
    cd = orc.dev_channels
    gi = np.random.default_rng(11).uniform(-1, 1, size=(16, 16, cd)).astype(np.float32)
    dev.grad_zero()
    dev.render_backward(2, 32, gi)
    orc.render_backward(2, 32, gi)
    _cmp_counters(dev.counters(), orc.counters())
    med = sc.desc.media[0]
    slot = med.sigma_t.buffer
    n = sc.buffers[slot].size
    _cmp_grid(dev.grad_get(slot, n), orc.grad_get(slot, n), "sigma_t grid")
    if albedo_grid:
        slot = med.albedo.buffer
        n = sc.buffers[slot].size
        _cmp_grid(dev.grad_get(slot, n), orc.grad_get(slot, n), "albedo grid")
    else:
        g, o = dev.grad_get_scalar(0), orc.grad_get_scalar(0)
        assert np.allclose(g, o, rtol=GRAD_TOL, atol=GRAD_TOL * np.abs(o).max())


def test_render_1_backward_cloud_polarized():
    sc = _cloud("llvm_ad_mono_polarized", "prb_volpathaos")
    dev, orc = _lib.DeviceScene(sc, 0), OracleScene(sc)
    dev.grad_zero()
    dev.render_1_backward(4, 64, [1.0, 0.5, -0.5, 0.1])
    orc.render_1_backward(4, 64, [1.0, 0.5, -0.5, 0.1])
    slot = sc.desc.media[0].sigma_t.buffer
    n = sc.buffers[slot].size
    _cmp_grid(dev.grad_get(slot, n), orc.grad_get(slot, n), "sigma_t grid")


def test_c2_slab_adjoint_null_boundary():
    """Homogeneous Rayleigh slab under a null top boundary and a diffuse ground: exercises the analytic
    NEE transmittance with the null-surface Mueller block in the same iteration (prb_volpathaos.py:416-486)."""
    sc = S.load_dict(scenes.c2_rayleigh_slab(spp=4, integrator='prb_volpathaos', albedo=0.9), "llvm_ad_mono_polarized")
    dev, orc = _lib.DeviceScene(sc, 0), OracleScene(sc)
    n = 1 << 16
    dev.grad_zero()
    dev.render_1_backward(0, n, [1.0, 0.7, -0.4, 0.0])
    orc.render_1_backward(0, n, [1.0, 0.7, -0.4, 0.0])
    g, o = dev.grad_get_scalar(0), orc.grad_get_scalar(0)
    print("scalar grads gpu", g, "oracle", o)
    assert np.allclose(g, o, rtol=GRAD_TOL, atol=GRAD_TOL * np.abs(o).max())
    _cmp_counters(dev.counters(), orc.counters())


def test_gradients_accumulate_until_zeroed():
    sc = _cloud("llvm_ad_mono", "prbvolpath")
    dev = _lib.DeviceScene(sc, 0)
    slot = sc.desc.media[0].sigma_t.buffer
    n = sc.buffers[slot].size
    dev.grad_zero()
    dev.render_1_backward(0, 16, [1.0])
    g1 = dev.grad_get(slot, n).copy()
    dev.render_1_backward(0, 16, [1.0])
    g2 = dev.grad_get(slot, n)
    assert np.allclose(g2, 2 * g1, rtol=1e-4, atol=1e-6 * np.abs(g1).max())
    dev.grad_zero()
    assert not dev.grad_get(slot, n).any()


def test_backward_needs_prb_integrator():
    sc = S.load_dict(scenes.c1_homogeneous_cube(res=8, spp=4), "llvm_ad_rgb")
    dev = _lib.DeviceScene(sc, 0)
    with pytest.raises(RuntimeError, match="prbvolpath"):
        dev.render_backward(0, 4, np.zeros((8, 8, 3), np.float32))


@pytest.mark.parametrize("rfilter,spp", [("gaussian", 8), ("tent", 8), ("gaussian", 2)])
def test_render_backward_through_reconstruction_filter(rfilter, spp):
    """Adjoint of splat + weight division: dL = sum_p grad[p] / W[p] * w_p (common.py:899-945)."""
    d = scenes.c1_homogeneous_cube(res=12, spp=spp, max_depth=8, rfilter=rfilter)
    d['integrator'] = {'type': 'prbvolpath', 'max_depth': 8, 'rr_depth': 5}
    sc = S.load_dict(d, "llvm_ad_rgb")
    dev, orc = _lib.DeviceScene(sc, 0), OracleScene(sc)
    gi = np.random.default_rng(5).uniform(-1, 1, size=(12, 12, 3)).astype(np.float32)
    dev.grad_zero()
    dev.render_backward(1, spp, gi)
    orc.render_backward(1, spp, gi)
    g, o = dev.grad_get_scalar(0), orc.grad_get_scalar(0)
    print("scalar grads gpu", g, "oracle", o)
    assert np.allclose(g, o, rtol=GRAD_TOL, atol=GRAD_TOL * np.abs(o).max())


def test_kernel_gradient_against_finite_differences_of_the_reference():
    """The CUDA adjoint kernel against the reference's own finite differences (render_1 of the unmodified volpathaos at albedo
    0.79 / 0.81; tests/golden/scalar_ref.npz fd_ref_albedo_*, tests/test_oracle_adjoint.py): prb_volpathaos' literal gradient of
    Stokes I w.r.t. the albedo is 0.8 x the true derivative, 4e6 lanes, within the reference's 5 %."""
    import os
    import scalar_cases
    gold = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "scalar_ref.npz"))
    lo, hi = (gold['fd_ref_albedo_' + t].reshape(4, 4)[:, 0].astype(np.float64) for t in ('minus', 'plus'))
    fd = (hi - lo) / 0.02
    sc = S.load_dict(scalar_cases.fd_scene(0.8, integrator='prb_volpathaos', spp=64), 'llvm_ad_mono_polarized')
    dev = _lib.DeviceScene(sc, 0)
    dev.grad_zero()
    dev.render_1_backward(3, 1 << 22, [1.0, 0, 0, 0])
    g = dev.grad_get_scalar(0)[0]
    print("kernel gradient %.6g, reference FD %.6g, quotient %.4f" % (g, fd[0], g / fd[0]))
    assert abs(g / fd[0] - 0.8) < 0.05 * 0.8


def test_kernel_grid_gradient_against_finite_differences_of_the_reference():
    """The C4 kind of gradient on the CUDA adjoint kernel against the reference's own finite differences (render_1 of the unmodified
    volpathaos on a 5^3 grid +/- 0.05 v; tests/golden/scalar_ref.npz fd_ref_grid_*, tests/test_oracle_adjoint.py): the literal
    prb_volpathaos sits 13 % above the true derivative (the oracle's unbiased switch gives 1.013), and the kernel must sit where the
    oracle's literal restatement sits."""
    import os
    import scalar_cases
    gold = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "scalar_ref.npz"))
    lo, hi = (gold['fd_ref_grid_' + t].reshape(4, 4)[:, 0].astype(np.float64) for t in ('minus', 'plus'))
    fd = (hi[0] - lo[0]) / 0.1
    sc = S.load_dict(scalar_cases.fd_grid_scene(0, integrator='prb_volpathaos'), 'llvm_ad_mono_polarized')
    dev = _lib.DeviceScene(sc, 0)
    dev.grad_zero()
    dev.render_1_backward(3, 1 << 22, [1.0, 0, 0, 0])
    g = np.asarray(dev.grad_get(sc.desc.media[0].sigma_t.buffer, 125), np.float64)
    q = float(g @ scalar_cases.fd_grid()[1].reshape(-1).astype(np.float64)) / fd
    print("kernel <gradient, v> / reference FD = %.4f" % q)
    assert abs(q - 1.13) < 0.05


def test_unbiased_extension_equals_the_oracle_and_the_reference_finite_differences():
    """prb_volpathaos with the integrator property `unbiased = true` (an extension, mi_b200.h unbiased_adjoint: the replayed radiance is
    divided by the step's weight as prbvolpath does): kernel equals the oracle with the same flag, and both agree with finite
    differences of the reference's own volpathaos -- albedo and the cells of a density grid -- within the reference's 5 %."""
    import os
    import scalar_cases
    gold = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "scalar_ref.npz"))

    def stokes_i(name):
        return gold[name].reshape(4, 4)[0, 0].astype(np.float64)
    # parity with the oracle, homogeneous scalars and a grid
    d = scalar_cases.fd_scene(0.8, integrator='prb_volpathaos')
    d['integrator']['unbiased'] = True
    sc = S.load_dict(d, 'llvm_ad_mono_polarized')
    assert sc.desc.integrator.unbiased_adjoint == 1
    dev, orc = _lib.DeviceScene(sc, 0), OracleScene(sc)
    dev.grad_zero()
    dev.render_1_backward(3, 4096, [0.3, -1.0, 0.5, 0.45])
    orc.render_1_backward(3, 4096, [0.3, -1.0, 0.5, 0.45])
    g, o = dev.grad_get_scalar(0), orc.grad_get_scalar(0)
    print("scalar grads gpu", g, "oracle", o)
    assert np.allclose(g, o, rtol=GRAD_TOL, atol=GRAD_TOL * np.abs(o).max())
    # against the reference's finite differences
    dev.grad_zero()
    dev.render_1_backward(3, 1 << 22, [1.0, 0, 0, 0])
    g = dev.grad_get_scalar(0)
    q_albedo = g[0] / ((stokes_i('fd_ref_albedo_plus') - stokes_i('fd_ref_albedo_minus')) / 0.02)
    q_sigma_t = g[3] / ((stokes_i('fd_ref_sigma_t_plus') - stokes_i('fd_ref_sigma_t_minus')) / 0.06)
    d = scalar_cases.fd_grid_scene(0, integrator='prb_volpathaos')
    d['integrator']['unbiased'] = True
    sc = S.load_dict(d, 'llvm_ad_mono_polarized')
    dev, orc = _lib.DeviceScene(sc, 0), OracleScene(sc)
    slot = sc.desc.media[0].sigma_t.buffer
    dev.grad_zero()
    dev.render_1_backward(5, 4096, [1.0, 0, 0, 0])
    orc.render_1_backward(5, 4096, [1.0, 0, 0, 0])
    gg, og = dev.grad_get(slot, 125), orc.grad_get(slot, 125)
    assert np.allclose(gg, og, rtol=GRAD_TOL, atol=GRAD_TOL * np.abs(og).max())
    dev.grad_zero()
    dev.render_1_backward(3, 1 << 22, [1.0, 0, 0, 0])
    gg = np.asarray(dev.grad_get(slot, 125), np.float64)
    q_grid = float(gg @ scalar_cases.fd_grid()[1].reshape(-1).astype(np.float64)) / ((stokes_i('fd_ref_grid_plus') - stokes_i('fd_ref_grid_minus')) / 0.1)
    print("kernel, unbiased, over the reference's FD: albedo %.4f, sigma_t %.4f, grid %.4f" % (q_albedo, q_sigma_t, q_grid))
    assert abs(q_albedo - 1) < 0.05 and abs(q_sigma_t - 1) < 0.05 and abs(q_grid - 1) < 0.05
