"""Differentiable mi.render / mi.render_1 (ad.py): the torch-autograd stand-in for dr.enable_grad / dr.backward / dr.grad
(src/python/python/util.py:352-670).  The gradient that arrives in `.grad` must be the one the adjoint entry points of the
C ABI produce for the same seed_grad / spp_grad (itself checked against the oracle in test_gpu_parity_adjoint.py), and it
must be a descent direction of the loss (finite-difference check on a scalar parameter, as test_ad_volpathaos.py:476-529 does)."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")

import mitsuba3_nasa_b200 as mi
from mitsuba3_nasa_b200 import scenes

pytestmark = pytest.mark.gpu


def _cloud(integrator, variant, res=16, spp=32):
    mi.set_variant(variant)
    cost, val = scenes.hg_table(0.7, 129)
    d = scenes.c3_cloud(res=res, spp=spp, grid=scenes.cloud_grid(12), phase={'type': 'tabphase', 'cost': cost, 'values': val},
                        optical_depth=5.0, max_depth=12, integrator=integrator)
    return mi.load_dict(d)


def test_render_backward_through_autograd_matches_the_abi():
    scene = _cloud('prbvolpath', 'llvm_ad_mono')
    params = mi.traverse(scene)
    key = 'object.interior_medium.sigma_t.data'
    x = params.enable_grad(key)
    img = mi.render(scene, params, seed=3, spp=32)
    assert isinstance(img, torch.Tensor) and img.shape == (16, 16, 1) and img.requires_grad
    w = torch.from_numpy(np.random.default_rng(1).uniform(0.5, 1.5, (16, 16, 1)).astype(np.float32))
    (img * w).sum().backward()
    assert x.grad is not None and x.grad.shape == x.shape and torch.isfinite(x.grad).all() and x.grad.abs().max() > 0
    # the same backward pass through the integrator API: seed_grad = sample_tea_32(seed, 1)[0] (util.py:617-619)
    dev = scene.device
    dev.grad_zero()
    scene.integrator().render_backward(scene, params, w.numpy(), seed=mi.sample_tea_32(3, 1)[0], spp=32)
    ref = dev.grad_get(scene.flat.desc.media[0].sigma_t.buffer, x.numel()).reshape(x.shape)
    assert np.abs(x.grad.numpy() - ref).max() <= 1e-4 * np.abs(ref).max()
    # the primal value is the ordinary render
    assert np.allclose(img.detach().numpy(), mi.render(scene, seed=3, spp=32), rtol=1e-5, atol=1e-6)


def test_render_1_polarized_gradient_and_non_differentiable_keys():
    scene = _cloud('prb_volpathaos', 'llvm_ad_mono_polarized')
    params = mi.traverse(scene)
    x = params.enable_grad('object.interior_medium.sigma_t.data')
    m = mi.render_1(scene, params, seed=1, spp=1 << 14)
    assert m.shape == (4, 4)
    m[0, 0].backward()                                              # d I / d sigma_t
    assert torch.isfinite(x.grad).all() and x.grad.abs().sum() > 0
    with pytest.raises(RuntimeError, match="not differentiable"):
        params.enable_grad('emitter.radiance.value')
    scene2 = _cloud('volpath', 'llvm_ad_mono')
    p2 = mi.traverse(scene2)
    p2.enable_grad('object.interior_medium.sigma_t.data')
    with pytest.raises(RuntimeError, match="prbvolpath"):
        mi.render(scene2, p2, spp=4)


def test_scalar_albedo_gradient_has_the_sign_and_size_of_finite_differences():
    mi.set_variant('llvm_ad_mono')
    d = scenes.c1_homogeneous_cube(res=8, spp=64, max_depth=16)
    d['integrator'] = {'type': 'prbvolpath', 'max_depth': 16, 'rr_depth': 50}
    d['medium1']['albedo'] = 0.7
    scene = mi.load_dict(d)
    params = mi.traverse(scene)
    key = 'medium1.albedo.value.value'
    a = params.enable_grad(key)
    spp = 4096
    mi.render(scene, params, seed=5, spp=spp).sum().backward()
    g = float(a.grad)
    eps = 0.02
    vals = []
    for v in (0.7 + eps, 0.7 - eps):                                # same seed: correlated samples, smooth difference
        plain = mi.traverse(scene)
        plain[key] = v
        plain.update()
        vals.append(float(np.asarray(mi.render(scene, seed=5, spp=spp)).sum()))
    fd = (vals[0] - vals[1]) / (2 * eps)
    assert fd > 0 and g > 0
    assert abs(g - fd) <= 0.15 * abs(fd)                            # different seed_grad + FD curvature
