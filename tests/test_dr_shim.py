"""Host logic of the `dr` shim (mitsuba3_nasa_b200/dr.py; reference: the Dr.Jit calls around mi.render in
src/python/python/util.py:352-510 and the fork's optimisation scripts).  No GPU: tensors only; the render round trip is in
tests/test_gpu_ad.py."""
import numpy as np
import pytest
import torch

from mitsuba3_nasa_b200 import dr


def test_enable_grad_backward_grad_on_plain_tensors():
    x = torch.tensor([1.0, 2.0, 3.0])
    t = dr.enable_grad(x)
    assert t is x and dr.grad_enabled(x)
    np.testing.assert_array_equal(dr.grad(x), np.zeros(3, np.float32))      # zeros before the first backward, as Dr.Jit
    loss = dr.mean(dr.sqr(x - 1.0))
    dr.backward(loss)
    np.testing.assert_allclose(dr.grad(x), 2.0 * (np.array([1.0, 2.0, 3.0]) - 1.0) / 3.0, rtol=1e-6)
    dr.set_grad(x, 0.0)
    np.testing.assert_array_equal(dr.grad(x), np.zeros(3, np.float32))
    assert float(dr.sum(dr.detach(x))) == 6.0 and not dr.detach(x).requires_grad


def test_rejects_what_is_not_a_parameter():
    with pytest.raises(TypeError):
        dr.enable_grad(3.0)
    with pytest.raises(TypeError):
        dr.grad([1.0, 2.0])


def test_params_proxy_without_enabled_gradient():
    class FakeParams:
        _torch = {}

    class Proxy:
        _mi_params, _mi_key = FakeParams(), 'medium.sigma_t.data'
    with pytest.raises(RuntimeError, match="not enabled"):
        dr.grad(Proxy())
    assert not dr.grad_enabled(Proxy())
