"""`dr` -- the handful of Dr.Jit entry points the fork's differentiable-rendering scripts use around mi.render
(src/python/python/util.py:352-510 `_RenderOp`; tutorials: `dr.enable_grad(params[key])`, `dr.backward(loss)`,
`dr.grad(params[key])`, `dr.set_grad`, `dr.detach`).  There is no Dr.Jit on this path: the autograd engine is PyTorch's
(ad.py) and these functions are a thin namespace over it, so that

    import mitsuba3_nasa_b200 as mi
    from mitsuba3_nasa_b200 import dr
    params = mi.traverse(scene)
    key = 'object.interior_medium.sigma_t.data'
    dr.enable_grad(params[key])
    img = mi.render(scene, params, spp=16)
    loss = dr.mean(dr.sqr(img - target))
    dr.backward(loss)
    g = dr.grad(params[key])

reads like the reference.  Only leaf parameters and simple losses are supported."""
import numpy as np


def _tensor_of(x):
    """The torch tensor behind a value: a ParamArray from params[key], or a tensor itself."""
    if hasattr(x, '_mi_params'):
        t = x._mi_params._torch.get(x._mi_key)
        if t is None:
            raise RuntimeError("dr: gradients of '%s' are not enabled (call dr.enable_grad(params[key]) first)" % x._mi_key)
        return t
    if hasattr(x, 'requires_grad'):
        return x
    raise TypeError("dr: expected params[key] or a tensor returned by mi.render, got %s" % type(x).__name__)


def enable_grad(*args):
    """dr.enable_grad(params[key], ...): marks parameters as differentiable (util.py: attached parameters reach
    render_backward).  Returns the leaf tensor of the last argument."""
    t = None
    for x in args:
        if hasattr(x, '_mi_params'):
            t = x._mi_params.enable_grad(x._mi_key)
        elif hasattr(x, 'requires_grad_'):
            t = x.requires_grad_(True)
        else:
            raise TypeError("dr.enable_grad: expected params[key] (from mi.traverse) or a tensor")
    return t


def grad_enabled(x):
    try:
        return bool(_tensor_of(x).requires_grad)
    except RuntimeError:
        return False


def backward(loss, flags=None):
    """dr.backward(loss): back-propagates to the attached parameters; mi.render's backward is one launch of the
    path-replay adjoint kernel (ad.py)."""
    loss.backward()


def grad(x):
    """dr.grad(params[key]) -> numpy array of the parameter's shape (zeros before the first backward, as Dr.Jit)."""
    t = _tensor_of(x)
    if t.grad is None:
        return np.zeros(tuple(t.shape), dtype=np.float32)
    return t.grad.detach().cpu().numpy()


def set_grad(x, value):
    import torch
    t = _tensor_of(x)
    t.grad = torch.as_tensor(np.broadcast_to(np.asarray(value, dtype=np.float32), tuple(t.shape)).copy())


def detach(x):
    if hasattr(x, 'detach'):
        return x.detach()
    return np.asarray(x)


def sqr(x):
    return x * x


def mean(x):
    return x.mean()


def sum(x):                      # noqa: A001  (mirrors dr.sum)
    return x.sum()
