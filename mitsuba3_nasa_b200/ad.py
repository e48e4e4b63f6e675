"""Differentiable rendering entry points: the role of `dr.enable_grad` / `dr.backward` / `dr.grad` in
mi.render (src/python/python/util.py:352-510 `_RenderOp`, :512-670 `render` / `render_1`).

The fork hands parameter arrays with `dr.enable_grad(params[key])` to `mi.render(scene, params, ...)`; `dr.backward(loss)`
then reaches `integrator.render_backward(scene, params, grad_out, sensor, seed_grad, spp_grad)`.  There is no Dr.Jit on
this path: the autograd engine is PyTorch's.  A parameter becomes differentiable by assigning a leaf tensor,

    x = params.enable_grad('object.interior_medium.sigma_t.data')     # torch tensor, requires_grad
    img = mi.render(scene, params, seed=0, spp=16)                    # torch tensor wired to x
    loss = ((img - target) ** 2).mean(); loss.backward(); x.grad

and the backward pass is one launch of the path-replay adjoint kernel (prbvolpath / prb_volpathaos), exactly the reference's
two-pass scheme.  Differentiable parameters: sigma_t / albedo of media, as grids (`...data`) or constants (`...value.value`);
other keys raise, as attached-but-unsupported parameters do in the reference.
"""
import ctypes as C

import numpy as np

from . import _abi as A


def _addr(x):
    return C.addressof(x)


def grad_source(scene, key):
    """Where the adjoint kernel leaves d(loss)/d(params[key]): ('grid', slot) | ('scalar', medium, which) | None."""
    flat = scene.flat
    acc = flat.params[key]
    d = flat.desc
    for m in range(d.n_media):
        med = d.media[m]
        for which, vol in (('sigma_t', med.sigma_t), ('albedo', med.albedo)):
            if acc['kind'] == 'buffer' and vol.type == A.VOLUME_GRID and vol.buffer == acc['slot']:
                return ('grid', acc['slot'])
            if acc['kind'] == 'volume_const' and _addr(acc['vol']) == _addr(vol):
                return ('scalar', m, which)
    return None


def read_gradient(scene, key, shape):
    import torch
    src = grad_source(scene, key)
    if src is None:
        raise RuntimeError("parameter '%s' is not differentiable in this backend (sigma_t / albedo of media are)" % key)
    dev = scene.device
    if src[0] == 'grid':
        n = scene.flat.buffers[src[1]].size
        return torch.from_numpy(dev.grad_get(src[1], n).reshape(shape).copy())
    g4 = dev.grad_get_scalar(src[1])                       # [albedo r, g, b, sigma_t]
    n = int(np.prod(shape)) if len(shape) else 1
    if src[2] == 'sigma_t':
        if n != 1:
            raise RuntimeError("gradient of a constant sigma_t is available for scalar values only")
        g = np.array(g4[3])
    else:
        g = g4[:3] if n == 3 else np.array(g4[0])          # one-channel constants accumulate in slot 0 (adjoint.cuh sink_volume)
    return torch.from_numpy(np.asarray(g, dtype=np.float64).reshape(shape).copy())


def make_render_op():
    import torch

    class RenderOp(torch.autograd.Function):
        @staticmethod
        def forward(ctx, scene, params, keys, seed, seed_grad, spp, spp_grad, one, *tensors):
            for k, t in zip(keys, tensors):
                params._assign(k, t.detach().cpu().numpy())
            params.update()
            integ = scene.integrator()
            ctx.cfg = (scene, params, keys, seed_grad, spp_grad or spp, one, [tuple(t.shape) for t in tensors], [t.dtype for t in tensors])
            with torch.no_grad():
                if one:
                    out = integ.render_1(scene, 0, seed, spp)
                    return torch.as_tensor(np.asarray(out, dtype=np.float64))
                return torch.from_numpy(np.ascontiguousarray(integ.render(scene, 0, seed, spp, develop=True)))

        @staticmethod
        def backward(ctx, grad_out):
            scene, params, keys, seed_grad, spp_grad, one, shapes, dtypes = ctx.cfg
            integ = scene.integrator()
            scene.device.grad_zero()
            g = grad_out.detach().cpu().numpy()
            if one:
                integ.render_1_backward(scene, params, g, 0, seed_grad, spp_grad)
            else:
                integ.render_backward(scene, params, g, 0, seed_grad, spp_grad)
            grads = [read_gradient(scene, k, s).to(dt) for k, s, dt in zip(keys, shapes, dtypes)]
            return (None,) * 8 + tuple(grads)

    return RenderOp


_op = None


def render_op():
    global _op
    if _op is None:
        _op = make_render_op()
    return _op
