"""`mie` phase function + size distributions: host side of the device tabulation.

Reference: src/phase/mie.cpp:65-115 (properties), :117-215 (eval_mie), include/mitsuba/render/mie.h:45-222,
include/mitsuba/render/sizedistr.h:14-137, src/render/sizedistr.cpp:8-48, src/sizedistrs/{lognormal,gamma,
powerlaw,monodisperse}.cpp, include/mitsuba/core/quad.h:27-95 (Gauss-Legendre nodes).

The reference's `mie` plugin evaluates the Lorenz-Mie series per sample and samples the sphere uniformly; its own
comments say "We use a tabulated version of the Mie phase function for sampling in practice" (mie.cpp:251-257).
Here the plugin IS that practice: at load time the series is evaluated on a table of cos(theta) nodes by
libmi_b200.so (mib_mie_tabulate, CUDA) and the walk uses the table like a `tabphase_polarized` plugin.
Extension properties: `n_cost` (number of uniformly spaced cos(theta) nodes, default 1801) or explicit `cost`.
The size-distribution densities (a few hundred values) are evaluated here in float32.
"""
import ctypes as C

import numpy as np

f32 = np.float32


class SizeDistribution:
    """SizeDistribution base (sizedistr.cpp:8-48): Gauss-Legendre nodes on [min_radius, max_radius]."""
    type = None

    def __init__(self, props):
        self.min_radius = f32(props.get('min_radius', 500.0))
        self.max_radius = f32(props.get('max_radius', 5000.0))
        self.g = int(props.get('gauss_points', 100))
        if self.g < 1:
            raise ValueError("gauss_legendre(): n must be >= 1")
        nodes, weights = np.polynomial.legendre.leggauss(self.g)          # float64, cast like quad.h:83-84
        shift = f32(0.5) * (self.max_radius + self.min_radius)
        scale = f32(0.5) * (self.max_radius - self.min_radius)
        self.nodes = (nodes.astype(f32) * scale + shift).astype(f32)
        self.weights = (weights.astype(f32) * scale).astype(f32)
        self.normalization = f32(1.0)
        self.is_monodisperse = False

    def compute_constant(self):
        integral = f32(0.0)
        vals = self.eval(self.nodes, normalize=False)
        for w, v in zip(self.weights, vals):                               # sequential float32 sum (sizedistr.cpp:30-44)
            integral = f32(integral + f32(w * v))
        self.normalization = f32(1.0) / integral

    def eval(self, r, normalize=True):
        raise NotImplementedError

    def eval_gauss_all(self):
        return self.nodes, self.weights, self.eval(self.nodes).astype(f32)


class LogNormalSizeDistr(SizeDistribution):
    type = 'lognormal'

    def __init__(self, props):
        super().__init__(props)
        self.mean_radius = f32(props.get('mean_radius', 1000.0))
        self.std = f32(props.get('std', 100.0))
        ln_std = np.log(self.std, dtype=f32)
        self.std_constant = f32(1.0) / (f32(2.0) * ln_std * ln_std)
        self.compute_constant()

    def eval(self, r, normalize=True):
        r = np.asarray(r, f32)
        a = np.log(r, dtype=f32) - np.log(self.mean_radius, dtype=f32)
        value = np.exp(-(a * a) * self.std_constant, dtype=f32) / r
        return (self.normalization * value).astype(f32) if normalize else value.astype(f32)


class GammaSizeDistr(SizeDistribution):
    type = 'gamma'

    def __init__(self, props):
        super().__init__(props)
        self.a = f32(props.get('a', 600.0))
        self.b = f32(props.get('b', 0.25))
        self.compute_constant()

    def eval(self, r, normalize=True):
        r = np.asarray(r, f32)
        value = np.power(r, (f32(1) - f32(3) * self.b) / self.b, dtype=f32) * np.exp(-r / (self.a * self.b), dtype=f32)
        return (self.normalization * value).astype(f32) if normalize else value.astype(f32)


class PowerLawSizeDistr(SizeDistribution):
    type = 'powerlaw'

    def __init__(self, props):
        super().__init__(props)
        self.exponent = f32(props.get('exponent', 3.0))
        self.compute_constant()

    def eval(self, r, normalize=True):
        value = np.power(np.asarray(r, f32), -self.exponent, dtype=f32)
        return (self.normalization * value).astype(f32) if normalize else value.astype(f32)


class MonodisperseSizeDistr(SizeDistribution):
    type = 'monodisperse'

    def __init__(self, props):
        self.radius = f32(props.get('radius', 1000.0))
        self.min_radius = self.max_radius = self.radius
        self.normalization = f32(1.0)
        self.is_monodisperse = True
        self.g = 1
        self.nodes = np.array([self.radius], f32)
        self.weights = np.array([1.0], f32)

    def eval(self, r, normalize=True):
        return np.isclose(np.asarray(r, f32), self.radius).astype(f32)


SIZE_DISTRIBUTIONS = {c.type: c for c in (LogNormalSizeDistr, GammaSizeDistr, PowerLawSizeDistr, MonodisperseSizeDistr)}


def mie_tabulate(cost, wavelength, size_distr, ior_med=(1.0, 0.0), ior_sph=(1.33, 0.0), nmax=-1, device=0):
    """-> (m11, m12, m33, m34) float32 arrays on the cos(theta) nodes `cost`, computed on the GPU."""
    from ._lib import lib
    l = lib()
    cost = np.ascontiguousarray(cost, f32)
    out = [np.empty(cost.size, f32) for _ in range(4)]
    if size_distr.is_monodisperse:
        radius = np.array([size_distr.min_radius], f32)
        w = sdf = None
        n_gauss = 1
    else:
        radius, w, sdf = [np.ascontiguousarray(a, f32) for a in size_distr.eval_gauss_all()]
        n_gauss = radius.size
    p = lambda a: a.ctypes.data_as(C.c_void_p) if a is not None else None
    rc = l.mib_mie_tabulate(C.c_int(device), C.c_float(wavelength), C.c_float(ior_med[0]), C.c_float(ior_med[1]),
                            C.c_float(ior_sph[0]), C.c_float(ior_sph[1]), C.c_int(int(nmax)), p(radius), p(w), p(sdf),
                            C.c_int(n_gauss), p(cost), C.c_int(cost.size), *[p(o) for o in out])
    if rc != 0:
        raise RuntimeError(l.mib_last_error().decode())
    return tuple(out)


def tabulate_mie_plugin(plugin, props, device=0):
    """Used by scene.py for {'type': 'mie', ...}: returns (cost, m11, m12, m33, m34)."""
    from .scene import PluginError, _Props
    distr = None
    for name, child in props.children_of('sizedistr'):
        if distr is not None:
            raise PluginError("Only a single size distribution can be specified")           # mie.cpp:71-72
        if child.type not in SIZE_DISTRIBUTIONS:
            raise PluginError("Plugin \"%s\" not found!" % child.type)
        cp = _Props(child)
        distr = SIZE_DISTRIBUTIONS[child.type](cp)
        cp.finish()
    if distr is None:
        distr = MonodisperseSizeDistr({'radius': props.get('radius', 1000.0)})              # mie.cpp:78-84
    nmax = int(props.get('nmax', -1))
    ior_med = (float(props.get('ior_med', 1.0)), float(props.get('ior_med_i', 0.0)))
    ior_sph = (float(props.get('ior_sph', 1.33)), float(props.get('ior_sph_i', 0.0)))
    wavelength = float(props.get('wavelength', -1.0))
    if nmax < -1:
        raise PluginError("The number of series terms must be positive or -1 (automatic)!")
    if ior_med[0] <= 0 or ior_sph[0] <= 0:
        raise PluginError("Indices of refraction must be positive!")
    if not wavelength > 0:
        raise PluginError("mie: 'wavelength' must be given (same units as the radii) to tabulate the phase matrix")
    cost = props.get('cost')
    if cost is None:
        cost = np.linspace(-1.0, 1.0, int(props.get('n_cost', 1801)), dtype=np.float64).astype(f32)
    else:
        props.get('n_cost')
        if isinstance(cost, str):
            cost = np.array([float(t) for t in cost.replace(',', ' ').split()], f32)
        cost = np.asarray(cost, f32)
    normalize = bool(props.get('normalize', True))
    m11, m12, m33, m34 = mie_tabulate(cost, wavelength, distr, ior_med, ior_sph, nmax, device)
    if normalize:
        # The tabulated plugins return the UNNORMALISED matrix with the NORMALISED piecewise-linear pdf of m11
        # (tabphase_polarized.cpp:119-205), i.e. every scatter multiplies the throughput by 2 pi * trapz(m11).  A
        # forward peak narrower than the node spacing (cloud droplets: m11(1) ~ 1e3, width ~ 1e-5 in cos theta) makes
        # that factor 4.2 on 1801 uniform nodes.  Rescale so that the discretised m11 is a normalised phase function.
        k = 1.0 / (2.0 * np.pi * np.trapezoid(m11.astype(np.float64), cost.astype(np.float64)))
        m11, m12, m33, m34 = [(a.astype(np.float64) * k).astype(f32) for a in (m11, m12, m33, m34)]
    return cost, m11, m12, m33, m34
