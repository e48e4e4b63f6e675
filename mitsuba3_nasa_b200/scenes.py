"""Synthetic scenes of BASELINE.json's configs, as load_dict dictionaries.

All data is generated deterministically in-repo (there is no network for datasets).
"""
import numpy as np

from .transforms import ScalarTransform4f as T


def c1_homogeneous_cube(res=256, spp=64, max_depth=64, rfilter='box'):
    """C1: homogeneous HG (g=0.7) medium cube, volpath, rgb, 256x256, 64 spp."""
    return {
        'type': 'scene',
        'integrator': {'type': 'volpath', 'max_depth': max_depth, 'rr_depth': 5},
        'sensor': {
            'type': 'perspective', 'fov': 45.0,
            'to_world': T.look_at(origin=[0, 0, 5], target=[0, 0, 0], up=[0, 1, 0]),
            'film': {'type': 'hdrfilm', 'width': res, 'height': res, 'pixel_format': 'rgb',
                     'component_format': 'float32', 'rfilter': {'type': rfilter}},
            'sampler': {'type': 'independent', 'sample_count': spp},
        },
        'medium1': {
            'type': 'homogeneous', 'sigma_t': 2.0,
            'albedo': {'type': 'rgb', 'value': [0.9, 0.8, 0.7]},
            'has_spectral_extinction': False,
            'phase': {'type': 'hg', 'g': 0.7},
        },
        'object': {'type': 'cube', 'bsdf': {'type': 'null'},
                   'interior': {'type': 'ref', 'id': 'medium1'}},
        'emitter': {'type': 'constant', 'radiance': 1.0},
    }


def c2_rayleigh_slab(spp=4, integrator='volpathaos', max_depth=64, scale=1.0e4, sigma_t=0.5, albedo=1.0,
                     depolarization=0.0, ground=0.3, sun=True):
    """C2: plane-parallel Rayleigh slab, distant sensor (view zenith 30 deg), sun at 30 deg.

    Layout of src/integrators/tests/test_ad_volpathaos.py:357-379 (null top at z=1, diffuse
    bottom at z=0).  The horizontal extent defaults to 1e4 instead of the test's 1e6: with
    1e6 the distant sensor's ray origin sits 2.8e6 units away and fp32 places the entry
    point with an error of ~0.25 slab thicknesses (SURVEY F4 discussion in DESIGN.md)."""
    d = {
        'type': 'scene',
        'integrator': {'type': integrator, 'max_depth': max_depth, 'rr_depth': 5},
        'sensor': {
            'type': 'distant', 'direction': [0.5, 0.0, -0.8660254], 'target': [0.0, 0.0, 1.0],
            'film': {'type': 'hdrfilm', 'width': 1, 'height': 1, 'rfilter': {'type': 'box'}},
            'sampler': {'type': 'independent', 'sample_count': spp},
        },
        'a_medium': {
            'type': 'homogeneous', 'sigma_t': sigma_t, 'albedo': albedo,
            'has_spectral_extinction': False,
            'phase': {'type': 'rayleigh', 'depolarization': depolarization},
        },
        'top': {'type': 'rectangle', 'bsdf': {'type': 'null'},
                'to_world': T.translate([0.0, 0.0, 1.0]) @ T.scale([scale, scale, 1.0]),
                'interior': {'type': 'ref', 'id': 'a_medium'}},
        'bottom': {'type': 'rectangle', 'bsdf': {'type': 'diffuse', 'reflectance': ground},
                   'to_world': T.scale([scale, scale, 1.0]),
                   'exterior': {'type': 'ref', 'id': 'a_medium'}},
    }
    if sun:
        d['light'] = {'type': 'directional', 'direction': [-0.5, 0.0, -0.8660254], 'irradiance': 1.0}
    else:
        d['light'] = {'type': 'constant', 'radiance': 1.0}
    return d


def aos_ocean(spp=4, integrator='volpathaos', max_depth=32, scale=100.0, rough=True, alpha=0.15, distribution='beckmann',
              sun=True, ocean_albedo=0.6, view_zenith_deg=30.0):
    """Coupled atmosphere-ocean system of the fork (the scenes volpathaos.cpp:232-290,312-337 is written
    for): a Rayleigh atmosphere slab (z in [0, 1]) over a (rough) dielectric ocean surface at z = 0, a scattering ocean layer
    (z in [-1, 0]) and a diffuse sea floor.  All interfaces are horizontal rectangles, as the integrator assumes."""
    import math
    cz = math.cos(math.radians(view_zenith_deg)); sz = math.sin(math.radians(view_zenith_deg))
    surface = ({'type': 'roughdielectric', 'int_ior': 'water', 'ext_ior': 'air', 'distribution': distribution, 'alpha': alpha}
               if rough else {'type': 'dielectric', 'int_ior': 'water', 'ext_ior': 'air'})
    d = {
        'type': 'scene',
        'integrator': {'type': integrator, 'max_depth': max_depth, 'rr_depth': 5},
        'sensor': {
            'type': 'distant', 'direction': [sz, 0.0, -cz], 'target': [0.0, 0.0, 1.0],
            'film': {'type': 'hdrfilm', 'width': 1, 'height': 1, 'rfilter': {'type': 'box'}},
            'sampler': {'type': 'independent', 'sample_count': spp},
        },
        'atmosphere': {'type': 'homogeneous', 'sigma_t': 0.3, 'albedo': 1.0, 'has_spectral_extinction': False,
                       'phase': {'type': 'rayleigh', 'depolarization': 0.03}},
        'ocean': {'type': 'homogeneous', 'sigma_t': 1.5, 'albedo': ocean_albedo, 'has_spectral_extinction': False,
                  'phase': {'type': 'hg', 'g': 0.8}},
        'top': {'type': 'rectangle', 'bsdf': {'type': 'null'},
                'to_world': T.translate([0.0, 0.0, 1.0]) @ T.scale([scale, scale, 1.0]),
                'interior': {'type': 'ref', 'id': 'atmosphere'}},
        'surface': {'type': 'rectangle', 'bsdf': surface,
                    'to_world': T.scale([scale, scale, 1.0]),
                    'interior': {'type': 'ref', 'id': 'ocean'}, 'exterior': {'type': 'ref', 'id': 'atmosphere'}},
        'floor': {'type': 'rectangle', 'bsdf': {'type': 'diffuse', 'reflectance': 0.2},
                  'to_world': T.translate([0.0, 0.0, -1.0]) @ T.scale([scale, scale, 1.0]),
                  'exterior': {'type': 'ref', 'id': 'ocean'}},
    }
    if sun:
        d['light'] = {'type': 'directional', 'direction': [-0.5, 0.2, -0.8], 'irradiance': 1.0}
    else:
        d['light'] = {'type': 'constant', 'radiance': 1.0}
    return d


def cloud_grid(n=256, seed=1234, blobs=64):
    """sum of Gaussian blobs clipped to [0,1], zeroed below 0.02; [z,y,x,1] fp32."""
    rng = np.random.default_rng(seed)
    centres = rng.uniform(0.2, 0.8, size=(blobs, 3))
    sigmas = rng.uniform(0.03, 0.12, size=blobs)
    amps = rng.uniform(0.2, 1.0, size=blobs)
    ax = (np.arange(n, dtype=np.float32) + 0.5) / n
    grid = np.zeros((n, n, n), dtype=np.float32)
    for c, s, a in zip(centres, sigmas, amps):
        gx = np.exp(-0.5 * ((ax - c[0]) / s) ** 2).astype(np.float32)
        gy = np.exp(-0.5 * ((ax - c[1]) / s) ** 2).astype(np.float32)
        gz = np.exp(-0.5 * ((ax - c[2]) / s) ** 2).astype(np.float32)
        grid += np.float32(a) * gz[:, None, None] * gy[None, :, None] * gx[None, None, :]
    grid = np.clip(grid, 0.0, 1.0)
    grid[grid < 0.02] = 0.0
    return grid[..., None].astype(np.float32)


def hg_table(g=0.85, n=1801):
    """Analytic stand-in table (cos theta nodes, HG values) for tabphase tests."""
    cost = np.linspace(-1.0, 1.0, n)
    val = (1 - g * g) / (4 * np.pi * (1 + g * g - 2 * g * cost) ** 1.5)
    return cost.astype(np.float32), val.astype(np.float32)


def c3_cloud(res=1024, spp=256, grid_res=256, max_depth=64, integrator='volpath', phase=None, grid=None,
             optical_depth=20.0, albedo=0.99, rfilter='box', majorant_resolution=0):
    """C3: heterogeneous gridvolume cloud in the [-1,1]^3 cube, tabulated Mie phase.
    majorant_resolution > 0 selects the backend's fast mode (supergrid of local majorants; not in the fork)."""
    if grid is None:
        grid = cloud_grid(grid_res)
    if phase is None:
        phase = {'type': 'mie', 'ior_sph': 1.33, 'wavelength': 0.55, 'n_cost': 1801,
                 'sizedistr': {'type': 'lognormal', 'mean_radius': 10.0, 'std': 1.4,
                               'min_radius': 1.0, 'max_radius': 40.0, 'gauss_points': 100}}
    gmax = float(grid.max())
    scale = optical_depth / (2.0 * max(gmax, 1e-8))   # majorant optical depth across the cube (length 2)
    return {
        'type': 'scene',
        'integrator': dict({'type': integrator, 'max_depth': max_depth, 'rr_depth': 5},
                           **({'majorant_resolution': majorant_resolution} if majorant_resolution else {})),
        'sensor': {
            'type': 'perspective', 'fov': 45.0,
            'to_world': T.look_at(origin=[0, 0, 5], target=[0, 0, 0], up=[0, 1, 0]),
            'film': {'type': 'hdrfilm', 'width': res, 'height': res, 'rfilter': {'type': rfilter}},
            'sampler': {'type': 'independent', 'sample_count': spp},
        },
        'object': {
            'type': 'cube', 'bsdf': {'type': 'null'},
            'interior': {
                'type': 'heterogeneous',
                'sigma_t': {'type': 'gridvolume', 'data': grid,
                            'to_world': T.translate([-1, -1, -1]) @ T.scale(2.0)},
                'scale': scale, 'albedo': albedo, 'has_spectral_extinction': False,
                'phase': phase,
            },
        },
        'emitter': {'type': 'constant', 'radiance': 1.0},
    }


def c4_cloud_adjoint(res=512, spp=64, grid_res=256, polarized=True, **kw):
    """C4: C3 scene at 512x512x64spp for the prb adjoint w.r.t. the density grid."""
    d = c3_cloud(res=res, spp=spp, grid_res=grid_res, max_depth=kw.pop('max_depth', 64),
                 integrator='prb_volpathaos' if polarized else 'prbvolpath', **kw)
    return d
