"""Scenes of the integrator-level pin (shared by tests/golden/make_scalar_golden.py and
tests/test_oracle_vs_scalar_ref.py).  Every case is a load_dict dictionary small enough for the reference's scalar
variants (oracle/_ref/scalar, built by oracle/build_ref_scalar.sh) to dump all of its samples in a second.

A case: name -> dict(variant=..., scene=<load_dict dictionary>, spp=..., seed=..., kind='samples' | 'render1').
Seeds follow the reference's scalar render loop (integrator.cpp:230, 1103-1123): the film is one block, pixel (x, y) is
pixel i = morton_encode(x, y) of the block and draws its spp samples from PCG32(initstate = seed * W * H + i, default
stream).  render_1 (integrator.cpp:510-551 with one thread): one stream seeded `seed`, spp samples."""
import os

import numpy as np

from mitsuba3_nasa_b200 import scenes
from mitsuba3_nasa_b200.transforms import ScalarTransform4f as T


def morton_encode(x, y):
    r = 0
    for b in range(16):
        r |= ((x >> b) & 1) << (2 * b) | ((y >> b) & 1) << (2 * b + 1)
    return r


def pixel_seeds(W, H, seed):
    """(row-major pixel index, per-pixel seed) in the order the reference visits the pixels."""
    px = [(morton_encode(x, y), y * W + x) for y in range(H) for x in range(W)]
    px.sort()
    return np.array([p for _, p in px], np.uint32), np.array([seed * W * H + i for i, _ in px], np.uint32)


def polarized_table(n=181):
    cost = np.linspace(-1, 1, n).astype(np.float32)
    m11 = (1.0 + 0.5 * cost ** 2 + 0.8 * np.exp(4 * (cost - 1))).astype(np.float32)
    m12 = (-0.4 * (1 - cost ** 2)).astype(np.float32)
    m33 = (0.9 * cost).astype(np.float32)
    m34 = (0.1 * (1 - cost ** 2)).astype(np.float32)
    k = np.float32(1.0 / (2 * np.pi * np.trapezoid(m11.astype(np.float64), cost.astype(np.float64))))
    return {'type': 'tabphase_polarized', 'cost': cost, 'm11': m11 * k, 'm12': m12 * k, 'm33': m33 * k, 'm34': m34 * k}


def _grid(n, seed, lo=0.0, channels=1):
    return np.random.default_rng(seed).uniform(lo, 1.0, (n, n, n, channels)).astype(np.float32)


def cloud_no_plateau(n=32):
    """The C3 blob field without its clipped plateau.  Where the field sits at its maximum over a whole cell, sigma_n =
    majorant - sigma_t is 0 or one ulp depending on how the eight trilinear weights round, and ratio tracking stops or goes
    on accordingly: the number of random draws then hangs on the last bit of the look-up position, which differs between the
    reference's scalar and JIT arithmetic (left-to-right sums vs fma chains).  A ramp makes the maximum a single texel."""
    g = scenes.cloud_grid(n)
    ramp = 1.0 - 0.3 * (np.arange(n, dtype=np.float32) + 0.5) / n
    return (g * ramp[None, None, :, None]).astype(np.float32)


MESH_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "meshes")


def mesh_scene(hetero=False, res=16):
    m = lambda n: os.path.join(MESH_DIR, n)
    if hetero:
        medium = {'type': 'heterogeneous', 'scale': 3.0, 'albedo': 0.9, 'phase': {'type': 'hg', 'g': 0.4},
                  'sigma_t': {'type': 'gridvolume', 'data': _grid(8, 11), 'to_world': T.translate([-1, -1, 0.2]) @ T.scale(2.0)}}
    else:
        medium = {'type': 'homogeneous', 'scale': 1.5, 'sigma_t': {'type': 'rgb', 'value': [0.5, 0.7, 1.0]},
                  'albedo': {'type': 'rgb', 'value': [0.9, 0.8, 0.6]}, 'phase': {'type': 'isotropic'}}
    return {
        'type': 'scene',
        'integrator': {'type': 'volpath', 'max_depth': 8, 'rr_depth': 5},
        'sensor': {'type': 'perspective', 'fov': 40.0, 'near_clip': 0.1, 'far_clip': 100.0,
                   'to_world': T.look_at(origin=[6.0, -5.0, 4.0], target=[0.0, 0.0, 0.8], up=[0, 0, 1]),
                   'film': {'type': 'hdrfilm', 'width': res, 'height': res, 'rfilter': {'type': 'box'}},
                   'sampler': {'type': 'independent', 'sample_count': 16}},
        'emitter': {'type': 'constant', 'radiance': {'type': 'rgb', 'value': [1.0, 0.9, 0.8]}},
        'medium1': medium,
        'container': {'type': 'obj', 'filename': m('box.obj'), 'to_world': T.translate([0, 0, 1.2]), 'bsdf': {'type': 'null'},
                      'interior': {'type': 'ref', 'id': 'medium1'}},
        'ball': {'type': 'obj', 'filename': m('sphere.obj'), 'to_world': T.translate([2.2, 0.5, 0.7]) @ T.scale(0.7),
                 'bsdf': {'type': 'diffuse', 'reflectance': {'type': 'rgb', 'value': [0.7, 0.3, 0.2]}}},
        'crate': {'type': 'obj', 'filename': m('box_no_normals.obj'), 'to_world': T.translate([-2.0, 1.5, 0.4]) @ T.rotate([0, 0, 1], 30.0) @ T.scale(0.4),
                  'bsdf': {'type': 'diffuse', 'reflectance': {'type': 'rgb', 'value': [0.2, 0.5, 0.7]}}},
        'floor': {'type': 'obj', 'filename': m('ground.obj'),
                  'bsdf': {'type': 'diffuse', 'reflectance': {'type': 'checkerboard', 'color0': {'type': 'rgb', 'value': [0.2, 0.2, 0.2]},
                                                               'color1': {'type': 'rgb', 'value': [0.8, 0.8, 0.8]},
                                                               'to_uv': T.scale([8, 8, 8])}}},
    }


PM_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "participating_media")


def xml_cases():
    """The fork's own render-test scenes (resources/data/tests/scenes/participating_media, copied with their meshes, grids,
    the museum environment map and the reference images into tests/golden/): name -> dict(variant, file, spp, seed).  The
    homogeneous one is pinned sample by sample; the two heterogeneous ones hold a plateau of 231 grid cells that sit exactly
    at the majorant, where the reference's scalar and JIT arithmetic decide null collisions differently,
    so only a fraction bound is asserted for them."""
    return {
        'xml_homo_null_envmap': dict(variant='scalar_rgb', file=os.path.join(PM_DIR, 'test_homo_null_envmap_volpath.xml'), spp=2, seed=0, max_flips=0.005),
        'xml_homo_dielectric_envmap': dict(variant='scalar_rgb', file=os.path.join(PM_DIR, 'test_homo_dielectric_envmap_volpath.xml'), spp=2, seed=0, max_flips=0.005),
        'xml_homo_roughdielectric_spot': dict(variant='scalar_rgb', file=os.path.join(PM_DIR, 'test_homo_roughdielectric_spot_volpath.xml'), spp=2, seed=0, max_flips=0.005),
        'xml_homo_area_light': dict(variant='scalar_rgb', file=os.path.join(PM_DIR, 'test_homo_area_light_volpath.xml'), spp=2, seed=0, max_flips=0.005),
        'xml_homo_obj_inside_medium': dict(variant='scalar_rgb', file=os.path.join(PM_DIR, 'test_homo_obj_inside_medium_volpath.xml'), spp=2, seed=0, max_flips=0.005),
        'xml_surf_area_light': dict(variant='scalar_rgb', file=os.path.join(PM_DIR, 'test_surf_area_light_volpath.xml'), spp=2, seed=0, max_flips=0.005),
        'xml_homo_spectral_sigmat_mis': dict(variant='scalar_rgb', file=os.path.join(PM_DIR, 'test_homo_spectral_sigmat_volpathmis.xml'), spp=2, seed=0, max_flips=0.005),
        'xml_hetero_spectral_sigmat_mis': dict(variant='scalar_rgb', file=os.path.join(PM_DIR, 'test_hetero_spectral_sigmat_volpathmis.xml'), spp=2, seed=0, max_flips=0.05),
        'xml_hetero_sigmat': dict(variant='scalar_rgb', file=os.path.join(PM_DIR, 'test_hetero_sigmat_volpath.xml'), spp=2, seed=0, max_flips=0.05),
    }


def fd_scene(albedo, integrator='volpathaos', spp=64, sigma_t=0.6):
    d = scenes.c1_homogeneous_cube(res=1, spp=spp, max_depth=12)
    d['integrator'] = {'type': integrator, 'max_depth': 12, 'rr_depth': 1000}
    d['medium1']['sigma_t'] = sigma_t
    d['medium1']['albedo'] = albedo
    d['medium1']['phase'] = {'type': 'rayleigh'}
    d['emitter'] = {'type': 'directional', 'direction': [0.3, -0.2, -1.0], 'irradiance': 1.0}
    d['sensor']['film']['pixel_format'] = 'luminance'
    return d


def fd_grid(sign=0, h=0.05):
    """5^3 density grid, its perturbation direction v, and grid + sign * h * v; voxel (0, 0, 0) pins the majorant."""
    rng = np.random.default_rng(3)
    grid = rng.uniform(0.2, 0.8, size=(5, 5, 5, 1)).astype(np.float32)
    grid[0, 0, 0, 0] = 1.0
    v = rng.uniform(0.0, 1.0, size=grid.shape).astype(np.float32)
    v[0, 0, 0, 0] = 0
    return (grid + np.float32(sign * h) * v).astype(np.float32), v


def fd_grid_scene(sign=0, integrator='volpathaos', spp=64):
    g, _ = fd_grid(sign)
    d = scenes.c3_cloud(res=1, spp=spp, max_depth=12, integrator=integrator, grid=g, optical_depth=4.0, albedo=0.9, phase={'type': 'rayleigh'})
    d['integrator']['rr_depth'] = 1000
    d['object']['interior']['scale'] = 2.0
    d['sensor']['film']['pixel_format'] = 'luminance'
    d['emitter'] = {'type': 'directional', 'direction': [0.3, -0.2, -1.0], 'irradiance': 1.0}
    return d


def cases():
    c = {}
    # C1 (BASELINE configs[0]) at 32 x 32 x 16: homogeneous HG cube, volpath, rgb
    c['c1_rgb'] = dict(variant='scalar_rgb', scene=scenes.c1_homogeneous_cube(res=32, spp=16), spp=16, seed=0, kind='samples')
    c['c1_mono_seed3'] = dict(variant='scalar_mono', scene=scenes.c1_homogeneous_cube(res=8, spp=8, max_depth=12), spp=8, seed=3, kind='samples')
    # C2 (configs[1]): polarized Rayleigh slab, volpathaos, distant sensor: render_1 at 2e4 samples + the first 2048 samples
    c['c2_pol_aos_render1'] = dict(variant='scalar_mono_polarized', scene=scenes.c2_rayleigh_slab(spp=4, depolarization=0.03), spp=20000, seed=1, kind='render1')
    c['c2_pol_aos'] = dict(variant='scalar_mono_polarized', scene=scenes.c2_rayleigh_slab(spp=4, depolarization=0.03), spp=2048, seed=1, kind='samples')
    c['c2_pol_volpath_sky'] = dict(variant='scalar_mono_polarized', scene=scenes.c2_rayleigh_slab(spp=4, integrator='volpath', sun=False), spp=1024, seed=2, kind='samples')
    c['c2_mono_aos'] = dict(variant='scalar_mono', scene=scenes.c2_rayleigh_slab(spp=4), spp=1024, seed=4, kind='samples')
    # heterogeneous 32^3 grid, tabulated polarized phase (the C3 / C4 plugin set), volpath, polarized
    d = scenes.c3_cloud(res=12, spp=8, grid=cloud_no_plateau(32), phase=polarized_table(), optical_depth=6.0)
    c['het_pol_tab'] = dict(variant='scalar_mono_polarized', scene=d, spp=8, seed=0, kind='samples')
    cost, val = scenes.hg_table(0.8, 257)
    d = scenes.c3_cloud(res=16, spp=8, grid=cloud_no_plateau(32), phase={'type': 'tabphase', 'cost': cost, 'values': val}, optical_depth=8.0)
    c['het_mono_tab'] = dict(variant='scalar_mono', scene=d, spp=8, seed=5, kind='samples')
    d = scenes.c3_cloud(res=10, spp=4, grid=_grid(10, 2), phase={'type': 'rayleigh'}, optical_depth=6.0, integrator='volpathaos')
    d['emitter'] = {'type': 'directional', 'direction': [0.3, -0.2, -1.0], 'irradiance': 1.0}
    c['het_pol_aos_sun'] = dict(variant='scalar_mono_polarized', scene=d, spp=4, seed=0, kind='samples')
    # rgb: 3-channel albedo grid + spectral extinction through a 3-channel sigma_t scale
    d = scenes.c3_cloud(res=10, spp=8, grid=_grid(8, 3), phase={'type': 'hg', 'g': 0.5}, optical_depth=5.0)
    d['object']['interior']['albedo'] = {'type': 'gridvolume', 'data': _grid(6, 4, lo=0.3, channels=3),
                                         'to_world': T.translate([-1, -1, -1]) @ T.scale(2.0)}
    c['het_rgb_albedo_grid'] = dict(variant='scalar_rgb', scene=d, spp=8, seed=1, kind='samples')
    d = scenes.c1_homogeneous_cube(res=10, spp=8, max_depth=16)
    d['medium1']['sigma_t'] = {'type': 'rgb', 'value': [0.3, 0.6, 0.9]}     # (the reference refuses rgb values above 1)
    d['medium1']['scale'] = 3.0
    d['medium1']['has_spectral_extinction'] = True
    c['homog_rgb_spectral_extinction'] = dict(variant='scalar_rgb', scene=d, spp=8, seed=2, kind='samples')
    # plugin coverage: sphere container, isotropic phase, hide_emitters, nearest filter, repeat / mirror wrap
    d = scenes.c1_homogeneous_cube(res=10, spp=8, max_depth=16)
    d['object'] = {'type': 'sphere', 'center': [0.1, -0.1, 0.0], 'radius': 1.2, 'bsdf': {'type': 'null'}, 'interior': {'type': 'ref', 'id': 'medium1'}}
    d['medium1']['phase'] = {'type': 'isotropic'}
    c['sphere_isotropic'] = dict(variant='scalar_rgb', scene=d, spp=8, seed=0, kind='samples')
    d = scenes.c1_homogeneous_cube(res=10, spp=8, max_depth=16)
    d['integrator']['hide_emitters'] = True
    c['hide_emitters'] = dict(variant='scalar_rgb', scene=d, spp=8, seed=0, kind='samples')
    for wrap in ('repeat', 'mirror'):
        d = scenes.c3_cloud(res=8, spp=8, grid=_grid(6, 5), phase={'type': 'isotropic'}, optical_depth=5.0)
        v = d['object']['interior']['sigma_t']
        v['wrap_mode'] = wrap
        v['to_world'] = T.translate([-0.4, -0.5, -0.3]) @ T.scale(0.7)     # the cube spans several periods of the texture
        d['object']['interior']['scale'] = 2.0
        c['wrap_' + wrap] = dict(variant='scalar_mono', scene=d, spp=8, seed=0, kind='samples')
    d = scenes.c3_cloud(res=8, spp=8, grid=_grid(6, 6), phase={'type': 'isotropic'}, optical_depth=5.0)
    d['object']['interior']['sigma_t']['filter_type'] = 'nearest'
    c['filter_nearest'] = dict(variant='scalar_mono', scene=d, spp=8, seed=0, kind='samples')
    # area emitter on a rectangle above the cube (NEE + MIS of a hit), diffuse floor
    d = scenes.c1_homogeneous_cube(res=10, spp=8, max_depth=12)
    del d['emitter']
    d['lamp'] = {'type': 'rectangle', 'to_world': T.translate([0, 0, 3.0]) @ T.rotate([1, 0, 0], 180.0) @ T.scale(0.8),
                 'emitter': {'type': 'area', 'radiance': {'type': 'rgb', 'value': [6.0, 5.0, 4.0]}}}
    d['floor'] = {'type': 'rectangle', 'to_world': T.translate([0, 0, -1.5]) @ T.scale(4.0),
                  'bsdf': {'type': 'diffuse', 'reflectance': {'type': 'rgb', 'value': [0.5, 0.6, 0.7]}}}
    c['area_emitter'] = dict(variant='scalar_rgb', scene=d, spp=8, seed=0, kind='samples')
    # the coupled atmosphere-ocean system: rough / smooth dielectric ocean surface under a Rayleigh atmosphere,
    # scattering ocean below, refractive NEE (volpathaos.cpp:232-290,312-337)
    c['ocean_rough_pol'] = dict(variant='scalar_mono_polarized', scene=scenes.aos_ocean(rough=True), spp=1024, seed=3, kind='samples')
    c['ocean_rough_ggx_mono'] = dict(variant='scalar_mono', scene=scenes.aos_ocean(rough=True, distribution='ggx', alpha=0.3), spp=512, seed=5, kind='samples')
    c['ocean_smooth_pol'] = dict(variant='scalar_mono_polarized', scene=scenes.aos_ocean(rough=False), spp=1024, seed=3, kind='samples')
    c['ocean_smooth_mono'] = dict(variant='scalar_mono', scene=scenes.aos_ocean(rough=False, view_zenith_deg=55.0), spp=512, seed=1, kind='samples')
    c['ocean_rough_pol_render1'] = dict(variant='scalar_mono_polarized', scene=scenes.aos_ocean(rough=True), spp=20000, seed=2, kind='render1')
    # triangle meshes (tests/golden/meshes, written by tests/golden/make_meshes.py): a box of quads with
    # per-face normals bounding a homogeneous medium, a smooth uv-sphere, a textured ground quad with a checkerboard, a box
    # without normals; perspective sensor
    c['mesh_medium'] = dict(variant='scalar_rgb', scene=mesh_scene(), spp=16, seed=0, kind='samples')
    c['mesh_hetero_mono'] = dict(variant='scalar_mono', scene=mesh_scene(hetero=True, res=12), spp=8, seed=3, kind='samples')
    d = mesh_scene(res=12)
    d['integrator']['type'] = 'volpathaos'
    d['emitter'] = {'type': 'directional', 'direction': [0.3, 0.2, -1.0], 'irradiance': 2.0}
    d['medium1']['phase'] = {'type': 'rayleigh'}
    c['mesh_aos_pol_sun'] = dict(variant='scalar_mono_polarized', scene=d, spp=16, seed=5, kind='samples')
    # environment map (src/emitters/envmap.cpp): importance sampling through the luminance warp, MIS of rays that leave the scene
    tex = os.path.join(os.path.dirname(MESH_DIR), "textures", "sky.exr")
    d = scenes.c1_homogeneous_cube(res=12, spp=16, max_depth=8)
    d['emitter'] = {'type': 'envmap', 'filename': tex}
    c['envmap_homog_rgb'] = dict(variant='scalar_rgb', scene=d, spp=16, seed=0, kind='samples')
    d = mesh_scene(hetero=True, res=12)
    d['emitter'] = {'type': 'envmap', 'filename': tex, 'scale': 1.5, 'to_world': T.rotate([1, 0, 0], 90.0) @ T.rotate([0, 1, 0], 40.0)}
    c['envmap_mesh_mono'] = dict(variant='scalar_mono', scene=d, spp=8, seed=2, kind='samples')
    # volpathmis (src/integrators/volpathmis.cpp with use_spectral_mis): spectrally varying and grey
    # extinction, heterogeneous medium with an albedo grid, area emitter + diffuse floor (surface NEE / emitter-hit MIS), mono
    import copy
    for k in ('c1_rgb', 'homog_rgb_spectral_extinction', 'het_rgb_albedo_grid', 'area_emitter', 'c1_mono_seed3', 'envmap_homog_rgb'):
        m = copy.deepcopy(c[k])
        m['scene']['integrator']['type'] = 'volpathmis'
        if k == 'c1_rgb':
            m['scene'] = scenes.c1_homogeneous_cube(res=12, spp=16)
            m['scene']['integrator']['type'] = 'volpathmis'
        c['mis_' + k] = m
    # volpathmis with use_spectral_mis = false (volpathmis.cpp:519-525,539-542,558-560): one weight spectrum instead of the matrices
    for k in ('homog_rgb_spectral_extinction', 'het_rgb_albedo_grid', 'area_emitter'):
        m = copy.deepcopy(c[k])
        m['scene']['integrator'].update(type='volpathmis', use_spectral_mis=False)
        c['mis_plain_' + k] = m
    # plastic / roughplastic (src/bsdfs/plastic.cpp, roughplastic.cpp): a plastic floor under the lamp of `area_emitter`, a
    # rough plastic ball beside the cube (ggx and beckmann, with and without a specular tint, nonlinear colour shift)
    def plastic_scene(integrator='volpath', distribution='beckmann', alpha=0.1, tint=False, nonlinear=False):
        d = copy.deepcopy(c['area_emitter']['scene'])
        d['integrator']['type'] = integrator
        d['floor']['bsdf'] = {'type': 'plastic', 'diffuse_reflectance': {'type': 'rgb', 'value': [0.5, 0.3, 0.2]}, 'nonlinear': nonlinear}
        d['ball'] = {'type': 'sphere', 'center': [1.6, 0.3, -0.6], 'radius': 0.7,
                     'bsdf': {'type': 'roughplastic', 'distribution': distribution, 'alpha': alpha, 'nonlinear': nonlinear,
                              'diffuse_reflectance': {'type': 'rgb', 'value': [0.2, 0.7, 0.7]}}}
        if tint:
            d['floor']['bsdf']['specular_reflectance'] = {'type': 'rgb', 'value': [0.9, 0.8, 0.6]}
            d['ball']['bsdf']['specular_reflectance'] = {'type': 'rgb', 'value': [0.7, 0.9, 0.8]}
            d['ball']['bsdf']['int_ior'] = 1.9
        return d
    c['plastic_rgb'] = dict(variant='scalar_rgb', scene=plastic_scene(), spp=16, seed=0, kind='samples')
    c['plastic_ggx_tint_nonlinear_rgb'] = dict(variant='scalar_rgb', scene=plastic_scene(distribution='ggx', alpha=0.35, tint=True, nonlinear=True), spp=16, seed=1, kind='samples')
    c['plastic_mono_mis'] = dict(variant='scalar_mono', scene=plastic_scene(integrator='volpathmis', alpha=0.25), spp=16, seed=2, kind='samples')
    c['plastic_pol'] = dict(variant='scalar_mono_polarized', scene=plastic_scene(alpha=0.2), spp=16, seed=3, kind='samples')
    # the four active rough-dielectric configurations of the reference's own adjoint test (test_ad_volpathaos.py:235-338: a
    # directly visible rough dielectric plane, `sample_visible = false`, lit by an area light or by the sun), primal side:
    # volpathaos in the polarized variant, radiancemeter, max_depth 3
    def ad_config(directional, alpha=0.1, distribution='beckmann', sample_visible=False):
        d = {'type': 'scene',
             'integrator': {'type': 'volpathaos', 'max_depth': 3},
             'plane': {'type': 'rectangle',
                       'bsdf': {'type': 'roughdielectric', 'int_ior': 1.34, 'ext_ior': 1.0, 'distribution': distribution,
                                'alpha': alpha, 'sample_visible': sample_visible}},
             'sensor': {'type': 'radiancemeter', 'to_world': T.look_at(origin=[0, 0, 4], target=[0, 0, 0], up=[0, 1, 0]),
                        'film': {'type': 'hdrfilm', 'rfilter': {'type': 'box'}, 'width': 1, 'height': 1, 'sample_border': False,
                                 'pixel_format': 'luminance', 'component_format': 'float32'},
                        'sampler': {'type': 'independent', 'sample_count': 512}}}
        if directional:
            d['emitter'] = {'type': 'directional', 'direction': [-0.5, 0, -0.866], 'irradiance': 1.0}
        else:
            d['light'] = {'type': 'rectangle', 'to_world': T.translate([1.25, 0.0, 1.0]) @ T.rotate([0, 1, 0], -90),
                          'emitter': {'type': 'area', 'radiance': {'type': 'rgb', 'value': [3.0, 3.0, 3.0]}}}
        return d
    c['ad_rough_dielectric_area'] = dict(variant='scalar_mono_polarized', scene=ad_config(False), spp=512, seed=0, kind='samples')
    c['ad_rough_dielectric_directional'] = dict(variant='scalar_mono_polarized', scene=ad_config(True), spp=512, seed=1, kind='samples')
    c['ad_rough_dielectric_ggx_area_mono'] = dict(variant='scalar_mono', scene=ad_config(False, alpha=0.3, distribution='ggx'), spp=512, seed=2, kind='samples')
    # finite differences of the REFERENCE for the adjoint tests (the methodology of test_ad_volpathaos.py:504-529, with the medium
    # configuration that file defines but leaves switched off, :340-437): render_1 of the unmodified volpathaos at albedo 0.8 -/+
    # 0.01, one random stream (common random numbers), sun-lit homogeneous Rayleigh cube, polarized.  (volpath itself renders a
    # directional light black in this fork: its shadow ray is spawned toward a point at infinity, volpath.cpp:352 with
    # directional.cpp's ds.dist = inf -- reproduced by the oracle, and the reason the fork has volpathaos.)
    for tag, albedo in (('minus', 0.79), ('plus', 0.81)):
        c['fd_ref_albedo_' + tag] = dict(variant='scalar_mono_polarized', scene=fd_scene(albedo), spp=200000, seed=1, kind='render1')
    for tag, sign in (('minus', -1), ('plus', +1)):
        c['fd_ref_grid_' + tag] = dict(variant='scalar_mono_polarized', scene=fd_grid_scene(sign), spp=400000, seed=1, kind='render1')
    for tag, sigma_t in (('minus', 0.57), ('plus', 0.63)):
        c['fd_ref_sigma_t_' + tag] = dict(variant='scalar_mono_polarized', scene=fd_scene(0.8, sigma_t=sigma_t), spp=200000, seed=1, kind='render1')
    # radiancemeter looking through the cube
    d = scenes.c1_homogeneous_cube(res=1, spp=64, max_depth=16)
    d['sensor'] = {'type': 'radiancemeter', 'origin': [0.2, 0.1, 4.0], 'direction': [-0.05, 0.0, -1.0],
                   'film': {'type': 'hdrfilm', 'width': 1, 'height': 1, 'rfilter': {'type': 'box'}},
                   'sampler': {'type': 'independent', 'sample_count': 64}}
    c['radiancemeter'] = dict(variant='scalar_rgb', scene=d, spp=64, seed=0, kind='samples')
    return c
