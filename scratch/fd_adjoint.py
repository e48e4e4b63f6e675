import sys, time, copy
import numpy as np
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
from mitsuba3_nasa_b200 import scene as S, scenes, _abi as A
from mitsuba3_nasa_b200 import transforms as T
import orc

def homog(variant, integ, sigma_t=1.5, albedo=0.8, res=8, spp=1024, phase=None):
    d = scenes.c1_homogeneous_cube(res=res, spp=spp, max_depth=16)
    d['integrator'] = {'type': integ, 'max_depth': 16, 'rr_depth': 1000}
    d['medium1']['sigma_t'] = sigma_t
    d['medium1']['albedo'] = albedo
    if phase: d['medium1']['phase'] = phase
    if 'mono' in variant:
        d['sensor']['film']['pixel_format'] = 'luminance'
    if integ == 'prb_volpathaos':
        d['emitter'] = {'type': 'directional', 'direction': [0.3, -0.2, -1.0], 'irradiance': 1.0}
    return S.load_dict(d, variant)

def run(variant, integ, phase=None):
    print('==', variant, integ, phase)
    spp = 4096
    f = homog(variant, integ, spp=spp, phase=phase)
    o = orc.OracleScene(f)
    g4 = [1.0, 0, 0, 0]
    t = time.time()
    o.render_1_backward(7, spp, g4, nthreads=16)
    gs = o.grad_get_scalar(0)
    print(' adjoint d/dalbedo', gs[:3], 'd/dsigma_t', gs[3], 'time', time.time() - t)
    base = o.render_1(7, spp, nthreads=16)[0]
    # FD albedo (same seed => exact derivative of the estimator for prbvolpath)
    for key, idx, h in (('albedo', 0, 1e-2), ('sigma_t', 3, 2e-2)):
        vals = []
        for sgn in (+1, -1):
            f2 = homog(variant, integ, spp=spp, phase=phase,
                       albedo=0.8 + (sgn * h if key == 'albedo' else 0), sigma_t=1.5 + (sgn * h if key == 'sigma_t' else 0))
            o2 = orc.OracleScene(f2)
            vals.append(o2.render_1(7, spp, nthreads=16)[0])
        fd = (vals[0] - vals[1]) / (2 * h)
        print('  FD', key, fd, ' adjoint', gs[idx], ' ratio', gs[idx] / fd if fd else None, 'base', base)

run('llvm_ad_mono', 'prbvolpath')
run('llvm_ad_mono_polarized', 'prb_volpathaos', {'type': 'rayleigh'})
