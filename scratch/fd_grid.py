import sys, time
import numpy as np
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
from mitsuba3_nasa_b200 import scene as S, scenes, _abi as A
import orc

def mk(variant, integ, grid, spp, albedo=0.9):
    d = scenes.c3_cloud(res=4, spp=spp, max_depth=16, integrator=integ, grid=grid, optical_depth=4.0, albedo=albedo,
                        phase={'type': 'hg', 'g': 0.3} if 'polarized' not in variant else {'type': 'rayleigh'})
    d['integrator']['rr_depth'] = 1000
    d['object']['interior']['scale'] = 2.0
    if 'polarized' in variant:
        d['emitter'] = {'type': 'directional', 'direction': [0.3, -0.2, -1.0], 'irradiance': 1.0}
    return S.load_dict(d, variant)

rng = np.random.default_rng(3)
n = 6
grid = rng.uniform(0.2, 0.8, size=(n, n, n, 1)).astype(np.float32)
grid[0, 0, 0, 0] = 1.0   # pins the majorant
v = rng.uniform(0.0, 1.0, size=grid.shape).astype(np.float32)
v[0, 0, 0, 0] = 0
for variant, integ in (('llvm_ad_mono', 'prbvolpath'), ('llvm_ad_mono_polarized', 'prb_volpathaos')):
    spp = 1 << 18
    f = mk(variant, integ, grid, spp)
    o = orc.OracleScene(f)
    slot = [k for k, a in f.buffers.items() if a.size == grid.size][0]
    t = time.time()
    o.render_1_backward(5, spp, [1, 0, 0, 0], nthreads=16)
    g = o.grad_get(slot, grid.size)
    h = 0.05
    vals = [orc.OracleScene(mk(variant, integ, grid + s * h * v, spp)).render_1(9, spp, nthreads=16)[0] for s in (1, -1)]
    fd = (vals[0] - vals[1]) / (2 * h)
    print(variant, integ, 'adj <g,v>', float(g @ v.reshape(-1)), 'FD', fd, 'ratio', float(g @ v.reshape(-1)) / fd, 'scalar', o.grad_get_scalar(0), 'time', time.time() - t)
