import sys; sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import numpy as np
from mitsuba3_nasa_b200 import scene as S, scenes, mie as M
from test_mie import oracle_tabulate
from orc import OracleScene
d = M.LogNormalSizeDistr({'mean_radius': 10.0, 'std': 1.4, 'min_radius': 1.0, 'max_radius': 40.0, 'gauss_points': 100})
cost = np.linspace(-1,1,1801).astype(np.float32)
m11,m12,m33,m34 = oracle_tabulate(cost, 0.55, d)
print('m11 range', m11.min(), m11.max(), '2pi int', 2*np.pi*np.trapezoid(m11.astype(float), cost.astype(float)))
print('m12/m11 range', (m12/m11).min(), (m12/m11).max(), 'm33/m11', (m33/m11).min(), (m33/m11).max(), 'm34/m11', (m34/m11).min(), (m34/m11).max())
np.savez('/tmp/mie_c3.npz', cost=cost, m11=m11, m12=m12, m33=m33, m34=m34)
phase = {'type':'tabphase_polarized','cost':cost,'m11':m11,'m12':m12,'m33':m33,'m34':m34}
for integ in ('volpathaos','prb_volpathaos'):
    dd = scenes.c3_cloud(res=16, spp=16, grid_res=32, max_depth=64, integrator=integ, phase=phase)
    sc = S.load_dict(dd, 'llvm_ad_mono_polarized')
    o = OracleScene(sc)
    lanes = o.sample_lanes(0, 16, 0, 4096)
    print(integ, 'lanes max', np.abs(lanes).max(axis=0), 'mean', lanes.mean(axis=0), 'nan', np.isnan(lanes).sum())
