"""Short driver for the C4 adjoint: python profiles/adj_run.py [spp] [grid_res]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import mitsuba3_nasa_b200 as mi
from mitsuba3_nasa_b200 import scenes
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 16
gr = int(sys.argv[2]) if len(sys.argv) > 2 else 256
pol = (sys.argv[3] != 'mono') if len(sys.argv) > 3 else True
mi.set_variant('llvm_ad_mono_polarized' if pol else 'llvm_ad_mono')
scene = mi.load_dict(scenes.c4_cloud_adjoint(spp=spp, grid_res=gr, polarized=pol, majorant_resolution=int(os.environ.get('MAJORANT_RES', '0'))))
dev = scene.device
for k in range(2):
    dev.grad_zero()
    dev.render_1_backward(k, spp, [1.0, 0, 0, 0])
    c = dev.counters()
    print('c4 %s spp %d: %.1f ms %.2f Msamples/s lookups/sample %.1f' % ('pol' if pol else 'mono', spp, c['kernel_ms'], c['n_samples'] / c['kernel_ms'] / 1e3, c['n_lookups'] / c['n_samples']))
slot = scene.flat.desc.media[0].sigma_t.buffer
g = dev.grad_get(slot, scene.flat.buffers[slot].size)
print('grad abs sum', np.abs(g.astype(np.float64)).sum(), 'max', np.abs(g).max(), 'nan', np.isnan(g).sum())
