"""Tiny adjoint launch for debugging: python profiles/adj_debug.py [res] [spp] [grid] [mono|pol]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import mitsuba3_nasa_b200 as mi
from mitsuba3_nasa_b200 import scenes
res = int(sys.argv[1]) if len(sys.argv) > 1 else 8
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 1
gr = int(sys.argv[3]) if len(sys.argv) > 3 else 8
pol = (sys.argv[4] != 'mono') if len(sys.argv) > 4 else False
mi.set_variant('llvm_ad_mono_polarized' if pol else 'llvm_ad_mono')
d = scenes.c3_cloud(res=res, spp=spp, grid=scenes.cloud_grid(gr), phase={'type': 'hg', 'g': 0.5}, optical_depth=4.0, max_depth=8,
                    integrator='prb_volpathaos' if pol else 'prbvolpath')
scene = mi.load_dict(d)
dev = scene.device
print('scene ready', flush=True)
dev.grad_zero()
dev.render_1_backward(0, spp, [1.0, 0, 0, 0])
c = dev.counters()
print('OK', {k: c[k] for k in ('n_samples', 'n_real', 'n_null', 'kernel_ms')}, flush=True)
