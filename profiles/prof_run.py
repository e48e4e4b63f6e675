"""Short single-GPU driver for ncu captures: python profiles/prof_run.py {c2|c3|c1} [samples_log2]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import mitsuba3_nasa_b200 as mi
from mitsuba3_nasa_b200 import scenes

which = sys.argv[1] if len(sys.argv) > 1 else 'c2'
MR = int(os.environ.get('MAJORANT_RES', '0'))      # fast mode: majorant supergrid resolution (0 = parity mode)
lg = int(sys.argv[2]) if len(sys.argv) > 2 else 24
if which == 'c2':
    mi.set_variant('llvm_ad_mono_polarized')
    scene = mi.load_dict(scenes.c2_rayleigh_slab(spp=4))
    for k in range(3):
        out = scene.device.render_1(k, 1 << lg)
        c = scene.device.counters()
        print('c2', out, c['kernel_ms'], 'ms', c['n_samples'] / c['kernel_ms'] / 1e3, 'Msamples/s')
elif which == 'c3':
    mi.set_variant('llvm_ad_mono')
    cost, val = scenes.hg_table(0.85, 1801)
    d = scenes.c3_cloud(res=512, spp=max(1, (1 << lg) // (512 * 512)), grid_res=256, phase={'type': 'tabphase', 'cost': cost, 'values': val}, majorant_resolution=MR)
    scene = mi.load_dict(d)
    for k in range(3):
        scene.device.render(k, 0 or d['sensor']['sampler']['sample_count'], develop=False, to_host=False)
        c = scene.device.counters()
        print('c3', c['kernel_ms'], 'ms', c['n_samples'] / c['kernel_ms'] / 1e3, 'Msamples/s', c)
elif which == 'c3pol':
    # BASELINE C3 as specified: polarized, Mie x lognormal tabulated phase
    mi.set_variant('llvm_ad_mono_polarized')
    d = scenes.c3_cloud(res=512, spp=max(1, (1 << lg) // (512 * 512)), grid_res=256, majorant_resolution=MR)
    scene = mi.load_dict(d)
    for k in range(3):
        scene.device.render(k, d['sensor']['sampler']['sample_count'], develop=False, to_host=False)
        c = scene.device.counters()
        print('c3pol', c['kernel_ms'], 'ms', c['n_samples'] / c['kernel_ms'] / 1e3, 'Msamples/s', c)
else:
    mi.set_variant('llvm_ad_rgb')
    scene = mi.load_dict(scenes.c1_homogeneous_cube())
    for k in range(3):
        scene.device.render(k, 64, to_host=False)
        c = scene.device.counters()
        print('c1', c['kernel_ms'], 'ms', c['n_samples'] / c['kernel_ms'] / 1e3, 'Msamples/s')

if len(sys.argv) > 3 and sys.argv[3] == 'check':
    # pool kernel vs one-thread-one-path kernel at this size: per-lane results are schedule independent, so the event
    # counters must agree exactly and the film / Stokes sums up to the order of the float atomics
    res = {}
    for k in ('pool', 'thread'):
        os.environ['MIB_FORWARD_KERNEL'] = k
        if which == 'c2':
            out = scene.device.render_1(7, 1 << lg)
        else:
            out = np.asarray(scene.device.render(7, d['sensor']['sampler']['sample_count'] if which != 'c1' else 64, develop=False, to_host=True), dtype=np.float64)
        c = scene.device.counters()
        res[k] = (out, {kk: c[kk] for kk in ('n_samples', 'n_lookups', 'n_real', 'n_null', 'n_nee_steps', 'n_surface')})
    same = res['pool'][1] == res['thread'][1]
    err = np.abs(res['pool'][0] - res['thread'][0]).max() / max(np.abs(res['thread'][0]).max(), 1e-30)
    print('CHECK', which, 'counters identical' if same else 'COUNTERS DIFFER %s %s' % (res['pool'][1], res['thread'][1]), 'max rel diff of sums %.3g' % err)
