import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tests'))
import numpy as np
from mitsuba3_nasa_b200 import scene as S, _lib
from orc import OracleScene
import scalar_cases

def run(tag, d, variant):
    sc = S.load_dict(d, variant)
    dev, orc = _lib.DeviceScene(sc, 0), OracleScene(sc)
    n = 12 * 12 * 16
    g = dev.sample_lanes(5, 16, 0, n).astype(np.float64); o = orc.sample_lanes(5, 16, 0, n).astype(np.float64)
    scale = np.maximum(np.abs(o).max(axis=1, keepdims=True), 1e-6)
    bad = (np.abs(g - o) / scale > 1e-4).any(axis=1)
    print("%-40s %-24s differing %4d  exact %s" % (tag, variant, bad.sum(), np.array_equal(g, o)))
    for i in np.nonzero(bad)[0][:3]:
        print("    lane", i, "gpu", g[i], "oracle", o[i])

def base(phase='rayleigh', integ='volpathaos', sun=True, drop=()):
    d = scalar_cases.mesh_scene(res=12)
    d['integrator']['type'] = integ
    if sun: d['emitter'] = {'type': 'directional', 'direction': [0.3, 0.2, -1.0], 'irradiance': 2.0}
    d['medium1']['phase'] = {'type': phase}
    for k in drop: del d[k]
    return d

run("aos sun rayleigh", base(), "llvm_ad_mono_polarized")
run("aos sun rayleigh", base(), "llvm_ad_mono")
run("aos sun isotropic", base(phase='isotropic'), "llvm_ad_mono_polarized")
run("aos sky rayleigh", base(sun=False), "llvm_ad_mono_polarized")
run("volpath sun rayleigh", base(integ='volpath'), "llvm_ad_mono_polarized")
run("volpath sky rayleigh", base(integ='volpath', sun=False), "llvm_ad_mono_polarized")
run("aos sun rayleigh, no ball", base(drop=('ball',)), "llvm_ad_mono_polarized")
run("aos sun rayleigh, no ball/crate", base(drop=('ball', 'crate')), "llvm_ad_mono_polarized")
run("aos sun rayleigh, container only", base(drop=('ball', 'crate', 'floor')), "llvm_ad_mono_polarized")
