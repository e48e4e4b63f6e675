"""Load balance of the lane stripes, measured on ONE GPU: the C4 adjoint (512 x 512 x spp) is run once per emulated rank with
mib_scene_set_lane_stripes(rank, N, chunk) and the kernel times are compared.  efficiency = t(1 rank) / (N * max_r t(r)).

    python profiles/dist_balance.py [spp] [chunk_pixels ...]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import mitsuba3_nasa_b200 as mi
from mitsuba3_nasa_b200 import scenes

spp = int(sys.argv[1]) if len(sys.argv) > 1 else 16
chunks = [int(a) for a in sys.argv[2:]] or [64]
mi.set_variant('llvm_ad_mono_polarized')
scene = mi.load_dict(scenes.c4_cloud_adjoint(spp=spp, grid_res=256, polarized=True))
dev = scene.device


def run():
    dev.grad_zero()
    dev.render_1_backward(0, spp, [1.0, 0, 0, 0])
    c = dev.counters()
    return c['kernel_ms'], c['n_samples']


run()
t1, n1 = run()
print("1 rank: %.2f ms, %d samples" % (t1, n1))
for cp in chunks:
    for N in (2, 4, 8):
        ts = []
        for r in range(N):
            dev.set_lane_stripes(r, N, cp * spp)
            t, n = run()
            ts.append(t)
        dev.set_lane_stripes(0, 1, 1)
        print("chunk %3d px, N=%d: per-rank ms min %.2f max %.2f mean %.2f -> efficiency %.3f (ideal-balance efficiency %.3f)"
              % (cp, N, min(ts), max(ts), np.mean(ts), t1 / (N * max(ts)), t1 / (N * np.mean(ts))))
