"""C5: sample-count sweep on the C3 Mie cloud (BASELINE configs[4]): python profiles/c5_sweep.py [max_log10=9] [variant=mono|pol] [majorant_res=0]

Renders the 1024x1024 C3 scene at 1e7, 1e8, ... samples (spp = samples / 2^20, rounded), in passes of at most 2^32 lanes with
seed + pass as the reference does (integrator.cpp:566-578), and prints one JSON line per sample count: device time of the
walk kernels (CUDA events inside the library), Msamples/s, look-ups per sample, mean radiance (converges as 1/sqrt(N)).
Under torchrun every rank renders its lane slice and rank 0 reports the max-over-ranks time (weak data, strong scaling here:
the total work is fixed).
"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import mitsuba3_nasa_b200 as mi
from mitsuba3_nasa_b200 import scenes

max_log = int(sys.argv[1]) if len(sys.argv) > 1 else 9
pol = len(sys.argv) > 2 and sys.argv[2] == 'pol'
mr = int(sys.argv[3]) if len(sys.argv) > 3 else 0
rank, world = int(os.environ.get('RANK', 0)), int(os.environ.get('WORLD_SIZE', 1))
local = int(os.environ.get('LOCAL_RANK', 0))
if world > 1:
    import torch, torch.distributed as dist
    torch.cuda.set_device(local)
    dist.init_process_group('nccl', device_id=torch.device('cuda', local))
mi.set_variant('llvm_ad_mono_polarized' if pol else 'llvm_ad_mono')
res = 1024
scene = mi.load_dict(scenes.c3_cloud(res=res, spp=16, grid_res=256, majorant_resolution=mr), device=local)
dev = scene.device
dev.render(123, 1, develop=False, to_host=False)                     # warm-up
for lg in range(7, max_log + 1):
    spp_total = max(1, round(10 ** lg / (res * res)))
    max_spp_pass = (1 << 32) // (res * res) - 1                       # < 2^32 lanes per pass
    passes = [max_spp_pass] * (spp_total // max_spp_pass) + ([spp_total % max_spp_pass] if spp_total % max_spp_pass else [])
    ms, lookups, n, acc = 0.0, 0, 0, None
    t0 = time.perf_counter()
    for p, spp in enumerate(passes):
        if world > 1:
            dev.set_lane_stripes(rank, world, 64 * spp)                   # interleaved stripes of 64 pixels: every rank samples the whole image
        raw = dev.render(p, spp, develop=False, to_host=(lg <= 8))     # seed + pass
        c = dev.counters()
        ms += c['kernel_ms']; lookups += c['n_lookups']; n += c['n_samples']
        if lg <= 8:
            acc = np.asarray(raw, dtype=np.float64) if acc is None else acc + raw
    wall = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([ms, float(n), float(lookups)], dtype=torch.float64, device='cuda')
        tm = t.clone(); dist.all_reduce(tm, op=dist.ReduceOp.MAX); dist.all_reduce(t, op=dist.ReduceOp.SUM)
        ms, n, lookups = float(tm[0]), float(t[1]), float(t[2])
    if rank == 0:
        line = {"config": "C5", "variant": 'mono_polarized' if pol else 'mono', "majorant_resolution": mr, "n_gpus": world,
                "samples": int(n), "spp": spp_total, "passes": len(passes), "kernel_ms": ms, "msamples_per_s": n / ms / 1e3,
                "lookups_per_sample": lookups / n, "wall_s": wall}
        if acc is not None and world == 1:
            line["mean_radiance"] = float(acc[..., 0].sum() / max(acc[..., -1].sum() if not pol else n, 1))
        print(json.dumps(line), flush=True)
if world > 1:
    dist.destroy_process_group()
