#!/usr/bin/env python
"""bench.py -- headline benchmark of the volumetric hot path (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W            # our arm (CUDA, sm_100a)
    python bench.py --impl reference --gpus N --steps K ...  # CPU reference arm (oracle port)

A "step" is one pass of the hot path over one batch of synthetic samples.
Workload at N=1 = BASELINE.json configs[1]: plane-parallel Rayleigh slab,
mono_polarized, distant sensor, 1e8 samples, Stokes IQUV via render_1
(through volpathaos).  Weak scaling: every rank renders its own
1e8-lane slice of one N*1e8-lane wavefront; the only collective is the
all-reduce of the 4 Stokes sums.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

SAMPLES_PER_GPU = 100_000_000
C2_KW = dict(spp=4, integrator='volpathaos', max_depth=64)
# Algorithmic flop constants per event (counted from the oracle source; DESIGN.md "Roofline")
FLOPS = dict(collision=90, real_pol=700, nee_setup_pol=380, nee_step=90, surface=320, sample=150)
FP32_PEAK_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12   # 74.4: 148 SMs x 128 lanes x FMA at clocks.max.sm


def algorithmic_flops(c):
    return (FLOPS['collision'] * (c['n_real'] + c['n_null'] + c['n_surface']) + FLOPS['real_pol'] * c['n_real'] +
            FLOPS['nee_setup_pol'] * c['n_real'] + FLOPS['nee_step'] * c['n_nee_steps'] +
            FLOPS['surface'] * c['n_surface'] + FLOPS['sample'] * c['n_samples'])


def algorithmic_bytes(c, film_channels=4):
    # 32 B per grid lookup (8 taps x 4 B) + 4*C B per sample of film traffic
    return 32 * c['n_lookups'] + 4 * film_channels * c['n_samples']


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.f = tempfile.NamedTemporaryFile('w+', suffix='.csv', delete=False)
        try:
            self.p = subprocess.Popen(['nvidia-smi', '-i', str(gpu_index), '--query-gpu=' + self.Q,
                                       '--format=csv,noheader,nounits', '-lms', '100'], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        rows = [r.strip().split(', ') for r in open(self.f.name).read().strip().splitlines() if r.strip()]
        os.unlink(self.f.name)
        sm, reasons = [], set()
        for r in rows:
            if len(r) < 9:
                continue
            try:
                sm.append(float(r[1])); out["sm_max_mhz"] = float(r[2])
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.strip().lower().startswith('active'):
                    reasons.add(name)
        if sm:
            out["sm_mhz"] = float(np.median(sm))
            out["samples"] = len(sm)
        out["reasons"] = sorted(reasons)
        return out


_JSON_FD = None


def _capture_stdout():
    """Native libraries (NCCL's version banner, CUDA warnings) write to fd 1: route everything to stderr and keep the
    original stdout for the ONE JSON line of the contract."""
    global _JSON_FD
    if _JSON_FD is None:
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    sys.stdout.flush()
    os.write(_JSON_FD if _JSON_FD is not None else 1, (json.dumps(line) + "\n").encode())


def _lib_handle():
    from mitsuba3_nasa_b200 import _lib
    l = _lib.lib()
    l.mib_sizeof_device_scene.restype = __import__('ctypes').c_size_t
    return l


def dist_env():
    return int(os.environ.get('RANK', 0)), int(os.environ.get('LOCAL_RANK', 0)), int(os.environ.get('WORLD_SIZE', 1))


def time_oracle(scene_flat, n_samples, seed, nthreads):
    sys.path.insert(0, os.path.join(ROOT, 'tests'))
    from orc import OracleScene
    o = OracleScene(scene_flat)
    t0 = time.perf_counter()
    out = o.render_1(seed, n_samples, nthreads=nthreads)
    dt = time.perf_counter() - t0
    return n_samples / dt / 1e6, dt, out


def run_reference(args):
    """CPU arm: the reference's algorithm for this path on the host cores (the oracle port; the
    reference itself needs libLLVM / cmake + external libraries and cannot be built here)."""
    rank, _, world = dist_env()
    if rank != 0:
        return
    from mitsuba3_nasa_b200 import scene as S, scenes
    flat = S.load_dict(scenes.c2_rayleigh_slab(**C2_KW), 'llvm_ad_mono_polarized')
    cores = os.cpu_count() or 1
    n = 250_000          # bounded sample of the 1e8-sample workload per step
    for w in range(args.warmup):
        time_oracle(flat, n, 1000 + w, cores)
    t = []
    for k in range(args.steps):
        v, dt, _ = time_oracle(flat, n, k, cores)
        t.append(dt)
    value = n * args.steps / sum(t) / 1e6
    line = {
        "impl": "reference", "metric": "Msamples/s polarized volpath fwd (render_1, Stokes IQUV)", "value": value,
        "unit": "Msamples/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * sum(t) / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": "C2 plane-parallel Rayleigh slab, mono_polarized, volpathaos, distant sensor, render_1 IQUV",
                   "samples_per_step": n, "full_workload_samples": SAMPLES_PER_GPU},
        "cpu_baseline": {"value": value, "unit": "Msamples/s", "cores": cores, "kind": "port",
                         "sample": "%d samples per step of the 1e8-sample C2 workload, std::thread over lane ranges" % n},
        "e2e": {"value": value, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def run_ours(args):
    import ctypes
    import torch
    rank, local_rank, world = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    import mitsuba3_nasa_b200 as mi
    from mitsuba3_nasa_b200 import scenes, _abi
    mi.set_variant('llvm_ad_mono_polarized')
    scene = mi.load_dict(scenes.c2_rayleigh_slab(**C2_KW), device=local_rank)
    dev = scene.device
    n_total = SAMPLES_PER_GPU * world            # one wavefront of N*1e8 lanes (< 2^32), 1x1 film
    lanes = (rank * SAMPLES_PER_GPU, (rank + 1) * SAMPLES_PER_GPU)
    dev.set_lane_range(*lanes)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allreduce(vals, op="sum"):
        t = torch.tensor(vals, dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM if op == "sum" else dist.ReduceOp.MAX)
        return t.cpu().numpy()

    # ---- device-resident timing: K steps, CUDA events on the launching stream (inside the C ABI) ----
    for w in range(args.warmup):
        dev.render_1(1000 + w, n_total, normalize=False)
    barrier()
    clocks = ClockSampler(local_rank) if rank == 0 else None
    kernel_ms, launches, stokes = 0.0, 0, np.zeros(4)
    cnt_sum = None
    t0 = time.perf_counter()
    for k in range(args.steps):
        s = dev.render_1(k, n_total, normalize=False)
        c = dev.counters()
        kernel_ms += c['kernel_ms']
        launches += c['n_launches']
        stokes = s
        cnt_sum = c if cnt_sum is None else {kk: (cnt_sum[kk] + c[kk]) for kk in c}
    barrier()
    wall = time.perf_counter() - t0
    clk = clocks.stop() if clocks else None
    t_dev = float(allreduce([kernel_ms], "max")[0]) / 1e3
    stokes_all = allreduce(list(stokes)) / float(n_total)             # the film reduce (4 doubles)
    value = n_total * args.steps / t_dev / 1e6

    # ---- end-to-end through the public API with host buffers (desc upload + commit + D2H every step) ----
    params = mi.traverse(scene)
    barrier()
    t0 = time.perf_counter()
    for k in range(args.steps):
        params['a_medium.albedo.value.value'] = 1.0          # host-side parameter write -> H2D of the scene block
        params.update()
        m = mi.render_1(scene, spp=n_total, seed=k)          # returns the 4x4 Mueller / Stokes column on the host
        allreduce([float(x) for x in np.asarray(m)[:, 0]])  # film reduce across ranks (4 doubles)
    barrier()
    t_e2e = float(allreduce([time.perf_counter() - t0], "max")[0])
    e2e = n_total * args.steps / t_e2e / 1e6

    # ---- adjoint (C4), all ranks: lanes sharded, gradients all-reduced over NCCL ----
    try:
        adjoint = adjoint_bench(local_rank, rank, world) if not args.no_adjoint else None
    except Exception as e:
        adjoint = {"error": str(e)}
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    c1 = {kk: cnt_sum[kk] / args.steps for kk in cnt_sum}     # per launch (this rank)
    per_launch_s = kernel_ms / 1e3 / args.steps
    flops = algorithmic_flops(c1)
    achieved_tf = flops / per_launch_s / 1e12
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
    except Exception:
        pass
    line = {
        "metric": "Msamples/s polarized volpath fwd (render_1, Stokes IQUV)", "value": value, "unit": "Msamples/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_dev / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "C2 plane-parallel Rayleigh slab, mono_polarized, volpathaos, distant sensor, render_1 IQUV",
                   "samples_per_gpu": SAMPLES_PER_GPU, "max_depth": 64, "rr_depth": 5, "sharding": "lane range per rank",
                   "l2": "analytic scene: no input buffers to cache, L2 flush not applicable"},
        "stokes": [float(x) for x in stokes_all],
        "wall_s_timed_region": wall,
        "clocks": clk,
        "e2e": {"value": e2e, "unit": "Msamples/s",
                "h2d_bytes_per_step": ctypes.sizeof(_abi.SceneDesc) + int(_lib_handle().mib_sizeof_device_scene()),
                "d2h_bytes_per_step": 4 * 8 + ctypes.sizeof(_abi.Counters)},
        "gpu_launches": int(launches),
        "roofline": {"bound": "fp32", "achieved": achieved_tf, "peak": FP32_PEAK_TFLOPS, "unit": "TFLOP/s",
                     "frac": achieved_tf / FP32_PEAK_TFLOPS, "traffic": int(5466880 * SAMPLES_PER_GPU / float(1 << 24)),
                     "traffic_note": "dram bytes read + written by one k_forward<2,1> launch: 0.77 MB + 4.69 MB at 2^24 samples in the ncu --set full capture "
                                     "(profiles/r01_ncu_i_c2_kforward_thread.txt), scaled to this launch's sample count; the scene is analytic, the "
                                     "writes are local-memory spills (224 B stack frame)",
                     "kernel": "k_forward<mono_polarized, volpathaos> (one thread per path; the library picks the path-pool kernels for heterogeneous media)",
                     "note": "homogeneous medium: no grid traffic, FP32-pipe bound; peak = 148 SM x 128 lanes x 2 x 1.965 GHz "
                             "(derived, not in MEASURED_PEAKS.json); flops = documented per-event constants x exact event counters",
                     "events_per_launch": c1, "hbm_peak_gbs_measured": peaks.get("hbm_gbs")},
    }
    line["adjoint"] = adjoint
    # ---- secondary: heterogeneous grid workload (memory side) on rank 0 ----
    try:
        line["roofline_grid"] = grid_roofline(local_rank, peaks)
    except Exception as e:  # never lose the headline line
        line["roofline_grid"] = {"error": str(e)}
    # ---- CPU baseline (oracle port) on a bounded sample, rank 0, N=1 only ----
    if world == 1 and not args.no_cpu:
        from mitsuba3_nasa_b200 import scene as S
        flat = S.load_dict(scenes.c2_rayleigh_slab(**C2_KW), 'llvm_ad_mono_polarized')
        cores = os.cpu_count() or 1
        n_cpu = 2_000_000
        v, dt, out = time_oracle(flat, n_cpu, 0, cores)
        line["cpu_baseline"] = {"value": v, "unit": "Msamples/s", "cores": cores, "kind": "port",
                                "sample": "%d of the 1e8 samples (%.1f s), std::thread over lane ranges" % (n_cpu, dt),
                                "stokes": [float(x) for x in out]}
    emit(line)
    if world > 1:
        dist.destroy_process_group()


def grid_roofline(device, peaks):
    """C3-shaped workload (256^3 grid, global majorant, Mie x lognormal tabulated phase) for the memory-side roofline:
    the path-pool kernel (k_forward_pool), mono volpath and -- as BASELINE configs[2] reads -- mono_polarized volpath."""
    import torch
    import mitsuba3_nasa_b200 as mi
    from mitsuba3_nasa_b200 import scenes
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    hbm = peaks.get("hbm_gbs", 6650.0)
    out = {}
    for variant, channels in (('llvm_ad_mono', 2), ('llvm_ad_mono_polarized', 4)):
        mi.set_variant(variant)
        d = scenes.c3_cloud(res=512, spp=16, grid_res=256)   # BASELINE C3 scene (Mie x lognormal table, 1801 nodes), 512^2 x 16 spp
        scene = mi.load_dict(d, device=device)
        ms, cnt = [], None
        for k in range(4):
            flush.zero_()                                      # evict the 64 MiB grid from the 126 MB L2
            torch.cuda.synchronize()
            scene.device.render(k, 16, develop=False, to_host=False)
            cnt = scene.device.counters()
            if k > 0:
                ms.append(cnt['kernel_ms'])
        t = float(np.mean(ms)) / 1e3
        b = algorithmic_bytes(cnt, channels)
        leg = {"bound": "hbm", "achieved": b / t / 1e9, "peak": hbm, "unit": "GB/s", "frac": b / t / 1e9 / hbm, "traffic": None,
               "peak_source": "measured (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback",
               "kernel": "k_forward_pool<%s, volpath>" % variant.replace('llvm_ad_', ''),
               "workload": "C3 scene (256^3 grid, Mie x lognormal table) at 512x512x16 spp, %s" % variant,
               "msamples_per_s": cnt['n_samples'] / t / 1e6, "ms": t * 1e3, "l2": "flushed between iterations (256 MiB write)",
               "events": cnt}
        try:
            fscene = mi.load_dict(scenes.c3_cloud(res=512, spp=16, grid_res=256, majorant_resolution=16), device=device)
            fms = []
            for k in range(3):
                flush.zero_()
                torch.cuda.synchronize()
                fscene.device.render(k, 16, develop=False, to_host=False)
                fc = fscene.device.counters()
                if k > 0:
                    fms.append(fc['kernel_ms'])
            leg["fast_mode"] = {"majorant_resolution": 16, "msamples_per_s": fc['n_samples'] / (float(np.mean(fms)) / 1e3) / 1e6,
                                "lookups_per_sample": fc['n_lookups'] / fc['n_samples'],
                                "note": "majorant supergrid + DDA, roulette in shadow walks: unbiased, but not the reference's random sequence"}
        except Exception as e:
            leg["fast_mode"] = {"error": str(e)}
        if variant == 'llvm_ad_mono':
            out = leg
        else:
            out["polarized"] = leg
    return out


def adjoint_bench(device, rank, world):
    """C4: path-replay adjoint of the C3 scene w.r.t. the 256^3 density grid (prb_volpathaos, mono_polarized,
    render_1_backward with d loss / d I = 1), 512x512x64 spp sharded over the ranks + gradient all-reduce."""
    import torch
    import mitsuba3_nasa_b200 as mi
    from mitsuba3_nasa_b200 import scenes
    from mitsuba3_nasa_b200.dist import ShardedScene
    mi.set_variant('llvm_ad_mono_polarized')
    scene = mi.load_dict(scenes.c4_cloud_adjoint(), device=device)
    sh = ShardedScene(scene)
    dev = scene.device
    spp = 64
    dev.grad_zero()
    sh.render_backward([1.0, 0.0, 0.0, 0.0], seed=99, spp=4, render_1=True)       # warm-up (also allocates the gradient)
    dev.grad_zero()
    torch.cuda.synchronize()
    if world > 1:
        torch.distributed.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    sh.render_backward([1.0, 0.0, 0.0, 0.0], seed=0, spp=spp, render_1=True)
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    c = dev.counters()
    t = torch.tensor([c['kernel_ms'], wall * 1e3], dtype=torch.float64, device="cuda")
    if world > 1:
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
    kernel_ms, wall_ms = [float(x) for x in t.cpu().numpy()]
    n = 512 * 512 * spp
    slot = sh.grad_slots()[0]
    g = dev.grad_get(slot, scene.flat.buffers[slot].size)
    # bytes: primal pass look-ups + backward look-ups (32 B each) + gradient read-modify-write through 8 taps (64 B)
    alg_bytes = 32 * c['n_lookups'] + 64 * (c['n_lookups'] / 2)
    fast = None
    try:
        # backend fast mode (majorant supergrid 16^3 + DDA, not the reference's random sequence): reported next to, never instead of, parity mode
        fscene = mi.load_dict(scenes.c4_cloud_adjoint(majorant_resolution=16), device=device)
        fsh = ShardedScene(fscene)
        fsh.render_backward([1.0, 0.0, 0.0, 0.0], seed=99, spp=4, render_1=True)
        fscene.device.grad_zero()
        fsh.render_backward([1.0, 0.0, 0.0, 0.0], seed=0, spp=spp, render_1=True)
        torch.cuda.synchronize()
        fc = fscene.device.counters()
        tf = torch.tensor([fc['kernel_ms']], dtype=torch.float64, device="cuda")
        if world > 1:
            torch.distributed.all_reduce(tf, op=torch.distributed.ReduceOp.MAX)
        fg = fscene.device.grad_get(slot, fscene.flat.buffers[slot].size)
        fast = {"majorant_resolution": 16, "msamples_per_s": n / (float(tf[0]) / 1e3) / 1e6, "kernel_ms_max_over_ranks": float(tf[0]),
                "lookups_per_sample": fc['n_lookups'] / max(fc['n_samples'], 1), "grad_abs_sum": float(np.abs(fg.astype(np.float64)).sum())}
    except Exception as e:
        fast = {"error": str(e)}
    return {"workload": "C4: C3 scene 512x512x64 spp, prb_volpathaos, render_1_backward wrt the 256^3 sigma_t grid", "kernel": "k_adjoint_pool<mono_polarized>",
            "fast_mode": fast, "lookups_per_sample": c['n_lookups'] / max(c['n_samples'], 1),
            "msamples_per_s": n / (kernel_ms / 1e3) / 1e6, "kernel_ms_max_over_ranks": kernel_ms,
            "wall_ms_with_gradient_allreduce": wall_ms, "gradient_bytes": int(g.size * 4),
            "grad_abs_sum": float(np.abs(g.astype(np.float64)).sum()), "events_this_rank": c,
            "algorithmic_gbs_this_rank": alg_bytes / (c['kernel_ms'] / 1e3) / 1e9,
            "note": "one sample = primal replay + backward pass of one lane (both passes timed)"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=3)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--no-cpu', action='store_true', help='skip the cpu_baseline leg')
    ap.add_argument('--no-adjoint', action='store_true', help='skip the C4 adjoint leg')
    args = ap.parse_args()
    _capture_stdout()
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_ours(args)


if __name__ == '__main__':
    main()
