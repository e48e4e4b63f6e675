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
METRIC = "Msamples/s polarized volpath fwd+adjoint at 1/2/4/8 GPU; % HBM/FP32 roofline"
WORKLOAD = ("C2 (BASELINE configs[1]): plane-parallel Rayleigh slab, mono_polarized, volpathaos, distant sensor, render_1 Stokes IQUV, "
            "1e8 samples per GPU; first-class `adjoint` key: C4 (configs[3]) prb_volpathaos w.r.t. the 256^3 density grid")
CONFIG = {"workload": WORKLOAD, "samples_per_gpu": SAMPLES_PER_GPU, "max_depth": 64, "rr_depth": 5, "sharding": "lane range per rank",
          "l2": "analytic scene: no input buffers to cache, L2 flush not applicable (grid legs: 256 MiB write between iterations)"}
# Algorithmic flops per event (FMA = 2), counted on the oracle source and recorded in BASELINE.md section 2a.  Homogeneous
# media carry no trilinear look-up; the look-up is charged per grid look-up (n_lookups) on heterogeneous scenes only.
FLOPS = dict(free_flight=36, lookup=88, real_pol=950, real_unpol=180, nee_setup_pol=111, nee_setup_unpol=90, nee_step=40,
             surface_pol=360, surface_unpol=70, sample_pol=195, sample_unpol=50)
FP32_PEAK_DERIVED_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12   # 74.4: 148 SMs x 128 lanes x FMA at clocks.max.sm (fallback only)


def algorithmic_flops(c, polarized=True):
    p = 'pol' if polarized else 'unpol'
    return (FLOPS['free_flight'] * (c['n_real'] + c['n_null'] + c['n_surface']) + FLOPS['lookup'] * c['n_lookups'] +
            (FLOPS['real_' + p] + FLOPS['nee_setup_' + p]) * c['n_real'] + FLOPS['nee_step'] * c['n_nee_steps'] +
            FLOPS['surface_' + p] * c['n_surface'] + FLOPS['sample_' + p] * c['n_samples'])


def algorithmic_bytes(c, film_channels=4):
    # 32 B per grid lookup (8 taps x 4 B) + 4*C B per sample of film traffic
    return 32 * c['n_lookups'] + 4 * film_channels * c['n_samples']


def lib_sha16():
    """Identity of the kernel build: sha256 over the sources libmi_b200.so is compiled from (csrc/*.cu, *.cuh, *.inc, the Makefile and
    the C header), in name order.  (The binary itself is no stable key: two nvcc builds of the same sources differ.)"""
    import glob
    import hashlib
    h = hashlib.sha256()
    src = os.path.join(ROOT, 'mitsuba3_nasa_b200', 'csrc')
    files = sorted(glob.glob(os.path.join(src, '*.cu')) + glob.glob(os.path.join(src, '*.cuh')) + glob.glob(os.path.join(src, '*.inc')))
    for f in files + [os.path.join(src, 'Makefile'), os.path.join(ROOT, 'include', 'mi_b200.h')]:
        h.update(os.path.basename(f).encode() + b'\0' + open(f, 'rb').read())
    return h.hexdigest()[:16]


def measured_traffic(kernel_key, n_samples):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch from the `ncu --set full` capture of THIS build
    (profiles/r02_traffic.json, written by profiles/ncu_traffic.py next to the capture): bytes per sample x this launch's
    samples.  None when the capture on file is of other kernel sources (lib_sha16 differs) or absent."""
    try:
        rec = json.load(open(os.path.join(ROOT, 'profiles', 'r02_traffic.json')))
        if rec.get('lib_sha16') != lib_sha16() or kernel_key not in rec['kernels']:
            return None
        k = rec['kernels'][kernel_key]
        return int(k['dram_bytes'] / k['n_samples'] * n_samples)
    except Exception:
        return None


def measured_issue(kernel_key):
    """What bounds these kernels is instruction issue, not a memory or FP32 roof: issue-slot utilisation, active lanes per
    instruction, L2 hit rate and resident warps per scheduler from the same `ncu --set full` capture of THIS build as
    measured_traffic (None when the capture on file is of other kernel sources)."""
    try:
        rec = json.load(open(os.path.join(ROOT, 'profiles', 'r02_traffic.json')))
        if rec.get('lib_sha16') != lib_sha16() or kernel_key not in rec['kernels']:
            return None
        out = dict(rec['kernels'][kernel_key].get('issue') or {})
        out['source'] = "ncu --set full capture of this build (%s; summaries profiles/r02_ncu_d_*.txt, profiles/r02_traffic.json)" % rec['kernels'][kernel_key].get('report', '')
        return out
    except Exception:
        return None


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


class ReferenceCPU:
    """The reference's OWN CPU implementation of the path: the unmodified volpathaos.cpp / render_1 of ksalesin/mitsuba3-nasa
    in its scalar_mono_polarized variant (oracle/_ref/scalar, built by oracle/build_ref_scalar.sh; the llvm_ad variants need
    libLLVM, which this image does not have).  Multi-threaded by the reference itself (integrator.cpp:510-551).  Falls back to
    the oracle port (kind "port") when the prebuilt reference is not on the box."""

    def __init__(self):
        import mitsuba3_nasa_b200 as mi
        from mitsuba3_nasa_b200 import scenes
        self.driver = os.path.join(ROOT, 'oracle', '_ref', 'scalar', 'ref_driver')
        self.cores = os.cpu_count() or 1
        self.kind = 'port'
        self.tmp = tempfile.mkdtemp(prefix='mib_ref_')
        d = scenes.c2_rayleigh_slab(**C2_KW)
        if os.path.exists(self.driver):
            try:
                self.xml = os.path.join(self.tmp, 'c2.xml')
                open(self.xml, 'w').write(mi.dict_to_xml(d, self.tmp))
                self._run(0, 1000)
                self.kind = 'reference'
            except Exception as e:
                print("bench.py: the prebuilt scalar reference did not run (%s; stderr: %s); using the oracle port"
                      % (e, (getattr(e, 'stderr', '') or '').strip()[-400:]), file=sys.stderr)
        if self.kind == 'port':
            from mitsuba3_nasa_b200 import scene as S
            self.flat = S.load_dict(d, 'llvm_ad_mono_polarized')

    def _run(self, seed, n):
        out = os.path.join(self.tmp, 'o.bin')
        # libmitsuba.so carries the RUNPATH of the build tree it came from; its own dependencies sit next to it
        env = dict(os.environ, LD_LIBRARY_PATH=os.path.dirname(self.driver) + os.pathsep + os.environ.get('LD_LIBRARY_PATH', ''))
        r = subprocess.run([self.driver, 'scalar_mono_polarized', self.xml, 'render1', str(seed), str(n), out, str(self.cores)],
                           stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, text=True, check=True, env=env)
        secs = [float(l.split()[1]) for l in r.stderr.splitlines() if l.startswith('RENDER1_SECONDS')]
        if not secs:
            raise RuntimeError("no timing line from ref_driver")
        stokes = np.fromfile(out, np.float32).reshape(4, 4)[:, 0]
        return secs[-1], stokes

    def time(self, seed, n):
        """-> (seconds of the render_1 call alone, Stokes vector)"""
        if self.kind == 'reference':
            return self._run(seed, n)
        v, dt, out = time_oracle(self.flat, n, seed, self.cores)
        return dt, np.asarray(out)

    def describe(self, n):
        if self.kind == 'reference':
            return ("%d samples per step of the 1e8-sample C2 workload: SamplingIntegrator::render_1 of the unmodified reference, "
                    "scalar_mono_polarized variant (oracle/_ref/scalar), %d host threads, time of the render_1 call" % (n, self.cores))
        return "%d samples per step of the 1e8-sample C2 workload: oracle port, std::thread over lane ranges" % n


def run_reference(args):
    """CPU arm: the reference's own CPU implementation of the path on the box's host cores (ReferenceCPU)."""
    rank, _, world = dist_env()
    if rank != 0:
        return
    ref = ReferenceCPU()
    n = 1_000_000        # bounded sample of the 1e8-sample workload per step (>= 1e6)
    for w in range(args.warmup):
        ref.time(1000 + w, n)
    t = []
    for k in range(args.steps):
        dt, _ = ref.time(k, n)
        t.append(dt)
    value = n * args.steps / sum(t) / 1e6
    line = {
        "impl": "reference", "metric": METRIC, "value": value,
        "unit": "Msamples/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * sum(t) / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": CONFIG,
        "cpu_baseline": {"value": value, "unit": "Msamples/s", "cores": ref.cores, "kind": ref.kind, "sample": ref.describe(n)},
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

    # ---- measured roofs of this device (rank 0): FP32 FMA microbenchmark, L2 streaming read ----
    from mitsuba3_nasa_b200 import _lib
    fp32_peak = l2_peak = None
    if rank == 0:
        try:
            fp32_peak, l2_peak = _lib.fp32_peak(local_rank, 5), _lib.l2_peak(local_rank, 3)
        except Exception as e:
            print("bench.py: peak microbenchmarks failed: %s" % e, file=sys.stderr)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
    except Exception:
        pass
    peaks["l2_gbs_measured_here"] = l2_peak

    # ---- adjoint (C4), all ranks: lanes sharded, gradients all-reduced over NCCL ----
    try:
        adjoint = adjoint_bench(local_rank, rank, world, peaks) if not args.no_adjoint else None
    except Exception as e:
        adjoint = {"error": str(e)}
    # ---- strong scaling legs (fixed total work shared by the ranks): C2 at 1e8 samples in total, C5 at 1e9 ----
    strong = {}
    try:
        if not args.no_strong:
            strong = strong_scaling(local_rank, rank, world, dev, barrier, allreduce)
    except Exception as e:
        strong = {"error": str(e)}
    if adjoint and "msamples_per_s" in adjoint:
        strong["c4_adjoint_512x512x64_msamples_per_s"] = adjoint["msamples_per_s"]
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    c1 = {kk: cnt_sum[kk] / args.steps for kk in cnt_sum}     # per launch (this rank)
    per_launch_s = kernel_ms / 1e3 / args.steps
    flops = algorithmic_flops(c1, polarized=True)
    achieved_tf = flops / per_launch_s / 1e12
    peak_tf = fp32_peak or FP32_PEAK_DERIVED_TFLOPS
    is_pool = ctypes.c_int(0)
    _lib_handle().mib_scene_forward_kernel(dev.h, ctypes.byref(is_pool))
    kernel_name = ("k_forward_pool" if is_pool.value else "k_forward") + "<mono_polarized, volpathaos>"
    line = {
        "metric": METRIC, "value": value, "unit": "Msamples/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_dev / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": CONFIG,
        "value_is": "forward leg of the metric (C2, configs[1]); the adjoint leg is the first-class `adjoint` key (C4, configs[3])",
        "stokes": [float(x) for x in stokes_all],
        "wall_s_timed_region": wall,
        "clocks": clk,
        "e2e": {"value": e2e, "unit": "Msamples/s",
                "h2d_bytes_per_step": ctypes.sizeof(_abi.SceneDesc) + int(_lib_handle().mib_sizeof_device_scene()),
                "d2h_bytes_per_step": 4 * 8 + ctypes.sizeof(_abi.Counters)},
        "gpu_launches": int(launches),
        "roofline": {"bound": "fp32", "achieved": achieved_tf, "peak": peak_tf, "unit": "TFLOP/s",
                     "frac": achieved_tf / peak_tf,
                     "traffic": measured_traffic("c2_forward", c1['n_samples']),
                     "issue": measured_issue("c2_forward"),
                     "traffic_note": "dram bytes of one launch from the ncu --set full capture of this very build (profiles/r02_traffic.json, "
                                     "keyed by the sha256 of the kernel sources), scaled to this launch's samples; null when no capture of this build is on file",
                     "peak_source": "measured in this run: independent FFMA chains on all SMs (mib_debug_fp32_peak), best of 5" if fp32_peak else
                                    "fallback: 148 SM x 128 lanes x 2 x 1.965 GHz (derived)",
                     "peak_derived_at_max_clock": FP32_PEAK_DERIVED_TFLOPS,
                     "kernel": kernel_name,
                     "note": "homogeneous medium: no grid traffic, FP32-pipe bound; flops = per-event constants counted on the oracle source "
                             "(BASELINE.md section 2a; no trilinear look-up on a homogeneous scene) x exact event counters",
                     "flops_per_event": FLOPS, "events_per_launch": c1, "hbm_peak_gbs_measured": peaks.get("hbm_gbs")},
        "lib_sha16": lib_sha16(),
    }
    line["adjoint"] = adjoint
    line["strong_scaling"] = strong
    # ---- secondary: heterogeneous grid workloads (memory side) on rank 0 ----
    try:
        line["roofline_grid"] = grid_roofline(local_rank, peaks)
    except Exception as e:  # never lose the headline line
        line["roofline_grid"] = {"error": str(e)}
    if world == 1 and not args.no_full:
        try:
            line["full_size_configs"] = full_size_configs(local_rank)
        except Exception as e:
            line["full_size_configs"] = {"error": str(e)}
    # ---- CPU baseline on a bounded sample, rank 0, N=1 only: the reference's own scalar build when it is on the box ----
    if world == 1 and not args.no_cpu:
        ref = ReferenceCPU()
        n_cpu = 4_000_000
        ref.time(999, 200_000)
        dt, out = ref.time(0, n_cpu)
        line["cpu_baseline"] = {"value": n_cpu / dt / 1e6, "unit": "Msamples/s", "cores": ref.cores, "kind": ref.kind,
                                "sample": ref.describe(n_cpu) + " (%.1f s)" % dt,
                                "stokes": [float(x) for x in np.asarray(out, dtype=np.float64)[:4]]}
    emit(line)
    if world > 1:
        dist.destroy_process_group()


def _timed_render(scene, seed, spp, flush=None):
    import torch
    if flush is not None:
        flush.zero_()                                      # evict the grid from the 126 MB L2
        torch.cuda.synchronize()
    scene.device.render(seed, spp, develop=False, to_host=False)
    return scene.device.counters()


def grid_roofline(device, peaks):
    """C3-shaped workload (256^3 grid, global majorant, Mie x lognormal tabulated phase) for the memory-side roofline:
    the path-pool kernel (k_forward_pool), mono volpath and -- as BASELINE configs[2] reads -- mono_polarized volpath.
    Three roofs per leg: HBM (algorithmic bytes), L2 (the sectors the gathers move: 8 taps x 32 B per look-up,
    against the L2 streaming rate measured in this run) and what ncu says bounds the kernel (issue slots; profiles/)."""
    import torch
    import mitsuba3_nasa_b200 as mi
    from mitsuba3_nasa_b200 import scenes
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    hbm = peaks.get("hbm_gbs", 6650.0)
    l2 = peaks.get("l2_gbs_measured_here")
    out = {}
    for variant, channels in (('llvm_ad_mono', 2), ('llvm_ad_mono_polarized', 4)):
        mi.set_variant(variant)
        d = scenes.c3_cloud(res=512, spp=16, grid_res=256)   # BASELINE C3 scene (Mie x lognormal table, 1801 nodes), 512^2 x 16 spp
        scene = mi.load_dict(d, device=device)
        ms, cnt = [], None
        for k in range(4):
            cnt = _timed_render(scene, k, 16, flush)
            if k > 0:
                ms.append(cnt['kernel_ms'])
        t = float(np.mean(ms)) / 1e3
        b = algorithmic_bytes(cnt, channels)
        sectors = 256 * cnt['n_lookups']
        leg = {"bound": "hbm", "achieved": b / t / 1e9, "peak": hbm, "unit": "GB/s", "frac": b / t / 1e9 / hbm,
               "traffic": measured_traffic("c3_forward_" + variant.replace('llvm_ad_', ''), cnt['n_samples']),
               "issue": measured_issue("c3_forward_" + variant.replace('llvm_ad_', '')),
               "peak_source": "measured (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback",
               "l2_roof": {"achieved": sectors / t / 1e9, "peak": l2, "unit": "GB/s", "frac": (sectors / t / 1e9 / l2) if l2 else None,
                           "note": "8 taps x 32-byte sectors per look-up (the gathers are L2 hits, ncu lts__t_sector_hit_rate ~90 %) against the L2 "
                                   "streaming-read rate measured in this run (mib_debug_l2_peak)"},
               "limiter": "instruction issue, not memory: see profiles/r02_ncu_*_c3pol_kforward_pool.txt (issue slots busy, stall reasons)",
               "kernel": "k_forward_pool<%s, volpath>" % variant.replace('llvm_ad_', ''),
               "workload": "C3 scene (256^3 grid, Mie x lognormal table) at 512x512x16 spp, %s" % variant,
               "msamples_per_s": cnt['n_samples'] / t / 1e6, "ms": t * 1e3, "l2": "flushed between iterations (256 MiB write)",
               "events": cnt}
        try:
            fscene = mi.load_dict(scenes.c3_cloud(res=512, spp=16, grid_res=256, majorant_resolution=16), device=device)
            fms = []
            for k in range(3):
                fc = _timed_render(fscene, k, 16, flush)
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


def full_size_configs(device):
    """BASELINE configs at their configured size, one timed launch each after a small warm-up launch (N = 1):
    C1 256 x 256 x 64 spp rgb; C3 1024 x 1024 x 256 spp mono and mono_polarized."""
    import mitsuba3_nasa_b200 as mi
    from mitsuba3_nasa_b200 import scenes
    out = {}
    mi.set_variant('llvm_ad_rgb')
    s = mi.load_dict(scenes.c1_homogeneous_cube(), device=device)
    _timed_render(s, 7, 64)
    c = _timed_render(s, 0, 64)
    out["c1_rgb_256x256x64"] = {"msamples_per_s": c['n_samples'] / c['kernel_ms'] / 1e3, "kernel_ms": c['kernel_ms'], "n_samples": c['n_samples']}
    for variant in ('llvm_ad_mono', 'llvm_ad_mono_polarized'):
        mi.set_variant(variant)
        s = mi.load_dict(scenes.c3_cloud(), device=device)          # 1024 x 1024, 256 spp, 256^3 grid, Mie x lognormal table
        _timed_render(s, 7, 1)
        c = _timed_render(s, 0, 256)
        out["c3_%s_1024x1024x256" % variant.replace('llvm_ad_', '')] = {
            "msamples_per_s": c['n_samples'] / c['kernel_ms'] / 1e3, "kernel_ms": c['kernel_ms'], "n_samples": c['n_samples'],
            "lookups_per_sample": c['n_lookups'] / c['n_samples'],
            "hbm_frac": algorithmic_bytes(c, 2 if variant.endswith('mono') else 4) / (c['kernel_ms'] / 1e3) / 1e9 / 6464.6}
    return out


def strong_scaling(device, rank, world, c2_dev, barrier, allreduce):
    """Fixed total work shared by the ranks, so that value(N) / value(1) is a speed-up:
    C2 with 1e8 samples IN TOTAL (1e8 / N per rank) and C5 (BASELINE configs[4]) with 1e9 samples on the polarized C3 scene."""
    import mitsuba3_nasa_b200 as mi
    from mitsuba3_nasa_b200 import scenes
    from mitsuba3_nasa_b200.dist import ShardedScene
    out = {}
    share = SAMPLES_PER_GPU // world
    c2_dev.set_lane_range(rank * share, (rank + 1) * share)
    c2_dev.render_1(77, SAMPLES_PER_GPU, normalize=False)
    barrier()
    c2_dev.render_1(0, SAMPLES_PER_GPU, normalize=False)
    ms = float(allreduce([c2_dev.counters()['kernel_ms']], "max")[0])
    out["c2_1e8_total_msamples_per_s"] = share * world / ms / 1e3
    mi.set_variant('llvm_ad_mono_polarized')
    spp = 954                                                # 1024 x 1024 x 954 = 1.0003e9 samples
    scene = mi.load_dict(scenes.c3_cloud(spp=spp), device=device)
    sh = ShardedScene(scene)
    sh.render(seed=7, spp=1)
    barrier()
    sh.render(seed=0, spp=spp)
    ms = float(allreduce([scene.device.counters()['kernel_ms']], "max")[0])
    out["c5_1e9_polarized_msamples_per_s"] = 1024 * 1024 * spp / ms / 1e3
    out["note"] = "kernel time, max over ranks; the film all-reduce of C5 (16 MiB) is outside this figure"
    return out


def adjoint_bench(device, rank, world, peaks):
    """C4 (BASELINE configs[3]): path-replay adjoint of the C3 scene w.r.t. the 256^3 density grid (prb_volpathaos,
    mono_polarized, render_1_backward with d loss / d I = 1), 512x512x64 spp sharded over the ranks + gradient all-reduce.
    One sample = primal replay + backward pass of one lane (both passes timed).  `e2e`: the same through the public API with
    HOST buffers -- the 64 MiB grid goes host -> device, mi.render_1 runs the primal pass, backward the adjoint kernel, and
    the 64 MiB gradient comes back to the host -- every step."""
    import torch
    import mitsuba3_nasa_b200 as mi
    from mitsuba3_nasa_b200 import scenes
    from mitsuba3_nasa_b200.dist import ShardedScene
    mi.set_variant('llvm_ad_mono_polarized')
    d = scenes.c4_cloud_adjoint()
    scene = mi.load_dict(d, device=device)
    sh = ShardedScene(scene)
    dev = scene.device
    spp = 64
    dev.grad_zero()
    sh.render_backward([1.0, 0.0, 0.0, 0.0], seed=99, spp=4, render_1=True)       # warm-up (also allocates the gradient)
    dev.grad_zero()
    torch.cuda.synchronize()
    if world > 1:
        torch.distributed.barrier()
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
    hbm = peaks.get("hbm_gbs", 6650.0)
    gbs = alg_bytes / (c['kernel_ms'] / 1e3) / 1e9
    out = {"workload": "C4: C3 scene 512x512x64 spp, prb_volpathaos, render_1_backward wrt the 256^3 sigma_t grid", "kernel": "k_adjoint_pool<mono_polarized>",
           "value": n / (kernel_ms / 1e3) / 1e6, "unit": "Msamples/s",
           "lookups_per_sample": c['n_lookups'] / max(c['n_samples'], 1),
           "msamples_per_s": n / (kernel_ms / 1e3) / 1e6, "kernel_ms_max_over_ranks": kernel_ms,
           "wall_ms_with_gradient_allreduce": wall_ms, "gradient_bytes": int(g.size * 4),
           "grad_abs_sum": float(np.abs(g.astype(np.float64)).sum()), "events_this_rank": c,
           "roofline": {"bound": "hbm", "achieved": gbs, "peak": hbm, "unit": "GB/s", "frac": gbs / hbm,
                        "traffic": measured_traffic("c4_adjoint", c['n_samples']),
                        "issue": measured_issue("c4_adjoint"),
                        "note": "32 B per look-up of both passes + 64 B gradient read-modify-write per backward look-up"},
           "note": "one sample = primal replay + backward pass of one lane (both passes timed)"}
    # ---- end to end with host buffers, this rank's share ----
    try:
        key = 'object.interior_medium.sigma_t.data'
        params = mi.traverse(scene)
        grid_host = np.array(params[key], dtype=np.float32)
        steps = 2
        sh._set_slice(spp)
        torch.cuda.synchronize()
        if world > 1:
            torch.distributed.barrier()
        t0 = time.perf_counter()
        for k in range(steps):
            x = torch.from_numpy(grid_host).requires_grad_(True)
            params[key] = x                                              # host array -> H2D at update() inside mi.render_1
            m = mi.render_1(scene, params, seed=k, spp=spp)              # primal pass (prb_volpathaos), result on the host
            m[0, 0].backward() if m.ndim == 2 else m[0].backward()       # adjoint kernel; gradient D2H into x.grad
            assert x.grad is not None
        torch.cuda.synchronize()
        te = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        if world > 1:
            torch.distributed.all_reduce(te, op=torch.distributed.ReduceOp.MAX)
        out["e2e"] = {"value": n * steps / float(te[0]) / 1e6, "unit": "Msamples/s", "steps": steps,
                      "h2d_bytes_per_step": int(grid_host.nbytes), "d2h_bytes_per_step": int(grid_host.nbytes) + 16 * 8,
                      "note": "mi.render_1(scene, params) + backward() through torch autograd (ad.py): grid H2D, primal kernel, adjoint kernel, "
                              "gradient D2H; includes the primal render the device-timed figure does not"}
        dev.set_lane_range(0, 0)
    except Exception as e:
        out["e2e"] = {"error": str(e)}
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
        out["fast_mode"] = {"majorant_resolution": 16, "msamples_per_s": n / (float(tf[0]) / 1e3) / 1e6, "kernel_ms_max_over_ranks": float(tf[0]),
                            "lookups_per_sample": fc['n_lookups'] / max(fc['n_samples'], 1), "grad_abs_sum": float(np.abs(fg.astype(np.float64)).sum())}
    except Exception as e:
        out["fast_mode"] = {"error": str(e)}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=3)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--no-cpu', action='store_true', help='skip the cpu_baseline leg')
    ap.add_argument('--no-adjoint', action='store_true', help='skip the C4 adjoint leg')
    ap.add_argument('--no-strong', action='store_true', help='skip the strong-scaling legs (C2 1e8 total, C5 1e9)')
    ap.add_argument('--no-full', action='store_true', help='skip the full-size C1 / C3 launches')
    args = ap.parse_args()
    _capture_stdout()
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_ours(args)


if __name__ == '__main__':
    main()
