"""Multi-GPU path on real devices (needs >= 2 GPUs; skipped on a single-GPU box): ShardedScene over NCCL must
reproduce the single-GPU film, render_1 sums and gradients up to fp32 summation order."""
import os
import socket
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpu():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


def _scenes():
    from mitsuba3_nasa_b200 import scenes
    grid = np.random.default_rng(4).uniform(0.05, 1.0, (10, 10, 10, 1)).astype(np.float32)
    d = scenes.c3_cloud(res=12, spp=16, max_depth=12, integrator='prb_volpathaos', grid=grid, optical_depth=5.0,
                        phase={'type': 'rayleigh'})
    d['emitter'] = {'type': 'directional', 'direction': [0.3, -0.2, -1.0], 'irradiance': 1.0}
    return scenes.c1_homogeneous_cube(res=12, spp=8, max_depth=8, rfilter='gaussian'), d


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    import mitsuba3_nasa_b200 as mi
    from mitsuba3_nasa_b200.dist import ShardedScene
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        c1, c4 = _scenes()
        mi.set_variant('llvm_ad_rgb')
        img = ShardedScene(mi.load_dict(c1, device=rank)).render(seed=3, spp=8)
        mi.set_variant('llvm_ad_mono_polarized')
        sc = mi.load_dict(c4, device=rank)
        sh = ShardedScene(sc)
        stokes = sh.render_1(seed=1, spp=16)
        for _ in range(2):                       # twice: the second call must not see leftovers of the first
            sc.device.grad_zero()
            sh.render_backward([1.0, 0.5, 0.0, 0.0], seed=2, spp=16, render_1=True)
        slot = sh.grad_slots()[0]
        g = sc.device.grad_get(slot, sc.flat.buffers[slot].size)
        gs = sc.device.grad_get_scalar(0)
        np.savez(os.path.join(out_dir, "rank%d.npz" % rank), img=img, stokes=stokes, g=g, gs=gs)
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(_ngpu() < 2, reason="needs two GPUs")
def test_sharded_scene_matches_single_gpu(tmp_path):
    import torch.multiprocessing as mp
    import mitsuba3_nasa_b200 as mi
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    r0, r1 = [np.load(os.path.join(str(tmp_path), "rank%d.npz" % r)) for r in (0, 1)]
    for k in ("img", "stokes", "g", "gs"):
        assert np.array_equal(r0[k], r1[k]), k + " differs between ranks after the all-reduce"
    c1, c4 = _scenes()
    mi.set_variant('llvm_ad_rgb')
    img = mi.load_dict(c1).device.render(3, 8)
    assert np.allclose(r0["img"], img, rtol=2e-5, atol=1e-6)
    mi.set_variant('llvm_ad_mono_polarized')
    sc = mi.load_dict(c4)
    assert np.allclose(r0["stokes"], sc.device.render_1(1, 16), rtol=1e-9, atol=1e-12)
    sc.device.grad_zero()
    sc.device.render_1_backward(2, 16, [1.0, 0.5, 0.0, 0.0])
    slot = sc.flat.desc.media[0].sigma_t.buffer
    g = sc.device.grad_get(slot, sc.flat.buffers[slot].size)
    assert np.abs(r0["g"] - g).max() <= 1e-5 * np.abs(g).max()
    assert np.allclose(r0["gs"], sc.device.grad_get_scalar(0), rtol=1e-9, atol=1e-12)
