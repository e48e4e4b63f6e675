"""Host-side multi-GPU logic on CPU: lane sharding + the reductions of mitsuba3_nasa_b200/dist.py,
run with world_size 2 over gloo.  The per-rank renderer here is the oracle (test infrastructure);
what is under test is the sharding arithmetic and the reduce-then-develop order."""
import os
import socket
import sys

import numpy as np
import pytest

from mitsuba3_nasa_b200.dist import lane_slice

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("n_pixels,spp,world", [(1, 100000000, 8), (1, 7, 2), (256 * 256, 64, 8), (10, 3, 4), (3, 5, 8), (1, 1, 2)])
def test_lane_slices_tile_the_wavefront(n_pixels, spp, world):
    total = n_pixels * spp
    prev = 0
    for r in range(world):
        b, e = lane_slice(n_pixels, spp, r, world)
        assert b == prev and e >= b
        if n_pixels >= world:
            assert b % spp == 0 and e % spp == 0          # whole pixels per rank
        prev = e
    assert prev == total
    sizes = [lane_slice(n_pixels, spp, r, world)[1] - lane_slice(n_pixels, spp, r, world)[0] for r in range(world)]
    unit = spp if n_pixels >= world else 1
    assert max(sizes) - min(sizes) <= unit                # balanced


def test_lane_slice_rejects_bad_rank():
    with pytest.raises(ValueError):
        lane_slice(4, 4, 2, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import torch.distributed as dist
    from mitsuba3_nasa_b200 import scene as S, scenes
    from mitsuba3_nasa_b200.dist import lane_slice as ls
    from orc import OracleScene
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # forward film: 12x12 pixels, 8 spp, sharded by pixel
        sc = S.load_dict(scenes.c1_homogeneous_cube(res=12, spp=8, max_depth=8), "llvm_ad_rgb")
        o = OracleScene(sc)
        b, e = ls(12 * 12, 8, rank, world)
        raw, _ = o.render(3, 8, lanes=(b, e) if e > b else (12 * 12 * 8, 12 * 12 * 8), nthreads=2)
        t = torch.from_numpy(raw.astype(np.float32))
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        # render_1 on a 1x1 film sharded along the sample index
        sc2 = S.load_dict(scenes.c2_rayleigh_slab(spp=4), "llvm_ad_mono_polarized")
        o2 = OracleScene(sc2)
        n = 4001
        b2, e2 = ls(1, n, rank, world)
        sums = o2.render_1(0, n, lanes=(b2, e2), nthreads=2, normalize=False)
        t2 = torch.from_numpy(sums.copy())
        dist.all_reduce(t2, op=dist.ReduceOp.SUM)
        # adjoint: gradients of the lane slices add up
        sc3 = S.load_dict(scenes.c1_homogeneous_cube(res=6, spp=8, max_depth=8) | {
            'integrator': {'type': 'prbvolpath', 'max_depth': 8, 'rr_depth': 5}}, "llvm_ad_rgb")
        o3 = OracleScene(sc3)
        b3, e3 = ls(36, 8, rank, world)
        gi = np.random.default_rng(1).uniform(-1, 1, (6, 6, 3)).astype(np.float32)
        o3.render_backward(2, 8, gi, lanes=(b3, e3), nthreads=2)
        t3 = torch.from_numpy(o3.grad_get_scalar(0).copy())
        dist.all_reduce(t3, op=dist.ReduceOp.SUM)
        if rank == 0:
            np.savez(os.path.join(out_dir, "reduced.npz"), film=t.numpy(), sums=t2.numpy(), grads=t3.numpy())
    finally:
        dist.destroy_process_group()


def test_world_size_2_gloo_matches_single_process(tmp_path):
    import torch.multiprocessing as mp
    from mitsuba3_nasa_b200 import scene as S, scenes
    from orc import OracleScene
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    got = np.load(os.path.join(str(tmp_path), "reduced.npz"))

    sc = S.load_dict(scenes.c1_homogeneous_cube(res=12, spp=8, max_depth=8), "llvm_ad_rgb")
    raw, _ = OracleScene(sc).render(3, 8, nthreads=2)
    assert np.allclose(got["film"], raw, rtol=1e-6, atol=1e-7)
    assert np.array_equal(got["film"][..., 3], raw[..., 3])            # sample weights: exact counts

    sc2 = S.load_dict(scenes.c2_rayleigh_slab(spp=4), "llvm_ad_mono_polarized")
    sums = OracleScene(sc2).render_1(0, 4001, nthreads=2, normalize=False)
    assert np.allclose(got["sums"], sums, rtol=1e-12, atol=0)

    sc3 = S.load_dict(scenes.c1_homogeneous_cube(res=6, spp=8, max_depth=8) | {
        'integrator': {'type': 'prbvolpath', 'max_depth': 8, 'rr_depth': 5}}, "llvm_ad_rgb")
    o3 = OracleScene(sc3)
    gi = np.random.default_rng(1).uniform(-1, 1, (6, 6, 3)).astype(np.float32)
    o3.render_backward(2, 8, gi, nthreads=2)
    assert np.allclose(got["grads"], o3.grad_get_scalar(0), rtol=1e-10, atol=1e-14)


def test_stripe_lanes_partition_the_wavefront():
    """Load-balanced sharding: interleaved stripes of 64 pixels tile the lane range exactly once (host mirror of
    mib_scene_set_lane_stripes), for sizes that are not multiples of the chunk."""
    from mitsuba3_nasa_b200.dist import stripe_lanes
    for n_pix, spp, world in ((1000, 3, 4), (64 * 7 + 5, 8, 2), (512, 1, 8), (10, 4, 3)):
        parts = [stripe_lanes(n_pix, spp, r, world) for r in range(world)]
        allv = np.sort(np.concatenate(parts))
        assert np.array_equal(allv, np.arange(n_pix * spp))
        # whole pixels per rank, and every rank's lanes are spread over the whole range
        for p in parts:
            if p.size:
                assert (p // spp).size == p.size and np.all(np.bincount(p // spp)[np.unique(p // spp)] == spp)
