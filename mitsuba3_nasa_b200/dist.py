"""Multi-GPU driver: one process per GPU, lanes sharded across ranks.

The walk has no exchange step: a lane (pixel, sample) is independent of every other lane
and its PCG32 seed depends only on the global lane index, so rank r simply renders the
lane slice [begin_r, end_r) of the same wavefront against a full replica of the scene.
The only collectives are the reductions the reference does implicitly when all lanes
live in one array:
  * forward image : sum of the raw accumulation film (before the weight division of
    HDRFilm::develop, hdrfilm.cpp:304-418) -> one all_reduce of H*W*C floats;
  * render_1      : sum of 4 doubles (integrator.cpp:665-705);
  * adjoint       : sum of each parameter gradient (64 MiB for a 256^3 grid).
They run through torch.distributed (NCCL on GPUs, gloo in the CPU tests) directly on the
device buffers owned by libmi_b200.so (wrapped zero-copy through __cuda_array_interface__).
"""
import numpy as np


def lane_slice(n_pixels, spp, rank, world):
    """Contiguous slice of the global lane range [0, n_pixels*spp) owned by `rank`.

    Slices are aligned to whole pixels so that every pixel is accumulated by exactly one
    rank; a film with fewer pixels than ranks (the 1x1 film of render_1 runs) is sliced
    along the sample index instead.  Returns (begin, end); the slices tile the range.
    """
    if world < 1 or not (0 <= rank < world):
        raise ValueError("lane_slice: bad rank %r / world %r" % (rank, world))
    total = int(n_pixels) * int(spp)
    if n_pixels >= world:
        per, rem = divmod(int(n_pixels), world)
        b = rank * per + min(rank, rem)
        e = b + per + (1 if rank < rem else 0)
        return b * spp, e * spp
    per, rem = divmod(total, world)
    b = rank * per + min(rank, rem)
    return b, b + per + (1 if rank < rem else 0)


# pixels per stripe chunk.  Measured on the C4 adjoint by running every rank's share on one GPU (profiles/dist_balance.py,
# profiles/r02_dist_balance.log): slowest / mean rank time 1.17 (N=2), 1.15 (N=4) with 64-pixel chunks -- with the rotation below a
# 512-pixel row still gives a rank the same column bands in every row when N < 8 -- against 1.03 / 1.01 with 4-pixel chunks.
STRIPE_PIXELS = 4


def stripe_lanes(n_pixels, spp, rank, world, chunk_pixels=STRIPE_PIXELS):
    """Lanes of `rank` under striping (host mirror of mib_scene_set_lane_stripes, for tests and documentation): chunk c of
    chunk_pixels * spp lanes belongs to rank (c + c // world) % world.  Every rank samples the whole image, so the slowest rank is not the
    one that happens to own the rows through the cloud (contiguous slices: 37 % strong-scaling efficiency on C5 at 8 GPUs)."""
    total, chunk = int(n_pixels) * int(spp), int(chunk_pixels) * int(spp)
    lanes = np.arange(total, dtype=np.int64)
    c = lanes // chunk
    g = c // world                                    # chunks are dealt in groups of `world`, rotated by the group index so
    return lanes[(c - g * world + g) % world == rank]  # that a rank does not always get the same column band of the image


class _DevArray:
    """Zero-copy view of a device buffer owned by the CUDA library."""

    def __init__(self, ptr, count, typestr):
        self.__cuda_array_interface__ = {"shape": (int(count),), "typestr": typestr, "data": (int(ptr), False),
                                         "version": 3, "strides": None}


def _as_tensor(ptr, count, dtype, device):
    import torch
    typestr = {"float32": "<f4", "float64": "<f8"}[dtype]
    return torch.as_tensor(_DevArray(ptr, count, typestr), device=torch.device("cuda", device))


def _sync(device):
    """The collectives run on torch's NCCL stream, the kernels on the library's own stream: the buffers may only be
    touched again (grad_zero, develop, the next render) once the collective has finished."""
    import torch
    torch.cuda.synchronize(device)


def _world():
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


class ShardedScene:
    """Wraps an api.Scene: same render / render_1 / render_backward calls, sharded over the process group."""

    def __init__(self, scene):
        self.scene = scene
        self.rank, self.world = _world()

    def _set_slice(self, spp):
        sn = self.scene.flat.desc.sensor
        n_pix = sn.crop_width * sn.crop_height
        if self.world > 1 and n_pix >= 4 * STRIPE_PIXELS * self.world:
            # films: interleaved stripes of STRIPE_PIXELS pixels, every rank samples the whole image (load balance)
            self.scene.device.set_lane_stripes(self.rank, self.world, STRIPE_PIXELS * spp)
            return
        self.scene.device.set_lane_stripes(0, 1, 1)
        b, e = lane_slice(n_pix, spp, self.rank, self.world)
        if b == e:                        # empty slice: (total, total); (0, 0) would mean "everything"
            b = e = n_pix * spp
        self.scene.device.set_lane_range(b, e)

    def grad_slots(self):
        """Buffer slots that carry a gradient: grids a medium differentiates through."""
        d = self.scene.flat.desc
        slots = []
        for m in range(d.n_media):
            for v in (d.media[m].sigma_t, d.media[m].albedo):
                if v.type == 1 and v.buffer not in slots:      # MIB_VOLUME_GRID
                    slots.append(int(v.buffer))
        return slots

    def render(self, seed=0, spp=0):
        """Image identical (up to fp32 summation order) to the single-GPU render; valid on every rank."""
        import torch.distributed as dist
        spp = self.scene._spp(spp)
        self._set_slice(spp)
        dev = self.scene.device
        dev.render(int(seed), spp, develop=False, to_host=False)
        if self.world > 1:
            (raw_ptr, raw_n), _ = dev.film_dev()
            dist.all_reduce(_as_tensor(raw_ptr, raw_n, "float32", dev.device_index), op=dist.ReduceOp.SUM)
            _sync(dev.device_index)
        dev.film_develop()
        return dev.film_read(developed=True)

    def render_1(self, seed=0, spp=0):
        import torch
        import torch.distributed as dist
        spp = self.scene._spp(spp)
        self._set_slice(spp)
        dev = self.scene.device
        sums = np.asarray(dev.render_1(int(seed), spp, normalize=False), np.float64)
        if self.world > 1:
            t = torch.from_numpy(sums.copy()).to(torch.device("cuda", dev.device_index))
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
            sums = t.cpu().numpy()
        sn = self.scene.flat.desc.sensor
        # integrator.cpp:702-705: the normalisation factor is a float32 reciprocal (same as mib_render_1)
        nf = np.float32(1.0) / (np.float32(sn.crop_width * sn.crop_height) * np.float32(spp))
        return sums * float(nf)

    def render_backward(self, grad_in, seed=0, spp=0, render_1=False):
        """Accumulates the gradients of this rank's lanes, then sums every gradient buffer over the ranks.
        Call grad_zero() on every rank before, as for the single-GPU API."""
        import torch.distributed as dist
        spp = self.scene._spp(spp)
        self._set_slice(spp)
        dev = self.scene.device
        if render_1:
            dev.render_1_backward(int(seed), spp, grad_in)
        else:
            dev.render_backward(int(seed), spp, grad_in)
        if self.world > 1:
            for slot in self.grad_slots():
                ptr, n = dev.grad_dev(slot)
                dist.all_reduce(_as_tensor(ptr, n, "float32", dev.device_index), op=dist.ReduceOp.SUM)
            ptr, n = dev.grad_scalar_dev()
            dist.all_reduce(_as_tensor(ptr, n, "float64", dev.device_index), op=dist.ReduceOp.SUM)
            _sync(dev.device_index)
