"""Multi-GPU sharding of a render: one process per GPU, scene + BVH replicated, disjoint ITERATION ranges.

The reference's per-path random stream is a function of (x, y, iteration, W, H) only
(Kernel/PathTracer_FullKernel_header.cl:255-264), so any partition of the iteration range traces exactly the paths a
single-device run traces. The only exchange step is one sum of the accumulators (imageColor float4, imageRayNb float,
the three uint32 histograms) to rank 0: `torch.distributed` reduce over NCCL (NVLink) on GPUs, gloo in the CPU tests.
imageV (variance accumulator, read only by the reference's dead supersampling path) is not linearly summable and is
not reduced.

Progressive rendering (1 spp per displayed frame) cannot be split by iterations. There each rank renders interleaved
stripes of image rows (`PathTracerCUDA.set_pixel_stripes`; the random streams are addressed by pixel, so the shards
trace the paths of the whole-frame render) and the display frames, which are 0 outside a rank's stripes, are summed to
rank 0 as uint8 (`compose_stripes`): 6.2 MB per 1080p frame instead of 41.5 MB of float accumulators.
"""

STRIPE_ROWS = 8


def stripe_rows(rank, world, height, rows_per_stripe=STRIPE_ROWS):
    """Image rows of `rank`'s shard: y with (y // rows_per_stripe) % world == rank (what ptc_set_pixel_stripes selects)."""
    if world <= 0 or not (0 <= rank < world) or rows_per_stripe <= 0:
        raise ValueError("bad rank/world/stripe %r/%r/%r" % (rank, world, rows_per_stripe))
    return [y for y in range(height) if (y // rows_per_stripe) % world == rank]


def compose_stripes(frame, dst=0, group=None):
    """Sum the uint8 display frames of all ranks to rank `dst` (every pixel is non-zero on at most one rank)."""
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return
    dist.reduce(frame, dst=dst, op=dist.ReduceOp.SUM, group=group)


def display_tensor(renderer, device):
    """torch uint8 tensor aliasing the renderer's device display frame (H*W*3), filled by renderer.convert_rgb8()."""
    import torch
    b = renderer.device_buffers()
    return torch.as_tensor(DeviceArray(b.rgb8, renderer.n_pixels * 3, "|u1"), device=device)


def iteration_range(rank, world, first_iteration, n_iterations):
    """Contiguous share of [first, first + n) for `rank`: sizes differ by at most one, earlier ranks get the extras."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank/world %r/%r" % (rank, world))
    base, extra = divmod(n_iterations, world)
    begin = first_iteration + rank * base + min(rank, extra)
    return begin, base + (1 if rank < extra else 0)


def reduce_accumulators(tensors, dst=0, group=None):
    """Sum every tensor of `tensors` (an iterable of torch tensors aliasing the accumulators) to rank `dst`.
    uint32 histograms must be passed as int32 / int64 views (NCCL has no uint32 sum in torch)."""
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return
    for t in tensors:
        dist.reduce(t, dst=dst, op=dist.ReduceOp.SUM, group=group)


class DeviceArray:
    """__cuda_array_interface__ view of a device buffer owned by a PathTracer_CUDA context (zero copy)."""

    def __init__(self, ptr, count, typestr="<f4"):
        self.__cuda_array_interface__ = {"shape": (int(count),), "typestr": typestr, "data": (int(ptr), False), "version": 2}


def accumulator_tensors(renderer, device):
    """torch tensors aliasing the renderer's device accumulators: (color[nPix*4] f32, count[nPix] f32,
    depths[int32], bbx[int32 5000], tri[int32 5000])."""
    import torch
    b = renderer.device_buffers()
    n = renderer.n_pixels
    mk = lambda p, c, t: torch.as_tensor(DeviceArray(p, c, t), device=device)
    return (mk(b.image_color, n * 4, "<f4"), mk(b.image_ray_nb, n, "<f4"), mk(b.ray_depths, renderer.ray_max_depth + 1, "<i4"),
            mk(b.ray_bbx, 5000, "<i4"), mk(b.ray_tri, 5000, "<i4"))
