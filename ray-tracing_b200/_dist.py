"""Multi-GPU host logic: one process per GPU (torch.distributed), the scene replicated on every rank, camera samples sharded by
contiguous samples-per-pixel slices (optionally also by row bands), partial framebuffers summed onto rank 0 with one reduce
(NCCL over NVLink on GPUs; gloo in the CPU tests).  Sample s of pixel p always draws from streamGen(seed, p, s), so the
image does not depend on how the work was sharded (up to the order of the double-precision sums).
"""
import torch
import torch.distributed as dist


def shard_samples(total_spp, rank, world):
    """Contiguous slice [s0, s0+n) of the frame's samples for `rank`; the remainder goes to the lowest ranks."""
    base, rem = divmod(int(total_spp), int(world))
    n = base + (1 if rank < rem else 0)
    s0 = rank * base + min(rank, rem)
    return s0, n


def shard_rows(height, rank, world):
    """Row band [r0, r1) for `rank` (tile sharding for very low spp)."""
    base, rem = divmod(int(height), int(world))
    r0 = rank * base + min(rank, rem)
    return r0, r0 + base + (1 if rank < rem else 0)


def plan(total_spp, height, rank, world):
    """spp sharding when every rank gets at least one sample, else row bands with all samples each."""
    if total_spp >= world:
        s0, n = shard_samples(total_spp, rank, world)
        return dict(s0=s0, spp=n, region=None)
    r0, r1 = shard_rows(height, rank, world)
    return dict(s0=0, spp=total_spp, region=(r0, r1, 0, 0) if r1 > r0 else None, empty=(r1 <= r0))


def reduce_framebuffer(partial, dst=0, group=None):
    """Sum of the ranks' partial radiance sums (float64 [H,W,3]) onto rank `dst`, in place."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(partial, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return partial


def render_sharded(scene, W, H, total_spp, depth, seed, device_buffer, stream=0, precision=0, flags=0, group=None):
    """This rank's share of a frame rendered into `device_buffer` (zeroed here), then reduced onto rank 0.
    `scene` is a DeviceScene on this rank's GPU; `device_buffer` a CUDA float64 tensor [H,W,3] on the same GPU."""
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    pl = plan(total_spp, H, rank, world)
    device_buffer.zero_()
    if pl["spp"] > 0 and not pl.get("empty", False):
        region = pl["region"]
        if region is not None:
            region = (region[0], region[1], 0, W)
        scene.render_device(device_buffer.data_ptr(), W, H, pl["spp"], stream=stream, depth=depth, seed=seed, s0=pl["s0"],
                            region=region, precision=precision, flags=flags, total=total_spp)
    return reduce_framebuffer(device_buffer, 0, group)
