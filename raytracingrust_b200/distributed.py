"""Multi-GPU render: sample slices per rank + ONE reduce of the fp32 accumulation buffers.

Every pixel-sample is independent (lib/core/render.rs:86-100) and the per-sample clamp happens
before accumulation (render.rs:95-99), so per-rank partial sums add linearly.  Philox is keyed by
the GLOBAL (pixel, sample) index, hence the image does not depend on the number of ranks except
for fp32 summation order.  One process per GPU; torch.distributed supplies the collective (NCCL
over NVLink on GPUs, gloo in the CPU tests).
"""
import torch
import torch.distributed as dist


def sample_slice(spp: int, rank: int, world: int):
    """Contiguous, balanced partition of global samples [0, spp): returns (begin, count)."""
    base, rem = divmod(spp, world)
    begin = rank * base + min(rank, rem)
    return begin, base + (1 if rank < rem else 0)


def render_distributed(accumulate, resolve, width, height, spp, group=None, dst=0):
    """accumulate(begin, count) -> torch tensor (H, W, 4) f32 holding this rank's partial sums
    (rgb) and sample count (a), on the device the process group reduces on.
    resolve(accum, spp) -> RGBA8 image (only called on `dst`).  Returns (rgba | None, accum)."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    begin, count = sample_slice(spp, rank, world)
    accum = accumulate(begin, count)
    assert accum.shape == (height, width, 4) and accum.dtype == torch.float32
    if world > 1:
        dist.reduce(accum, dst=dst, op=dist.ReduceOp.SUM, group=group)
    if rank == dst:
        return resolve(accum, spp), accum
    return None, accum


def render_distributed_gpu(ctx, dev_scene, scene, width, height, seed=0, group=None):
    """The GPU instantiation: rtr_render_accumulate_device + NCCL reduce + rtr_resolve_device.

    Three streams can be involved: the library launches on the context's stream, torch allocates and NCCL reduces on
    torch's current stream.  Unless the context was created ON torch's current stream (ctx.stream_handle), each step
    is fenced from the next: the accumulation must be complete before the reduce reads it, and the reduce before the
    resolve."""
    from .render import make_params
    spp = scene.settings.samples_per_pixel
    shared = ctx.stream_handle is not None and ctx.stream_handle == torch.cuda.current_stream().cuda_stream

    def accumulate(begin, count):
        acc = torch.empty((height, width, 4), dtype=torch.float32, device="cuda")
        if count == 0:
            return acc.zero_()
        if not shared:
            torch.cuda.current_stream().synchronize()  # the allocation (and any earlier use of the block) is settled
        p = make_params(scene, width, height, seed=seed, sample_begin=begin, sample_count=count)
        dev_scene.render_accumulate_device(p, acc.data_ptr(), want_stats=False)
        if not shared:
            ctx.synchronize()  # render kernels done before NCCL (torch's stream) reads acc
        return acc

    def resolve(acc, n):
        rgba = torch.empty((height, width, 4), dtype=torch.uint8, device="cuda")
        if not shared:
            torch.cuda.current_stream().synchronize()  # the reduce has landed in acc
        ctx.resolve_device(acc.data_ptr(), width, height, n, rgba.data_ptr())
        if not shared:
            ctx.synchronize()  # rgba is complete before torch reads it
        return rgba

    return render_distributed(accumulate, resolve, width, height, spp, group)
