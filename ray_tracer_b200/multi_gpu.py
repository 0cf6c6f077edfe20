"""Host-side plumbing of the multi-GPU decompositions. One process per GPU, scene replicated.

* sample split (throughput): rank r of N renders progressive frames f = first + r, first + r + N, ...
  (frames have independent seeds: shaders/raytrace.comp:562-564) into a rank-local float4 accumulator;
  one reduce(sum) onto rank 0 per batch.
* row bands (1-spp interactive frames): rank r renders rows [r*H/N, (r+1)*H/N); per-pixel seeds depend
  only on (x, y, W, frame), so bands are exact; bands are gathered on rank 0.

The renderer is passed in as an object with the methods of `ComputePass` (`render_frames`, `sync`), so the same
logic is exercised on CPU under gloo with a stand-in renderer (tests/test_multi_gpu_host.py). Nothing here
computes radiance. The product path for several GPUs is INSIDE the library (csrc/comm.cu: b200pt_comm_init_rank,
b200pt_render_frames_split, b200pt_reduce_accum, b200pt_render_frame_interleaved — used by bench.py and by
b200pt_headless --gpus N); these torch.distributed helpers remain for callers that already hold a process group.

Stream ordering: the renderer works on its context's stream, torch's collectives on torch's current stream. Both
helpers call `render.sync()` before the collective and synchronise torch's stream before returning, so they are
correct whichever stream the context was given.
"""
import torch
import torch.distributed as dist


def frame_share(first_frame, n_frames, rank, world):
    """Frames of [first_frame, first_frame + n_frames) that rank `rank` renders -> (first, count, stride)."""
    if n_frames <= rank:
        return first_frame + rank, 0, world
    return first_frame + rank, (n_frames - rank + world - 1) // world, world


def row_band(height, rank, world):
    """Rows [begin, end) of rank `rank`; the bands tile [0, height) exactly, sizes differ by at most one row."""
    base, extra = divmod(height, world)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


def render_frames_split(render, push_constants, first_frame, n_frames, accum, total=None, group=None):
    """Sample split of one batch. `render` = ComputePass-like, already accumulating into `accum`
    (a W*H*4 float32 tensor on this rank's device, attached with attach_accum). After the call rank 0's
    `total` (or `accum` when total is None) has gained the sum over all ranks; `accum` of the other ranks
    (and of rank 0 when `total` is given) is zeroed for the next batch."""
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    first, count, stride = frame_share(first_frame, n_frames, rank, world)
    if count:
        render.render_frames(push_constants, first, count, frame_stride=stride)
    _sync(render)                                   # the collective below runs on torch's stream, the renderer on its own
    if world == 1:
        if total is not None:
            total.add_(accum)
            accum.zero_()
        return
    dist.reduce(accum, dst=0, op=dist.ReduceOp.SUM, group=group)
    if rank == 0 and total is not None:
        total.add_(accum)
    if rank != 0 or total is not None:
        accum.zero_()
    _sync_torch(accum)                              # zero_ / add_ must have landed before the next batch renders into accum


def render_frame_bands(render, push_constants, frame, accum, group=None):
    """Row-band split of ONE progressive frame; rank 0 ends with the whole frame in `accum` (H x W x 4)."""
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    H = accum.shape[0]
    r0, r1 = row_band(H, rank, world)
    if r1 > r0:
        render.render_frames(push_constants, frame, 1, frame_stride=1, rows=(r0, r1))
    _sync(render)
    if world == 1:
        return
    # bands differ in size by at most one row: gather padded bands, copy the valid rows on rank 0
    pad_rows = (H + world - 1) // world
    mine = torch.zeros((pad_rows,) + tuple(accum.shape[1:]), dtype=accum.dtype, device=accum.device)
    mine[: r1 - r0] = accum[r0:r1]
    parts = [torch.empty_like(mine) for _ in range(world)] if rank == 0 else None
    dist.gather(mine, parts, dst=0, group=group)
    if rank == 0:
        for r in range(1, world):
            b0, b1 = row_band(H, r, world)
            accum[b0:b1] = parts[r][: b1 - b0]
    _sync_torch(accum)


def _sync(render):
    s = getattr(render, "sync", None)
    if s is not None:
        s()


def _sync_torch(t):
    if t.is_cuda:
        torch.cuda.current_stream(t.device).synchronize()
