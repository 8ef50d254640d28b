"""Multi-GPU sharding of the render pass: one process per GPU, no mid-render communication.

Every (pixel, sampleIndex) path is independent and its random stream depends only on (pixel, seed, sampleIndex)
(samplers.h:455-458), so both partitions below are exact:

  sample-pass sharding   rank r renders sample indices r, r+N, ... over the whole image; the fp64 film buffers
                         {rgbSum[3], weightSum} (film.h:251-267) are summed onto rank 0 with one reduce
  band sharding          rank r renders an interleaved set of horizontal bands of the image (BASELINE C4's "tiles
                         sharded over 8 GPUs"); supports are disjoint, the same reduce acts as the gather

The exchange step is a single `torch.distributed.reduce(SUM)` of H*W*4 doubles (NCCL over NVLink on the GPU box,
gloo in the CPU tests).  Inputs (scene, grids, cache, buffers) are replicated: every rank loads them itself.
"""


def sample_pass_shard(rank, world, spp_total):
    """-> (spp_begin, spp_end, spp_stride) for vspg_render: rank r takes r, r+world, ... below spp_total."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world {world}")
    return rank, spp_total, world


def band_shard(rank, world, height, band=8):
    """-> list of (y0, y1) row bands owned by `rank`: bands of `band` rows dealt round-robin so that every rank sees
    every part of the image (load balance across sky / medium / ground)."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world {world}")
    out = []
    for i, y0 in enumerate(range(0, height, band)):
        if i % world == rank:
            out.append((y0, min(height, y0 + band)))
    return out


def paths_of_shard(width, height, spp_total, rank, world, mode="samples", band=8):
    if mode == "samples":
        b, e, s = sample_pass_shard(rank, world, spp_total)
        return width * height * len(range(b, e, s))
    return width * sum(y1 - y0 for y0, y1 in band_shard(rank, world, height, band)) * spp_total


def render_shard(render, width, height, spp_total, rank, world, seed, mode="samples", band=8):
    """Drives `render(spp_begin, spp_end, spp_stride, window, seed)` (Context.render or a stand-in) over this rank's
    shard.  The caller clears the film first and combines afterwards with combine_film()."""
    if mode == "samples":
        b, e, s = sample_pass_shard(rank, world, spp_total)
        if b < e:
            render(b, e, s, (0, 0, width, height), seed)
    elif mode == "bands":
        for y0, y1 in band_shard(rank, world, height, band):
            render(0, spp_total, 1, (0, y0, width, y1), seed)
    else:
        raise ValueError(f"unknown sharding mode {mode!r}")


def combine_film(film_tensor, world, dst=0):
    """The one exchange step: sum the [H*W*4] fp64 film buffers onto rank `dst` (in place on dst)."""
    if world > 1:
        import torch.distributed as dist
        dist.reduce(film_tensor, dst=dst, op=dist.ReduceOp.SUM)
    return film_tensor
