"""Multi-GPU partitioning of the pixel/sample loop: samples-per-pixel are split across ranks, every
rank renders every pixel for its own sample range into a private fp32 sum buffer, and the buffers are summed onto
rank 0 with one reduce (NCCL on GPUs; the CPU tests run the same code over gloo).  The Philox counter is
(pixel, sample, ...), so the union of the ranks' sample sets -- hence the image -- does not depend on the world size.

Host logic only: no rendering code here."""


def sample_range(spp, rank, world):
    """Contiguous block [begin, begin+count) of the spp samples owned by `rank`; blocks tile [0, spp) exactly."""
    if world < 1 or not (0 <= rank < world) or spp < 0:
        raise ValueError("bad shard arguments")
    begin = spp * rank // world
    return begin, spp * (rank + 1) // world - begin


def ray_range(n, rank, world):
    """Contiguous chunk of a ray batch for the closest-hit harness (config 5): results are concatenated, no reduce."""
    return sample_range(n, rank, world)


def reduce_sums(accum, dst=0, group=None):
    """Sum the ranks' accumulation buffers onto `dst` (torch tensor, in place).  No-op without a process group."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(accum, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return accum


def band_ranges(height, rank, world, band_rows=16):
    """Image-tile partitioning (SURVEY 8(e) "partitioning 2", for frames whose accumulators should not be replicated):
    the frame is cut into bands of `band_rows` rows and dealt round-robin, band b to rank b % world -- interleaved, because
    cost varies strongly across a frame (open front vs glass and smoke in the Cornell box).  Returns this rank's
    [(row_begin, row_count), ...]; over all ranks the bands tile [0, height) exactly.  Each band is rendered with
    rtow_render_params.row_begin/row_count into a band-sized buffer; no arithmetic reduce is needed, the bands are
    gathered (gather_bands)."""
    if world < 1 or not (0 <= rank < world) or height < 0 or band_rows < 1:
        raise ValueError("bad band arguments")
    out = []
    for b, y in enumerate(range(0, height, band_rows)):
        if b % world == rank:
            out.append((y, min(band_rows, height - y)))
    return out


def gather_bands(frame, my_bands, band_tensors, width, dst=0, group=None, band_rows=16):
    """Assemble the full frame (H x W x C tensor on `dst`) from every rank's bands.  `band_tensors[i]` holds this rank's
    band my_bands[i] (rows x W x C).  Disjoint tiles: the bands are exchanged with one all_gather of a padded stack per
    rank (no summation).  Without a process group it just copies the local bands."""
    import torch
    import torch.distributed as dist
    H = frame.shape[0]
    world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
    rank = dist.get_rank(group) if world > 1 else 0
    if world == 1:
        for (y, n), t in zip(my_bands, band_tensors):
            frame[y:y + n].copy_(t)
        return frame
    n_bands = (H + band_rows - 1) // band_rows
    per_rank = (n_bands + world - 1) // world
    C = frame.shape[2]
    mine = torch.zeros(per_rank, band_rows, width, C, dtype=frame.dtype, device=frame.device)
    for i, t in enumerate(band_tensors):
        mine[i, :t.shape[0]].copy_(t)
    parts = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(parts, mine, group=group)
    if rank == dst:
        for r in range(world):
            for i, (y, n) in enumerate(band_ranges(H, r, world, band_rows)):
                frame[y:y + n].copy_(parts[r][i, :n])
    return frame
