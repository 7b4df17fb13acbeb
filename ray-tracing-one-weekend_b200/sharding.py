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
