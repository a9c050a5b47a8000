"""Multi-GPU sharding of one frame: one process per GPU, torch.distributed for
the plumbing. The path shards with NO data-path collective for row tiles (pixels are
independent) and ONE real exchange for sample ranges (a sum of per-rank partial colours):

  * row tiles      rank g renders rows [g*H/G, (g+1)*H/G) (``row_range``); the tiles are
                   gathered to rank 0 (``gather_rows``), or written straight into rank 0's
                   frame buffer when the output pointers are peer-mapped (rt_render takes
                   any device address).
  * sample ranges  rank g renders frame_count in [g*S/G, (g+1)*S/G) into an FP32 partial sum
                   (RT_FLAG_ACCUMULATE), ``reduce_samples`` sums the partials on rank 0 and
                   scales by 1/S. Position/normal come from the rank that owns sample 0.

The host logic is backend-agnostic (gloo on CPU in tests/test_sharding.py, nccl on GPUs).
"""
from __future__ import annotations


def row_range(height: int, rank: int, world: int, align: int = 4) -> tuple:
    """Rows of ``rank``: contiguous, 4-aligned boundaries (the kernel's tile height) so that
    no warp tile straddles two ranks; the last rank takes the remainder."""
    per = (height + world - 1) // world
    per = (per + align - 1) // align * align
    a = min(rank * per, height)
    b = min(a + per, height)
    return a, b


def sample_range(num_samples: int, rank: int, world: int) -> tuple:
    """[first, first+count) of the frame_count indices rendered by ``rank``."""
    base, rem = divmod(num_samples, world)
    first = rank * base + min(rank, rem)
    return first, base + (1 if rank < rem else 0)


def gather_rows(buf, height: int, world: int, dst: int = 0, group=None):
    """Gather the row tiles of a full-frame (H, W, C) tensor onto ``dst``: every rank sends its
    own rows; ``dst`` receives them in place. Equal-size tiles use one all_gather-free
    gather; ragged last tiles go point to point."""
    import torch.distributed as dist
    rank = dist.get_rank(group)
    if world == 1:
        return buf
    ops = []
    if rank == dst:
        for r in range(world):
            if r == dst:
                continue
            a, b = row_range(height, r, world)
            if b > a:
                ops.append(dist.P2POp(dist.irecv, buf[a:b], r, group))
    else:
        a, b = row_range(height, rank, world)
        if b > a:
            ops.append(dist.P2POp(dist.isend, buf[a:b], dst, group))
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()
    return buf


def reduce_samples(partial_sum, total_samples: int, dst: int = 0, group=None):
    """Sum the per-rank partial colour sums on ``dst`` and turn them into the mean (alpha
    restored to 1). One collective: the path's only real exchange step."""
    import torch.distributed as dist
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(partial_sum, dst=dst, op=dist.ReduceOp.SUM, group=group)
        if dist.get_rank(group) != dst:
            return partial_sum
    partial_sum[..., :3] *= 1.0 / float(total_samples)
    partial_sum[..., 3] = 1.0
    return partial_sum
