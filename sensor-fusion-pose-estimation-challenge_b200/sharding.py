"""Frame sharding of a batch of sweeps over the GPUs of one box.

Sweeps are independent (nothing in modules/lidar/process/extract_rosbag_lidar.py:18-86 or
modules/lidar/common/birds_eye_view_generator.py:31-54 carries state between frames), so rank r of
G owns the contiguous frames [r*B/G, (r+1)*B/G) and there is no data-path collective.  Pure host
arithmetic; torch.distributed is only needed by gather_frame_counts (a sanity all-gather used by
the tests and the bench, off the timed path).
"""
import numpy as np


def shard_frames(n_frames, world_size, rank):
    """(first_frame, n_local): contiguous split, remainder spread over the first ranks."""
    if world_size < 1 or not 0 <= rank < world_size or n_frames < 0:
        raise ValueError("bad shard request")
    base, rem = divmod(int(n_frames), int(world_size))
    first = rank * base + min(rank, rem)
    return first, base + (1 if rank < rem else 0)


def shard_points(frame_offsets, world_size, rank):
    """For a ragged batch (int64 offsets, n_frames+1): this rank's (first_frame, n_local,
    point_lo, point_hi, local_offsets) -- local_offsets start at 0."""
    offs = np.asarray(frame_offsets, dtype=np.int64)
    first, n_local = shard_frames(len(offs) - 1, world_size, rank)
    lo, hi = int(offs[first]), int(offs[first + n_local])
    return first, n_local, lo, hi, offs[first:first + n_local + 1] - lo


def gather_frame_counts(n_local, group=None):
    """All ranks' local frame counts (list of ints) through torch.distributed (gloo or nccl)."""
    import torch
    import torch.distributed as dist
    ws = dist.get_world_size(group)
    dev = "cuda" if dist.get_backend(group) == "nccl" else "cpu"
    mine = torch.tensor([int(n_local)], dtype=torch.int64, device=dev)
    out = [torch.zeros_like(mine) for _ in range(ws)]
    dist.all_gather(out, mine, group=group)
    return [int(t.item()) for t in out]
