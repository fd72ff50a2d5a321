"""Multi-GPU path on the CPU: the frame partition (no collective on the data path) and a
world-size-2 gloo run in which each rank rasterises its shard with the oracle; together the shards
must equal the single-process result frame for frame."""
import os
import socket
import sys

import numpy as np
import pytest

from conftest import ROOT, load_sub


def test_shard_frames_partition():
    sh = load_sub("sharding")
    for n in (0, 1, 7, 8, 1024, 1025):
        for ws in (1, 2, 3, 4, 8):
            got = [sh.shard_frames(n, ws, r) for r in range(ws)]
            assert got[0][0] == 0 and sum(c for _, c in got) == n
            for (f0, c0), (f1, _) in zip(got, got[1:]):
                assert f1 == f0 + c0
            assert max(c for _, c in got) - min(c for _, c in got) <= 1
    with pytest.raises(ValueError):
        sh.shard_frames(4, 2, 2)


def test_shard_points_ragged():
    sh = load_sub("sharding")
    offs = np.array([0, 5, 5, 12, 20, 21], np.int64)
    parts = [sh.shard_points(offs, 2, r) for r in range(2)]
    assert parts[0][:4] == (0, 3, 0, 12) and parts[1][:4] == (3, 2, 12, 21)
    assert parts[0][4].tolist() == [0, 5, 5, 12] and parts[1][4].tolist() == [0, 8, 9]


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    from conftest import load_sub as ls
    from oracle import lidar_oracle as lo
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sh, synth = ls("sharding"), ls("synth")
    pts, offs = synth.make_batch(5, "hdl32", seed0=300, n_az=40)
    first, n_local, lo_p, hi_p, loffs = sh.shard_points(offs, world, rank)
    counts = sh.gather_frame_counts(n_local)
    local = pts[lo_p:hi_p]
    out = []
    for f in range(n_local):
        P = synth.as_reference_array(local[loffs[f]:loffs[f + 1]])
        fv = lo.front_view(P)
        out.append((first + f, fv["winner"], lo.bev_counts(P, (2, 2), 120)))
    dist.barrier()
    dist.destroy_process_group()
    q.put((rank, counts, out))


def test_two_rank_gloo_shards_equal_single_process():
    import torch.multiprocessing as mp
    from oracle import lidar_oracle as lo
    synth = load_sub("synth")
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    pts, offs = synth.make_batch(5, "hdl32", seed0=300, n_az=40)
    seen = {}
    for rank, counts, out in res:
        assert counts == [3, 2]
        for f, win, cnt in out:
            seen[f] = (win, cnt)
    assert sorted(seen) == list(range(5))
    for f in range(5):
        P = synth.as_reference_array(pts[offs[f]:offs[f + 1]])
        assert np.array_equal(seen[f][0], lo.front_view(P)["winner"])
        assert np.array_equal(seen[f][1], lo.bev_counts(P, (2, 2), 120))
