"""CPU, world_size 2 over gloo: the frame-sharding host logic used for N > 1 (round-robin ownership, batched launch lists, gather
to rank 0 ordered by frame index, max-over-ranks timing reduction)."""
import os
import socket

import pytest


def _worker(rank, world, port, n_frames, q):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    import torch
    import torch.distributed as dist
    import mcslam_b200  # noqa: F401
    from mcslam_b200 import sharding as S
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = S.frames_for_rank(n_frames, rank, world)
    local = {}
    for b in S.batches(mine, 3):
        assert 1 <= len(b) <= 3
        for f in b:
            local[f] = ("frame", f, rank)
    out = S.gather_by_frame(local, n_frames, dist)
    t = torch.tensor([10.0 + rank], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dist.barrier()
    if rank == 0:
        q.put((out, float(t.item())))
    dist.destroy_process_group()


def test_frame_sharding_world2_gloo():
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    n = 11
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n, q)) for r in range(2)]
    for p in procs:
        p.start()
    out, tmax = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert [o[1] for o in out] == list(range(n))
    assert [o[2] for o in out] == [f % 2 for f in range(n)]
    assert tmax == 11.0


def test_sharding_single_process_and_errors():
    import mcslam_b200  # noqa: F401
    from mcslam_b200 import sharding as S
    assert S.frames_for_rank(10, 1, 4) == [1, 5, 9]
    assert S.frames_for_rank(3, 3, 4) == []
    assert S.batches([0, 1, 2, 3, 4], 2) == [[0, 1], [2, 3], [4]]
    assert S.gather_by_frame({1: "b", 0: "a"}, 2) == ["a", "b"]
    with pytest.raises(ValueError):
        S.gather_by_frame({0: "a"}, 2)
