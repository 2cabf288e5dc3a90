"""N > 1 host logic on CPU: env sharding and the observation all-gather over gloo, world_size 2 and 3."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gsworld_b200.dist import gather_observations, shard_envs


def test_shard_envs_partitions_exactly():
    for n in (1, 2, 7, 64, 65, 256):
        for world in (1, 2, 3, 4, 8):
            spans = [shard_envs(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            for (a, b), (c, d) in zip(spans, spans[1:]):
                assert b == c
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_envs(4, 4, 4)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, num_envs, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        lo, hi = shard_envs(num_envs, rank, world)
        H, W = 4, 6

        def frame(cam_id, env):  # every (camera, env) has a distinct constant value
            return torch.full((H, W, 3), (cam_id * 100 + env) % 251, dtype=torch.uint8)

        local = {c: torch.stack([frame(i, e) for e in range(lo, hi)]) if hi > lo
                 else torch.zeros((0, H, W, 3), dtype=torch.uint8) for i, c in enumerate(("right_cam", "wrist_cam"))}
        full = gather_observations(local, num_envs)
        ok = True
        for i, c in enumerate(("right_cam", "wrist_cam")):
            want = torch.stack([frame(i, e) for e in range(num_envs)])
            ok = ok and full[c].shape == want.shape and torch.equal(full[c], want)
        q.put((rank, ok))
    except Exception as e:  # report instead of leaving the parent to time out
        q.put((rank, repr(e)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,num_envs", [(2, 8), (2, 5), (3, 7)])
def test_gather_observations_gloo(world, num_envs):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, num_envs, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=90) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(r for r, _ in results) == list(range(world))
    assert all(ok is True for _, ok in results), results
