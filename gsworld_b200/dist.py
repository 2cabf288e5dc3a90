"""Environment sharding across the GPUs of one box.

The observation path has no cross-env dependency (reference loop body gs_world_wrapper.py:242 touches only
row i), so rank r of G simply owns a contiguous block of envs and renders it with its own libgsw context;
the scene is replicated.  The single exchange is an all-gather of the uint8 observations so that the
policy sees the whole batch (NCCL over NVLink on GPUs; the same code runs over gloo on CPU tensors, which
is how tests/test_dist_gloo.py covers it).  The reference has no distributed code at all.
"""
from __future__ import annotations

from typing import Dict, Tuple

import torch
import torch.distributed as dist


def shard_envs(num_envs: int, rank: int, world: int) -> Tuple[int, int]:
    """[lo, hi) of the envs rank `rank` owns; blocks differ by at most one env."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world of {world}")
    base, extra = divmod(num_envs, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def gather_observations(local: Dict[str, torch.Tensor], num_envs: int, group=None) -> Dict[str, torch.Tensor]:
    """All-gather {cam: uint8 (B_local,H,W,3)} into {cam: uint8 (num_envs,H,W,3)} on every rank.

    One collective per call: cameras are stacked into a single buffer; uneven shards are padded to the
    largest block and trimmed after the exchange."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return dict(local)
    rank = dist.get_rank(group)
    cams = sorted(local)
    lo, hi = shard_envs(num_envs, rank, world)
    bmax = -(-num_envs // world)
    first = local[cams[0]]
    H, W = first.shape[1], first.shape[2]
    send = torch.zeros((len(cams), bmax, H, W, 3), dtype=torch.uint8, device=first.device)
    for i, c in enumerate(cams):
        if local[c].shape[0] != hi - lo:
            raise ValueError(f"{c}: rank {rank} should hold {hi - lo} envs, got {local[c].shape[0]}")
        send[i, : hi - lo] = local[c]
    recv = torch.empty((world,) + tuple(send.shape), dtype=torch.uint8, device=first.device)
    dist.all_gather_into_tensor(recv.view(-1), send.view(-1), group=group)
    out = {}
    for i, c in enumerate(cams):
        parts = []
        for r in range(world):
            rlo, rhi = shard_envs(num_envs, r, world)
            parts.append(recv[r, i, : rhi - rlo])
        out[c] = torch.cat(parts, 0)
    return out
