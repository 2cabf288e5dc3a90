"""Environment sharding across the GPUs of one box.

The observation path has no cross-env dependency (reference loop body gs_world_wrapper.py:242 touches only
row i), so rank r of G simply owns a contiguous block of envs and renders it with its own libgsw context;
the scene is replicated.  The single exchange is that of the uint8 observations so that the policy sees the
whole batch: on GPUs through peer memory (PeerObservationExchange: copy-engine pushes over NVLink, overlapped
with the blend), else an all-gather (gather_observations: NCCL, or gloo on CPU tensors, which is how
tests/test_dist_gloo.py covers it).  The reference has no distributed code at all.
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


class PeerObservationExchange:
    """Every rank's uint8 observations in every rank's memory WITHOUT a collective kernel.

    Each rank owns `nbuf` symmetric buffers [world, *shard_shape] (torch symmetric memory: the same allocation
    on every GPU of the box, each mapped into every process over NVLink / NVSwitch).  Step k of rank r:

        begin(k)            the render stream waits until buffer k % nbuf may be written again
        local_out(k)        slot r of the rank's own buffer: the renderer composites straight into it (obs_out)
        destinations(k)     slot r of the same buffer on every OTHER rank: libgsw's copy stream pushes each finished
                            frame group there with the copy engines (no SMs, no NCCL kernel) while later frames
                            are still being composited (gsw_frames_desc.peer_rgb_u8, GSW_DELIVER_NO_JOIN)
        publish(k, engine)  on the exchange's own stream: wait for this rank's pushes, then a device-side barrier
                            over all ranks; the render stream is not involved and goes on with step k + 1
        wait_ready(k)       the calling stream waits for that barrier: gathered(k) then holds every rank's frames

    One-step slack, like a double-buffered all-gather: a rank may render step k + 1 while step k is still being
    exchanged; it rewrites a buffer only after the barrier of the buffer's previous use has passed.  Whoever reads gathered(k) must have its reads enqueued (on a stream ordered
    before the rank's next publish) before step k + 1 is published: the barrier of step k + 1 is what separates
    those reads from the writes of step k + nbuf.  The reference has no multi-GPU path; this
    replaces the NCCL all-gather of §8(e) where symmetric memory is available (`available()`),
    `gather_observations` stays as the fallback."""

    @staticmethod
    def available() -> bool:
        try:
            import torch.distributed._symmetric_memory  # noqa: F401
            return torch.cuda.is_available() and dist.is_initialized() and dist.get_backend() == "nccl"
        except Exception:
            return False

    def __init__(self, shard_shape, device, group=None, nbuf: int = 2):
        import torch.distributed._symmetric_memory as symm_mem
        self.group = group if group is not None else dist.group.WORLD
        self.rank, self.world = dist.get_rank(self.group), dist.get_world_size(self.group)
        self.shard_shape = tuple(int(x) for x in shard_shape)
        self.nbuf = nbuf
        full = (self.world,) + self.shard_shape
        self._local, self._hdl, self._dst = [], [], []
        for _ in range(nbuf):
            t = symm_mem.empty(full, dtype=torch.uint8, device=device)
            h = symm_mem.rendezvous(t, self.group)
            self._local.append(t)
            self._hdl.append(h)
            self._dst.append([h.get_buffer(p, full, torch.uint8)[self.rank] for p in range(self.world)
                              if p != self.rank])
        self.stream = torch.cuda.Stream(device=device)
        self._ready = [torch.cuda.Event() for _ in range(nbuf)]
        self._used = [False] * nbuf

    def begin(self, k: int):
        if self._used[k % self.nbuf]:  # barrier of the buffer's previous use
            torch.cuda.current_stream().wait_event(self._ready[k % self.nbuf])

    def local_out(self, k: int) -> torch.Tensor:
        return self._local[k % self.nbuf][self.rank]

    def destinations(self, k: int):
        return self._dst[k % self.nbuf]

    def publish(self, k: int, engine):
        engine.delivery_wait(self.stream)
        with torch.cuda.stream(self.stream):
            self._hdl[k % self.nbuf].barrier()
            self._ready[k % self.nbuf].record()
        self._used[k % self.nbuf] = True

    def wait_ready(self, k: int):
        torch.cuda.current_stream().wait_event(self._ready[k % self.nbuf])

    def gathered(self, k: int) -> torch.Tensor:
        return self._local[k % self.nbuf]
