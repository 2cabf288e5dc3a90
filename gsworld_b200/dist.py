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


def bind_to_gpu_numa_node(device_index: int) -> str:
    """Pin the calling process to the CPUs that are local to GPU `device_index` (its PCIe root's NUMA node).

    With one process per GPU on a two-socket box, pinned host buffers are first-touched, and the per-step H2D / D2H
    copies are driven, from whatever core the scheduler picked; half the ranks then cross the socket interconnect with
    every observation batch.  Call before allocating pinned memory.  Returns a short description (also when nothing
    could be done: no NVML, no sysfs entry, a single node)."""
    import os
    try:
        import pynvml
        pynvml.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        phys = int(vis.split(",")[device_index]) if vis and vis.split(",")[device_index].strip().isdigit() else device_index
        bus = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(phys)).busId
        bus = bus.decode() if isinstance(bus, bytes) else bus
        bdf = bus.lower()[-12:]  # 0000:xx:yy.z
        with open(f"/sys/bus/pci/devices/{bdf}/local_cpulist") as f:
            txt = f.read().strip()
        cpus = set()
        for part in txt.split(","):
            if "-" in part:
                a, b = part.split("-")
                cpus.update(range(int(a), int(b) + 1))
            elif part:
                cpus.add(int(part))
        allowed = os.sched_getaffinity(0)
        use = cpus & allowed
        if not use or use == allowed:
            return f"gpu {device_index} ({bdf}): no narrower local CPU set ({txt})"
        os.sched_setaffinity(0, use)
        return f"gpu {device_index} ({bdf}): bound to {len(use)} local CPUs ({txt})"
    except Exception as e:  # noqa: BLE001
        return f"gpu {device_index}: not bound ({e!r})"


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
    """Every rank's uint8 observations in the memory of the ranks that consume them, WITHOUT a collective kernel.

    Each rank owns `nbuf` symmetric buffers [world, *shard_shape] (torch symmetric memory: the same allocation
    on every GPU of the box, each mapped into every process over NVLink / NVSwitch).  `roots` names the ranks
    that need the whole batch: None = every rank (an all-gather), [0] = the policy rank only ("or
    gather-style if only rank 0 needs it"), which cuts the NVLink traffic into every other GPU to nothing.
    Step k of rank r:

        begin(k, engine)    the render stream waits until this rank's slot of buffer k % nbuf may be rewritten, and
                            libgsw's delivery streams are gated (gsw_delivery_gate) behind the exchange stream, so
                            the pushes of step k start only once the readers of the buffer's previous use are done
        local_out(k)        slot r of the rank's own buffer: the renderer composites straight into it (obs_out)
        destinations(k)     slot r of the same buffer on every OTHER root: libgsw's peer stream pushes each finished
                            frame group there with the copy engines (no SMs, no NCCL kernel) while later frames
                            are still being composited (gsw_frames_desc.peer_rgb_u8, GSW_DELIVER_NO_JOIN)
        publish(k, engine)  on the exchange's own stream: wait for this rank's pushes, then tell the consumers.
                            roots=None: one device-side barrier over all ranks.  roots given: a signal to every
                            root (symmetric-memory signal pad), and a root waits for one signal per producer.
                            The render stream is not involved and goes on with step k + 1.
        wait_ready(k)       the calling stream waits until gathered(k) holds every rank's frames (roots only)
        release(k)          the calling stream's reads of gathered(k) are enqueued: the buffer may be reused

    Reuse is enforced, not left to convention.  A buffer is rewritten by step k + nbuf.  Before that step's pushes
    may start, every consumer's reads of step k must have completed: release(k) records an event, the exchange
    stream waits for it before it enters the barrier (roots=None) or before it returns a credit signal to every
    producer (roots given), and begin(k + nbuf) orders the deliveries behind that.  A consumer that called
    wait_ready(k) and has not called release(k) when the buffer comes up for reuse makes publish() raise instead
    of racing.  One-step slack with nbuf = 2: render step k + 1, consume step k (wait_ready, read, release), publish
    step k + 1.  The reference has no multi-GPU path; `gather_observations` stays as the fallback."""

    _CH_DATA, _CH_CREDIT = 1, 2  # signal-pad channels (0 is the barrier's)
    _TIMEOUT_MS = 30000          # a device-side wait that outlives this traps instead of hanging the GPU

    @staticmethod
    def available() -> bool:
        try:
            import torch.distributed._symmetric_memory  # noqa: F401
            return torch.cuda.is_available() and dist.is_initialized() and dist.get_backend() == "nccl"
        except Exception:
            return False

    def __init__(self, shard_shape, device, group=None, nbuf: int = 2, roots=None):
        import torch.distributed._symmetric_memory as symm_mem
        self.group = group if group is not None else dist.group.WORLD
        self.rank, self.world = dist.get_rank(self.group), dist.get_world_size(self.group)
        self.shard_shape = tuple(int(x) for x in shard_shape)
        self.nbuf = nbuf
        self.roots = None if roots is None else sorted(set(int(r) for r in roots))
        if self.roots is not None and not all(0 <= r < self.world for r in self.roots):
            raise ValueError(f"roots {roots} outside world of {self.world}")
        self.is_root = self.roots is None or self.rank in self.roots
        full = (self.world,) + self.shard_shape
        self._local, self._hdl, self._dst = [], [], []
        targets = range(self.world) if self.roots is None else self.roots
        for _ in range(nbuf):
            t = symm_mem.empty(full, dtype=torch.uint8, device=device)
            h = symm_mem.rendezvous(t, self.group)
            self._local.append(t)
            self._hdl.append(h)
            self._dst.append([h.get_buffer(p, full, torch.uint8)[self.rank] for p in targets if p != self.rank])
        self.stream = torch.cuda.Stream(device=device)
        self._ready = [torch.cuda.Event() for _ in range(nbuf)]
        self._released = [None] * nbuf   # event recorded by release() of the buffer's last use
        self._reading = [False] * nbuf   # wait_ready() seen, release() not yet
        self._credit_due = [False] * nbuf  # roots mode, on a root: producers still wait for this buffer's credit
        self._used = [False] * nbuf

    # -- producer side --
    def begin(self, k: int, engine=None):
        b = k % self.nbuf
        cur = torch.cuda.current_stream()
        if self._used[b]:  # this rank's own pushes of the buffer's previous use are done (they read local_out)
            cur.wait_event(self._ready[b])
        if self._released[b] is not None:  # ... and so are the local consumer's reads of it
            cur.wait_event(self._released[b])
        if self.roots is not None and self._used[b]:
            # credit from every root: its consumer has released step k - nbuf (a root needs none from itself)
            with torch.cuda.stream(self.stream):
                for r in self.roots:
                    if r != self.rank:
                        self._hdl[b].wait_signal(r, self._CH_CREDIT, self._TIMEOUT_MS)
        if engine is not None:  # the pushes of step k wait for the barrier / credits above, the compositing does not
            engine.delivery_gate(self.stream)

    def local_out(self, k: int) -> torch.Tensor:
        return self._local[k % self.nbuf][self.rank]

    def destinations(self, k: int):
        return self._dst[k % self.nbuf]

    def publish(self, k: int, engine):
        b = k % self.nbuf
        nxt = (k + 1) % self.nbuf  # the buffer step k + 1 rewrites
        if self._reading[nxt]:
            raise RuntimeError(f"PeerObservationExchange: gathered({k + 1 - self.nbuf}) was handed out by wait_ready() "
                               f"and never release()d, but step {k + 1} is about to overwrite it")
        engine.delivery_wait(self.stream)
        with torch.cuda.stream(self.stream):
            if self.roots is None:
                # entering barrier k certifies that this rank no longer reads the buffer step k + 1 will rewrite
                if self._released[nxt] is not None:
                    self.stream.wait_event(self._released[nxt])
                self._hdl[b].barrier(0, self._TIMEOUT_MS)
            else:
                if self._credit_due[nxt]:  # nobody consumed that step: hand its credit back now (exactly one per use)
                    self._send_credit(nxt)
                for r in self.roots:
                    if r != self.rank:
                        self._hdl[b].put_signal(r, self._CH_DATA, self._TIMEOUT_MS)
                if self.is_root:
                    for p in range(self.world):
                        if p != self.rank:
                            self._hdl[b].wait_signal(p, self._CH_DATA, self._TIMEOUT_MS)
                    self._credit_due[b] = True
            self._ready[b].record()
        self._used[b] = True

    def _send_credit(self, b: int):  # on self.stream
        for p in range(self.world):
            if p != self.rank:
                self._hdl[b].put_signal(p, self._CH_CREDIT, self._TIMEOUT_MS)
        self._credit_due[b] = False

    # -- consumer side --
    def wait_ready(self, k: int):
        if not self.is_root:
            raise RuntimeError("wait_ready: this rank is not a root of the exchange")
        torch.cuda.current_stream().wait_event(self._ready[k % self.nbuf])
        self._reading[k % self.nbuf] = True

    def gathered(self, k: int) -> torch.Tensor:
        return self._local[k % self.nbuf]

    def release(self, k: int):
        """The current stream's reads of gathered(k) are all enqueued; the buffer may be rewritten after them."""
        b = k % self.nbuf
        ev = torch.cuda.Event()
        ev.record()
        self._released[b] = ev
        self._reading[b] = False
        if self.roots is not None and self.is_root and self._credit_due[b]:  # hand the credit back to every producer
            with torch.cuda.stream(self.stream):
                self.stream.wait_event(ev)
                self._send_credit(b)
