"""Feasibility probe (N >= 2, torchrun): torch symmetric memory rendezvous, peer-mapped buffers, copy-engine pushes."""
import os, time, torch, torch.distributed as dist
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
try:
    import torch.distributed._symmetric_memory as symm_mem
    n = 25 * 1024 * 1024
    t = symm_mem.empty((world, n), dtype=torch.uint8, device=dev)
    hdl = symm_mem.rendezvous(t, dist.group.WORLD)
    print(rank, "rendezvous ok", type(hdl).__name__, flush=True)
    mine = torch.full((n,), rank + 1, dtype=torch.uint8, device=dev)
    peers = [hdl.get_buffer(p, (world, n), torch.uint8) for p in range(world)]
    print(rank, "peer ptrs", [hex(b.data_ptr()) for b in peers], flush=True)
    s = torch.cuda.Stream()
    torch.cuda.synchronize(); dist.barrier()
    for it in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(s):
            e0.record()
            for p in range(world):
                peers[p][rank].copy_(mine, non_blocking=True)
            e1.record()
        s.synchronize()
        ms = e0.elapsed_time(e1)
        print(rank, f"push {world} x 25 MiB: {ms:.3f} ms  ({world * n / ms / 1e6:.0f} GB/s)", flush=True)
    hdl.barrier()
    torch.cuda.synchronize()
    ok = all(int(t[p][0]) == p + 1 and int(t[p][-1]) == p + 1 for p in range(world))
    print(rank, "all shards arrived:", ok, flush=True)
except Exception as e:
    print(rank, "symmetric memory unavailable:", repr(e)[:300], flush=True)
dist.barrier(); dist.destroy_process_group()
