"""Measures pinned-host <-> device copy rates for the observation-sized buffer (tuning aid for the e2e path)."""
import torch

dev = torch.device("cuda:0")
for mb in (3, 6, 25, 100):
    n = mb << 20
    h = torch.empty(n, dtype=torch.uint8).pin_memory()
    d = torch.empty(n, dtype=torch.uint8, device=dev)
    for name, dst, src in (("d2h", h, d), ("h2d", d, h)):
        for _ in range(3):
            dst.copy_(src, non_blocking=True)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            dst.copy_(src, non_blocking=True)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        print(f"{name} {mb} MiB: {ms:.3f} ms  {n / ms / 1e6:.1f} GB/s")
