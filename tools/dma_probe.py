import torch, time
dev = torch.device("cuda:0")
n = 25 << 20
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device=dev)
a = torch.randn(8192, 8192, device=dev, dtype=torch.bfloat16)
x = torch.randn(64 << 20, device=dev)
s2 = torch.cuda.Stream()
def t_copy(load):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if load == "gemm":
        for _ in range(6): a @ a
    elif load == "ew":
        for _ in range(12): x.mul_(1.0001)
    with torch.cuda.stream(s2):
        e0.record(); h.copy_(d, non_blocking=True); e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1)
for load in ("none", "gemm", "ew", "none"):
    ts = sorted(t_copy(load) for _ in range(7))
    print(f"D2H 25 MiB while GPU runs {load}: median {ts[3]:.3f} ms = {n/ts[3]/1e6:.1f} GB/s")
# eight chunk copies back to back vs one
def t_chunks(k):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    c = n // k
    with torch.cuda.stream(s2):
        e0.record()
        for i in range(k): h[i*c:(i+1)*c].copy_(d[i*c:(i+1)*c], non_blocking=True)
        e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1)
for k in (1, 5, 8, 16):
    ts = sorted(t_chunks(k) for _ in range(7))
    print(f"{k} chunk copies: median {ts[3]:.3f} ms")
