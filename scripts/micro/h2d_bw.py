import torch, time
for mb in (2, 16, 64, 170, 1024):
    n = mb << 20
    h = torch.empty(n, dtype=torch.uint8).pin_memory()
    d = torch.empty(n, dtype=torch.uint8, device="cuda")
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for _ in range(3): d.copy_(h, non_blocking=True)
        s.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = max(4, 2048 // mb)
        e0.record()
        for _ in range(reps): d.copy_(h, non_blocking=True)
        e1.record(); s.synchronize()
    print(f"{mb:5d} MiB copies: {n * reps / (e0.elapsed_time(e1) * 1e-3) / 1e9:6.1f} GB/s")
