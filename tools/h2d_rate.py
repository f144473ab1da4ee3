"""Floor for bench.py's e2e: how fast does this box move one 3 MiB pinned frame host -> device (back-to-back async copies)?"""
import torch, time
n = 64
h = torch.empty((n, 1024 * 1024 * 3), dtype=torch.uint8).pin_memory()
d = torch.empty_like(h, device="cuda")
s = torch.cuda.Stream()
for rep in range(3):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    with torch.cuda.stream(s):
        for i in range(n):
            d[i].copy_(h[i], non_blocking=True)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print("H2D %.1f us per 3 MiB frame, %.1f GB/s" % (dt / n * 1e6, h.numel() / dt / 1e9))
