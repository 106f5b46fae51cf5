"""Pinned host -> device copy rate of this box at the e2e step's transfer sizes (the roofline of `e2e` in bench.py)."""
import torch

dev = torch.device("cuda", 0)
for mb in (2.1, 4.2, 6.3, 14.7, 64.0):
    n = int(mb * 1e6)
    h = torch.empty(n, dtype=torch.uint8, pin_memory=True)
    d = torch.empty(n, dtype=torch.uint8, device=dev)
    for _ in range(5):
        d.copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50):
        d.copy_(h, non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    print(f"{mb:5.1f} MB: {50 * n / (e0.elapsed_time(e1) * 1e-3) / 1e9:.1f} GB/s")
