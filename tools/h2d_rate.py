"""Pinned host -> device copy rate for one bench step's worth of uint8 frames (64 x 752 x 480 bytes), the bound of bench.py's e2e leg.
Run on the GPU box: python tools/h2d_rate.py"""
import torch
n = 64 * 752 * 480
h = torch.empty(n, dtype=torch.uint8, pin_memory=True)
d = torch.empty(n, dtype=torch.uint8, device="cuda")
s = torch.cuda.Stream()
with torch.cuda.stream(s):
    for _ in range(5):
        d.copy_(h, non_blocking=True)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(s)
    for _ in range(50):
        d.copy_(h, non_blocking=True)
    b.record(s)
s.synchronize()
ms = a.elapsed_time(b) / 50
print(f"H2D {n / 1e6:.1f} MB pinned: {ms:.3f} ms per copy = {n / ms / 1e6:.1f} GB/s -> at most {64 / ms * 1e3:.0f} frames/s for 64-lane uint8 steps")
