"""Pinned host -> device copy bandwidth of this box (the e2e number is bound by it)."""
import torch
n = 160_000_000   # 1.28 GB of fp64
h = torch.empty(n, dtype=torch.float64).pin_memory()
d = torch.empty(n, dtype=torch.float64, device="cuda:0")
for chunks in (1, 7, 32):
    best = 1e9
    for _ in range(4):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        step = n // chunks
        for c in range(chunks):
            d[c * step:(c + 1) * step].copy_(h[c * step:(c + 1) * step], non_blocking=True)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    print(f"chunks={chunks}: {n * 8 / best / 1e6:.1f} GB/s ({best:.2f} ms)")
s2 = torch.cuda.Stream()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize()
e0.record()
half = n // 2
d[:half].copy_(h[:half], non_blocking=True)
with torch.cuda.stream(s2):
    d[half:].copy_(h[half:], non_blocking=True)
s2.synchronize()
e1.record()
torch.cuda.synchronize()
print(f"two streams: {n * 8 / e0.elapsed_time(e1) / 1e6:.1f} GB/s")
