"""Kernel-only timing of the int8-slice GEMM pieces (the slices are made once, outside the timed region)."""
import sys
import torch
from gaussian_processes_b200 import native

dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(3)
M, N, K = 65536, 500, 1600
A = torch.randn(M, K, generator=g, dtype=torch.float64, device=dev)
B = torch.randn(N, K, generator=g, dtype=torch.float64, device=dev)
for S in (8, 7):
    for _ in range(2):
        native.gemm_i8emu(A, B, S)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        native.gemm_i8emu(A, B, S)
    e1.record()
    torch.cuda.synchronize()
    print(f"S={S}: {e0.elapsed_time(e1) / 5:.3f} ms per call incl. slicing", flush=True)
