"""Timing of the int8-slice GEMM at the short-K moments shape (slicing included, constant across debug variants)."""
import os
import torch
from gaussian_processes_b200 import native
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(3)
M, N, K = 100000, 500, 500
A = torch.randn(M, K, generator=g, dtype=torch.float64, device=dev)
B = torch.randn(N, K, generator=g, dtype=torch.float64, device=dev)
for _ in range(2):
    native.gemm_i8emu(A, B, 8)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    native.gemm_i8emu(A, B, 8)
e1.record()
torch.cuda.synchronize()
print(f"dbg={os.environ.get('SGP_OZ_DEBUG', '0')}: {e0.elapsed_time(e1) / 5:.3f} ms per call incl. slicing", flush=True)
