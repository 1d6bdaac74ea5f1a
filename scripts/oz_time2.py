"""Timing of the int8-slice GEMM at the moments shape (K = 500)."""
import torch
from gaussian_processes_b200 import native
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(3)
M, N, K = 65536, 500, 500
A = torch.randn(M, K, generator=g, dtype=torch.float64, device=dev)
B = torch.randn(N, K, generator=g, dtype=torch.float64, device=dev)
for _ in range(3):
    native.gemm_i8emu(A, B, 8)
torch.cuda.synchronize()
