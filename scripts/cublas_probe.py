"""What does cuBLAS launch for fp64 GEMMs of the hot-path shapes? (kernel name / grid / block / regs / smem via ncu)"""
import torch
d = torch.device("cuda:0")
for (M, N, K) in [(8192, 8192, 8192), (65536, 512, 1600), (1600, 512, 65536)]:
    a = torch.randn(M, K, dtype=torch.float64, device=d)
    b = torch.randn(K, N, dtype=torch.float64, device=d)
    for _ in range(2):
        c = a @ b
    torch.cuda.synchronize()
    if M == 1600:
        at = torch.randn(K, M, dtype=torch.float64, device=d)
        c = at.T @ b
        torch.cuda.synchronize()
