"""Two lambda_moments calls at N=100000, q=500 (for an ncu capture of the int8-slice GEMM with the row-sum epilogue)."""
import torch
from gaussian_processes_b200 import native
d = torch.device("cuda:0")
g = torch.Generator(device=d).manual_seed(1)
N, q = 100000, 500
a = torch.randn(N, q, generator=g, dtype=torch.float64, device=d)
K_b = torch.randn(N, q, generator=g, dtype=torch.float64, device=d)
V = torch.eye(q, dtype=torch.float64, device=d) + 0.01 * torch.randn(q, q, generator=g, dtype=torch.float64, device=d)
m_b = torch.randn(q, generator=g, dtype=torch.float64, device=d)
Kvec = torch.rand(N, generator=g, dtype=torch.float64, device=d) + 600
for _ in range(2):
    native.lambda_moments(a, K_b, Kvec, m_b, V)
    native.matmul(a, V)
torch.cuda.synchronize()
print("done")
