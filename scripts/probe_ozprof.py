"""GPU probe (SGP_OZ_DEBUG=64): where CTA 0 of oz_gemm_kernel spends its cycles, for the four product shapes of config 2."""
import os
import sys
os.environ["SGP_OZ_DEBUG"] = os.environ.get("SGP_OZ_DEBUG", "64")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from gaussian_processes_b200 import native  # noqa: E402

dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(1)
for name, M, N, K, S, tr in (("X(UC)^T", 65536, 500, 1600, 7, False), ("K Mfull^T", 65536, 500, 500, 8, False),
                              ("X^T R", 1600, 500, 65536, 7, True)):
    if tr:
        A = torch.randn(K, M, generator=g, dtype=torch.float64, device=dev)
        B = torch.randn(K, N, generator=g, dtype=torch.float64, device=dev)
    else:
        A = torch.randn(M, K, generator=g, dtype=torch.float64, device=dev)
        B = torch.randn(N, K, generator=g, dtype=torch.float64, device=dev)
    for _ in range(3):
        native.gemm_i8emu(A, B, slices=S, transposed=tr)
    torch.cuda.synchronize()
    p = native.debug_get(dev, "oz_prof").cpu().flatten().tolist()
    nt = max(p[6], 1)
    clk = 1.92e3   # cycles per microsecond
    print(f"{name:10s} S={S} tiles(CTA0)={int(nt)}  per tile [us]: MMA loop {p[2]/nt/clk:6.2f} | wait TMEM-free {p[0]/nt/clk:5.2f} | wait operands {p[1]/nt/clk:5.2f}"
          f" || epilogue: wait MMAs {p[3]/nt/clk:5.2f} | drain {p[4]/nt/clk:5.2f} | scale+store {p[5]/nt/clk:5.2f}")
