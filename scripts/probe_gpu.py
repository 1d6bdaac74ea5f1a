"""One-off GPU probe: fp64 matmul peak (cuBLAS via torch), own DMMA GEMM throughput at the hot-path shapes."""
import json, sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gaussian_processes_b200 import native

d = torch.device("cuda:0")
res = {"gpu": torch.cuda.get_device_name(0)}

def timeit(fn, n=5, warm=2):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) * 1e-3)
    return best

n = 8192
a = torch.randn(n, n, dtype=torch.float64, device=d); b = torch.randn(n, n, dtype=torch.float64, device=d)
t = timeit(lambda: torch.matmul(a, b), n=5)
res["cublas_f64_8192_tflops"] = 2 * n**3 / t / 1e12
out = torch.zeros(n, n, dtype=torch.float64, device=d)
t = timeit(lambda: native.gemm_f64(a, b, 0, 0, out=out), n=5)
res["own_f64_8192_nt_tflops"] = 2 * n**3 / t / 1e12
t = timeit(lambda: native.gemm_f64(a, b, 0, 1, out=out), n=5)
res["own_f64_8192_nn_tflops"] = 2 * n**3 / t / 1e12
t = timeit(lambda: native.gemm_f64(a, b, 1, 1, out=out), n=5)
res["own_f64_8192_tn_tflops"] = 2 * n**3 / t / 1e12
del a, b, out
# hot-path shape G1: (65536 x 1600) x (1600 x 1600)
X = torch.randn(65536, 1600, dtype=torch.float64, device=d); Cm = torch.randn(1600, 1600, dtype=torch.float64, device=d)
Z = torch.zeros(65536, 1600, dtype=torch.float64, device=d)
t = timeit(lambda: native.gemm_f64(X, Cm, 0, 1, out=Z), n=5)
res["own_G1_65536x1600x1600_tflops"] = 2 * 65536 * 1600 * 1600 / t / 1e12
t = timeit(lambda: torch.matmul(X, Cm, out=Z), n=5)
res["cublas_G1_tflops"] = 2 * 65536 * 1600 * 1600 / t / 1e12
print(json.dumps(res, indent=1))
json.dump(res, open("gpurun_out/probe.json", "w"), indent=1)
