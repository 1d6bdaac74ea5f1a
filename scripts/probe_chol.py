"""GPU probe: device time of sgp_cholesky (no host read-back: flags_host = NULL, calls queue back to back), the one-launch
cluster kernel (chol_fused.cu) against the launch-per-step path; and sgp_spd_inverse as callers see it (with its sync)."""
import ctypes as C
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from gaussian_processes_b200 import _lib, native  # noqa: E402

dev = torch.device("cuda:0")
lib, h = _lib.lib, _lib.handle_for(0)


def timed(fn, reps=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


for n in (64, 100, 200, 300, 500, 750, 1000):
    g = torch.Generator(device=dev).manual_seed(n)
    W = torch.randn(n, 2 * n, generator=g, dtype=torch.float64, device=dev)
    A = W @ W.T / n + torch.eye(n, dtype=torch.float64, device=dev)
    ld = (n + 1) // 2 * 2
    work = torch.zeros(n, ld, dtype=torch.float64, device=dev)
    st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)

    def dev_only():
        work[:, :n].copy_(A)
        lib.sgp_cholesky(h.ptr, C.c_void_p(work.data_ptr()), ld, n, None, st)
    t_copy = timed(lambda: work[:, :n].copy_(A))
    row = []
    for fused in (1024, 0):
        native.set_option(0, "chol_fused", fused)
        row.append((timed(dev_only) - t_copy, timed(lambda: native.spd_inverse(A))))
    native.set_option(0, "chol_fused", 320)
    t_lib = timed(lambda: torch.linalg.cholesky(A))
    print(f"n = {n:5d}: cholesky on the device {1e3 * row[0][0]:7.1f} us (fused) vs {1e3 * row[1][0]:7.1f} us (stepwise), "
          f"torch.linalg.cholesky {1e3 * t_lib:7.1f} us | spd_inverse call {row[0][1]:.3f} vs {row[1][1]:.3f} ms")
