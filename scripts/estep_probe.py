"""GPU probe: estep_terms / lambda_moments on the int8-slice path against torch fp64, and E-step stage timings."""
import sys
import numpy as np
import torch
from gaussian_processes_b200 import native

d = torch.device("cuda:0")


def check(N, q, spread=1.0):
    g = torch.Generator().manual_seed(N + q)
    a = torch.randn(N, q, generator=g, dtype=torch.float64) * torch.exp(spread * torch.randn(N, 1, generator=g, dtype=torch.float64))
    r = torch.poisson(torch.full((N,), 0.8, dtype=torch.float64), generator=g)
    f = torch.exp(0.3 * torch.randn(N, generator=g, dtype=torch.float64))
    gv, G = native.estep_terms(a.to(d), r.to(d), f.to(d), -0.7)
    A = float(np.exp(-0.7))
    G_ref = A * A * (a.T * f) @ a
    bound = A * A * (a.abs().T * f) @ a.abs()
    e = ((G.cpu() - G_ref).abs() / bound)
    print(N, q, spread, "G err/bound max", float(e.max()), "argmax", divmod(int(e.argmax()), q),
          "rel", float((G.cpu() - G_ref).abs().max() / G_ref.abs().max()), flush=True)
    for S in (0,):
        native.set_option(0, "i8_slices", S)
        gv0, G0 = native.estep_terms(a.to(d), r.to(d), f.to(d), -0.7)
        print("   dmma err/bound", float(((G0.cpu() - G_ref).abs() / bound).max()), flush=True)
        native.set_option(0, "i8_slices", -1)


def timing(N=100000, q=500):
    g = torch.Generator(device=d).manual_seed(1)
    a = torch.randn(N, q, generator=g, dtype=torch.float64, device=d)
    K_b = torch.randn(N, q, generator=g, dtype=torch.float64, device=d)
    V = torch.eye(q, dtype=torch.float64, device=d)
    m_b = torch.randn(q, generator=g, dtype=torch.float64, device=d)
    Kvec = torch.rand(N, generator=g, dtype=torch.float64, device=d) + 600
    r = torch.ones(N, dtype=torch.float64, device=d)
    f = torch.rand(N, generator=g, dtype=torch.float64, device=d)

    def t(fn, n=5):
        fn(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n
    for S in (-1, 0):
        native.set_option(0, "i8_slices", S)
        print("i8_slices", S, "estep_terms ms", t(lambda: native.estep_terms(a, r, f, -0.7)),
              "lambda_moments ms", t(lambda: native.lambda_moments(a, K_b, Kvec, m_b, V)),
              "matmul(K,B) ms", t(lambda: native.matmul(a, V)),
              "fparam ms", t(lambda: native.fparam_eval(Kvec, Kvec, r, -3.0, -600.0, want_f=True)), flush=True)
    native.set_option(0, "i8_slices", -1)


if __name__ == "__main__":
    check(6000, 96)
    check(6000, 96, 0.0)
    check(9001, 131)
    check(20000, 500)
    if len(sys.argv) > 1:
        timing()
