"""Where the host-streamed evaluation spends its time: stage sums of sgp_elbo_grad_host at config 2."""
import sys, time, json
import torch
from gaussian_processes_b200 import native, synth

dev = torch.device("cuda:0")
N, n, m = 100000, 40, 500
prob = synth.make_problem(N, n, m, cfg=2)
Xh, rh = prob.X.pin_memory(), prob.r.pin_memory()
xt = prob.xtilde.to(dev)
Cm, mask = native.localker(prob.theta, n, dev)
Kt = native.acosker(prob.theta["sigma_0"], xt, xt, Cm, True)
ev, evec = torch.linalg.eigh(Kt, UPLO="L")
keep = ev > max(float(ev.max()) * 1e-4, 1e-4)
B = evec[:, keep].contiguous()
m_b, V_b = synth.make_variational_state(torch.diag(ev[keep]).cpu(), 2)
Bh, mh, Vh, xth = B.cpu().pin_memory(), m_b.cpu().pin_memory(), V_b.cpu().pin_memory(), prob.xtilde.pin_memory()
d = torch.empty_like(Xh, device=dev)
for _ in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); d.copy_(Xh, non_blocking=True); e1.record(); torch.cuda.synchronize()
    print(f"plain copy of X: {Xh.numel() * 8 / e0.elapsed_time(e1) / 1e6:.1f} GB/s", flush=True)
del d
import os
for k, v in [("overlap", float(os.environ.get("OVERLAP", "1"))), ("i8_slices", float(os.environ.get("I8", "-1")))]:
    native.set_option(0, k, v)
for slab in [int(a) for a in sys.argv[1:]] or [0]:
    native.set_slab_rows(0, slab)
    for _ in range(2):
        native.elbo_grad_host(Xh, rh, xth, n, prob.theta, Bh, mh, Vh, prob.logA, prob.lambda0, device=0)
    native.set_option(0, "profile", 1.0)
    native.get_profile(0, reset=True)
    t0 = time.perf_counter()
    for _ in range(3):
        native.elbo_grad_host(Xh, rh, xth, n, prob.theta, Bh, mh, Vh, prob.logA, prob.lambda0, device=0)
    dt = (time.perf_counter() - t0) / 3
    ms, calls = native.get_profile(0, reset=True)
    native.set_option(0, "profile", 0.0)
    print(f"slab={slab}: {dt * 1e3:.2f} ms per evaluation; stage ms:",
          json.dumps({k: round(v / calls, 2) for k, v in ms.items()}), flush=True)
