"""GPU probe: slice counts of the two gradient-only products (D = X^T R, Sigma2 = K^T diag(v) K): per-component
gradient error against the REAL reference's fixtures, and config-2 evaluation time, for each (xtr, sig2) setting."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402

from gaussian_processes_b200 import native, synth  # noqa: E402
from helpers import case_inputs, load  # noqa: E402

dev = torch.device("cuda:0")
settings = [(-1, -1), (6, -1), (-1, 7), (6, 7), (6, 6), (5, 7)]
cases = [("closure_c2_dense_n40", 4096), ("closure_c2_local_n40", 4096), ("closure_c3_m1000_n40", 2048),
         ("closure_rankdef_n10", 128), ("closure_c1_dense", 512), ("closure_c1_local", 512)]
res = {}
for xtr, sig2 in settings:
    native.set_option(0, "i8_slices_xtr", xtr)
    native.set_option(0, "i8_slices_sig2", sig2)
    worst = {}
    for name, slab in cases:
        fx = load(name)
        prob, st = case_inputs(fx)
        native.set_slab_rows(0, slab)
        native.set_option(0, "path", 2)
        if slab < 4096:
            native.set_option(0, "i8_slices", 8)
            native.set_option(0, "i8_slices_images", 7)
        out = native.elbo_grad(prob.X.to(dev), prob.r.to(dev), prob.xtilde.to(dev), prob.n_px_side, prob.theta,
                               st["B"].to(dev), st["m_b"].to(dev), st["V_b"].to(dev), st["logA"], st["lambda0"])
        native.set_option(0, "i8_slices", -1)
        native.set_option(0, "i8_slices_images", -1)
        native.set_option(0, "path", 0)
        native.set_slab_rows(0, 0)
        keys = [str(k) for k in fx["theta_keys"]]
        worst[name] = max(abs(out["grad"][k] - (fx["dloglik"][i] - fx["dKL"][i])) / max(abs(fx["dloglik"][i]), abs(fx["dKL"][i]))
                          for i, k in enumerate(keys))
    res[f"xtr{xtr}_sig{sig2}"] = {"grad_comp_err": worst}
# timing at config 2
N, n, m = 100000, 40, 500
prob = synth.make_problem(N, n, m, cfg=2, theta=synth.THETA_DENSE)
X, r, xt = prob.X.to(dev), prob.r.to(dev), prob.xtilde.to(dev)
Cm, mask = native.localker(prob.theta, n, dev)
Kt = native.acosker(prob.theta["sigma_0"], xt, xt, Cm, True)
ev, evec = torch.linalg.eigh(Kt, UPLO="L")
keep = ev > max(float(ev.max()) * 1e-4, 1e-4)
B = evec[:, keep].contiguous()
m_b, V_b = synth.make_variational_state(torch.diag(ev[keep]).cpu(), 2)
m_b, V_b = m_b.to(dev), V_b.to(dev)
for xtr, sig2 in settings:
    native.set_option(0, "i8_slices_xtr", xtr)
    native.set_option(0, "i8_slices_sig2", sig2)
    for _ in range(3):
        native.elbo_grad(X, r, xt, n, prob.theta, B, m_b, V_b, prob.logA, prob.lambda0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        native.elbo_grad(X, r, xt, n, prob.theta, B, m_b, V_b, prob.logA, prob.lambda0)
    e1.record()
    torch.cuda.synchronize()
    res[f"xtr{xtr}_sig{sig2}"]["ms"] = e0.elapsed_time(e1) / 10
native.set_option(0, "i8_slices_xtr", -1)
native.set_option(0, "i8_slices_sig2", -1)
for k, v in res.items():
    print(k, "ms %.3f" % v["ms"], " ".join(f"{n_.replace('closure_', '')}={e:.1e}" for n_, e in v["grad_comp_err"].items()))
