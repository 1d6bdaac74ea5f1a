"""GPU probe: config-2 evaluation time and error (vs the 8-slice result) for i8_slices_images in {8, 7, 6}."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from gaussian_processes_b200 import native, synth  # noqa: E402

dev = torch.device("cuda:0")
N, n, m = int(os.environ.get("N", 100000)), 40, 500
prob = synth.make_problem(N, n, m, cfg=2, theta=synth.THETA_DENSE)
X, r, xt = prob.X.to(dev), prob.r.to(dev), prob.xtilde.to(dev)
Cm, mask = native.localker(prob.theta, n, dev)
Kt = native.acosker(prob.theta["sigma_0"], xt, xt, Cm, True)
ev, evec = torch.linalg.eigh(Kt, UPLO="L")
keep = ev > max(float(ev.max()) * 1e-4, 1e-4)
B = evec[:, keep].contiguous()
m_b, V_b = synth.make_variational_state(torch.diag(ev[keep]).cpu(), 2)
m_b, V_b = m_b.to(dev), V_b.to(dev)
res = {}
ref = None
for S in (8, 7, 6):
    native.set_option(0, "i8_slices_images", S)
    for _ in range(3):
        out = native.elbo_grad(X, r, xt, n, prob.theta, B, m_b, V_b, prob.logA, prob.lambda0)
    torch.cuda.synchronize()
    native.set_option(0, "profile", 1)
    native.get_profile(0, reset=True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        out = native.elbo_grad(X, r, xt, n, prob.theta, B, m_b, V_b, prob.logA, prob.lambda0)
    e1.record()
    torch.cuda.synchronize()
    st, calls = native.get_profile(0, reset=True)
    native.set_option(0, "profile", 0)
    g = torch.tensor([out["grad"][k] for k in synth.THETA_KEYS], dtype=torch.float64)
    if ref is None:
        ref = (out["loglik"], g)
    res[S] = {"ms": e0.elapsed_time(e1) / 10, "stage_ms": {k: v / calls for k, v in st.items()},
              "loglik_rel_vs8": abs(out["loglik"] - ref[0]) / abs(ref[0]),
              "grad_rel_vs8": float(((g - ref[1]).abs() / ref[1].abs()).max())}
native.set_option(0, "i8_slices_images", -1)
print(json.dumps(res))
