"""GPU probe: wall time and launch count of one ELBO+grad evaluation at launch-bound sizes."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from gaussian_processes_b200 import native, synth  # noqa: E402

dev = torch.device("cuda:0")
out = {}
for name, N, n, m, theta in (("config1", 2000, 20, 100, synth.THETA_DENSE), ("config1_local", 2000, 20, 100, synth.theta_localised(20)),
                             ("same_m100", 100, 20, 100, synth.theta_localised(20)), ("same_m300", 300, 20, 300, synth.theta_localised(20)),
                             ("n40_N8192_m200", 8192, 40, 200, synth.THETA_DENSE)):
    prob = synth.make_problem(N, n, m, cfg=5, theta=theta)
    X, r, xt = prob.X.to(dev), prob.r.to(dev), prob.xtilde.to(dev)
    Cm, mask = native.localker(prob.theta, n, dev)
    um = xt if bool(mask.all()) else xt[:, mask].contiguous()
    Kt = native.acosker(prob.theta["sigma_0"], um, um, Cm, True)
    ev, evec = torch.linalg.eigh(Kt, UPLO="L")
    keep = ev > max(float(ev.max()) * 1e-4, 1e-4)
    B = evec[:, keep].contiguous()
    m_b, V_b = synth.make_variational_state(torch.diag(ev[keep]).cpu(), 5)
    m_b, V_b = m_b.to(dev), V_b.to(dev)

    def step():
        return native.elbo_grad(X, r, xt, n, prob.theta, B, m_b, V_b, prob.logA, prob.lambda0, logdetV=0.0)
    for _ in range(5):
        step()
    native.launch_count(0, reset=True)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    reps = 200
    for _ in range(reps):
        o = step()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / reps
    out[name] = {"ms": dt * 1e3, "launches": native.launch_count(0) / reps, "q": int(B.shape[1]), "p": o["p"], "flags": o["flags"]}
print(json.dumps(out))
