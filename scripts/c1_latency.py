"""Latency of one ELBO+grad evaluation at notebook scale (config 1: n = 20, N = 2000, m = 100) -- launch-bound regime."""
import sys, os, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gaussian_processes_b200 import native, synth

dev = torch.device("cuda:0")
res = {}
for name, (N, n, m, th) in {"c1_local": (2000, 20, 100, synth.theta_localised(20)),
                            "c1_dense": (2000, 20, 100, synth.THETA_DENSE),
                            "n3000_m300_local": (3000, 20, 300, synth.theta_localised(20))}.items():
    prob = synth.make_problem(N, n, m, cfg=1, theta=th)
    X, r, xt = prob.X.to(dev), prob.r.to(dev), prob.xtilde.to(dev)
    C, mask = native.localker(th, n, dev)
    um = xt[:, mask].contiguous()
    Kt = native.acosker(th["sigma_0"], um, um, C, True)
    ev, evec = torch.linalg.eigh(Kt, UPLO="L")
    keep = ev > max(float(ev.max()) * 1e-4, 1e-4)
    B = evec[:, keep].contiguous()
    m_b, V_b = synth.make_variational_state(torch.diag(ev[keep]).cpu(), 1)
    m_b, V_b = m_b.to(dev), V_b.to(dev)
    for _ in range(5):
        out = native.elbo_grad(X, r, xt, n, th, B, m_b, V_b, prob.logA, prob.lambda0)
    torch.cuda.synchronize()
    native.launch_count(0, reset=True)
    t0 = time.perf_counter()
    reps = 50
    for _ in range(reps):
        out = native.elbo_grad(X, r, xt, n, th, B, m_b, V_b, prob.logA, prob.lambda0)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / reps
    res[name] = {"N": N, "n": n, "m": m, "p": out["p"], "q": int(B.shape[1]), "ms_per_eval": dt * 1e3,
                 "launches_per_eval": native.launch_count(0) / reps}
print(json.dumps(res, indent=1))
