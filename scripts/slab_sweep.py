"""ELBO+grad evaluation time against the slab size of the image pass (config-2 sizes, dense theta)."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gaussian_processes_b200 import native, synth

dev = torch.device("cuda:0")
N, n, m = 100000, 40, 500
th = synth.THETA_DENSE
prob = synth.make_problem(N, n, m, cfg=2, theta=th)
X, r, xt = prob.X.to(dev), prob.r.to(dev), prob.xtilde.to(dev)
C, mask = native.localker(th, n, dev)
Kt = native.acosker(th["sigma_0"], xt, xt, C, True)
ev, evec = torch.linalg.eigh(Kt, UPLO="L")
keep = ev > max(float(ev.max()) * 1e-4, 1e-4)
B = evec[:, keep].contiguous()
m_b, V_b = synth.make_variational_state(torch.diag(ev[keep]).cpu(), 2)
m_b, V_b = m_b.to(dev), V_b.to(dev)
res = {}
for slab in [int(a) for a in sys.argv[1:]] or [32768, 50048, 65536, 100096]:
    native.set_slab_rows(0, slab)
    for _ in range(3):
        out = native.elbo_grad(X, r, xt, n, th, B, m_b, V_b, prob.logA, prob.lambda0)
    best = 1e9
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            out = native.elbo_grad(X, r, xt, n, th, B, m_b, V_b, prob.logA, prob.lambda0)
        e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / 5)
    res[slab] = {"ms": best, "loglik": out["loglik"]}
native.set_slab_rows(0, 0)
print(json.dumps(res, indent=1))
