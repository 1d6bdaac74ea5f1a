"""Stage split (CUDA events inside the library) of one ELBO+grad evaluation at the three theta operating points of
SURVEY 8(d), config-2 sizes."""
import sys, os, json, math
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gaussian_processes_b200 import native, synth

dev = torch.device("cuda:0")
N, n, m = 100000, 40, 500
prob = synth.make_problem(N, n, m, cfg=2, theta=synth.THETA_DENSE)
X, r, xt = prob.X.to(dev), prob.r.to(dev), prob.xtilde.to(dev)
res = {}
for name, th in (("dense", synth.THETA_DENSE), ("localised", synth.theta_localised(n)), ("notebook", dict(synth.THETA_NOTEBOOK))):
    C, mask = native.localker(th, n, dev)
    um = xt[:, mask].contiguous()
    Kt = native.acosker(th["sigma_0"], um, um, C, True)
    ev, evec = torch.linalg.eigh(Kt, UPLO="L")
    keep = ev > max(float(ev.max()) * 1e-4, 1e-4)
    B = evec[:, keep].contiguous()
    m_b, V_b = synth.make_variational_state(torch.diag(ev[keep]).cpu(), 2)
    m_b, V_b = m_b.to(dev), V_b.to(dev)
    for _ in range(3):
        out = native.elbo_grad(X, r, xt, n, th, B, m_b, V_b, prob.logA, prob.lambda0)
    native.set_option(0, "profile", 1.0)
    native.get_profile(0, reset=True)
    for _ in range(5):
        out = native.elbo_grad(X, r, xt, n, th, B, m_b, V_b, prob.logA, prob.lambda0)
    st, calls = native.get_profile(0, reset=True)
    native.set_option(0, "profile", 0.0)
    res[name] = {"p": out["p"], "q": int(B.shape[1]), "projected": out["projected"],
                 "stage_ms": {k: v / calls for k, v in st.items()}, "total_ms": sum(st.values()) / calls}
print(json.dumps(res, indent=1))
