"""GPU probe: a few evaluations at the reference's real data shape (for an ncu launch list)."""
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from gaussian_processes_b200 import native, synth  # noqa: E402

dev = torch.device("cuda:0")
n, N, m = 108, int(os.environ.get("N", 3160)), int(os.environ.get("M", 200))
pr = synth.make_problem(N, n, m, cfg=108, theta=synth.THETA_NOTEBOOK)
X, r, xt = pr.X.to(dev), pr.r.to(dev), pr.xtilde.to(dev)
C, mk = native.localker(pr.theta, n, dev)
u = xt[:, mk].contiguous()
ev, evec = torch.linalg.eigh(native.acosker(pr.theta["sigma_0"], u, u, C, True), UPLO="L")
kp = ev > max(float(ev.max()) * 1e-4, 1e-4)
B = evec[:, kp].contiguous()
mb, Vb = synth.make_variational_state(torch.diag(ev[kp]).cpu(), 108)
mb, Vb = mb.to(dev), Vb.to(dev)
for _ in range(int(os.environ.get("REPS", 4))):
    o = native.elbo_grad(X, r, xt, n, pr.theta, B, mb, Vb, pr.logA, pr.lambda0, logdetV=0.0)
torch.cuda.synchronize()
print(o["loglik"], o["p"], B.shape[1])
