"""GPU probe: what the FIRST call of each entry point costs in a fresh process (context, handles, work areas), config 2."""
import os
import sys
import time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
t00 = time.perf_counter()
import torch  # noqa: E402
t_import_torch = time.perf_counter() - t00
t0 = time.perf_counter()
from gaussian_processes_b200 import native, synth  # noqa: E402
t_import_pkg = time.perf_counter() - t0

dev = torch.device("cuda:0")


def timed(name, fn):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    out = fn()
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    out = fn()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print(f"{name:34s} first {1e3 * (t1 - t0):9.1f} ms   second {1e3 * (t2 - t1):8.2f} ms", flush=True)
    return out


print(f"import torch {t_import_torch:.2f} s, import package {t_import_pkg:.2f} s")
t0 = time.perf_counter()
torch.zeros(1, device=dev)
torch.cuda.synchronize()
print(f"CUDA context {time.perf_counter() - t0:.2f} s")
N, n, m = 100000, 40, 500
prob = synth.make_problem(N, n, m, cfg=2, theta=synth.THETA_DENSE)
X, r, xt = prob.X.to(dev), prob.r.to(dev), prob.xtilde.to(dev)
Cm, mask = timed("localker", lambda: native.localker(prob.theta, n, dev))
Kt = timed("acosker m x m", lambda: native.acosker(prob.theta["sigma_0"], xt, xt, Cm, True))
ev, evec = timed("torch.linalg.eigh", lambda: torch.linalg.eigh(Kt, UPLO="L"))
keep = ev > max(float(ev.max()) * 1e-4, 1e-4)
B = evec[:, keep].contiguous()
evk = ev[keep]
m_b, V_b = synth.make_variational_state(torch.diag(evk).cpu(), 2)
m_b, V_b = m_b.to(dev), V_b.to(dev)
K, Kvec = timed("acosker_kvec N x m", lambda: native.acosker_kvec(prob.theta["sigma_0"], X, xt, Cm, False))
K_b = timed("matmul K B", lambda: native.matmul(K, B))
a = K_b / evk[None, :]
lam_m, lam_var = timed("lambda_moments", lambda: native.lambda_moments(a, K_b, Kvec, m_b, V_b))
f = timed("rate", lambda: native.rate(lam_m, lam_var, prob.logA, prob.lambda0))
timed("fparam_eval", lambda: native.fparam_eval(lam_m, lam_var, r, prob.logA, 0.0, want_f=False))
g, G = timed("estep_terms", lambda: native.estep_terms(a, r, f, prob.logA))
W = torch.diag(1.0 / evk) + G
timed("spd_inverse", lambda: native.spd_inverse((W + W.T) * 0.5))
timed("elbo_grad", lambda: native.elbo_grad(X, r, xt, n, prob.theta, B, m_b, V_b, prob.logA, prob.lambda0))
p = torch.nn.Parameter(torch.tensor(0.1, device=dev))
opt = torch.optim.LBFGS([p], lr=0.1, max_iter=3)


def closure():
    opt.zero_grad()
    loss = (p - 1.0) ** 2
    p.grad = (2 * (p - 1.0)).detach()
    return loss


timed("torch.optim.LBFGS step (toy)", lambda: opt.step(closure))
