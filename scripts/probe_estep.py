"""GPU probe: the pieces of one E-step iteration (U:1877-1892) at config 2, each timed with CUDA events."""
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from gaussian_processes_b200 import native, synth  # noqa: E402

dev = torch.device("cuda:0")
N, n, m = 100000, 40, 500
prob = synth.make_problem(N, n, m, cfg=2, theta=synth.THETA_DENSE)
X, r, xt = prob.X.to(dev), prob.r.to(dev), prob.xtilde.to(dev)
Cm, mask = native.localker(prob.theta, n, dev)
Kt = native.acosker(prob.theta["sigma_0"], xt, xt, Cm, True)
ev, evec = torch.linalg.eigh(Kt, UPLO="L")
keep = ev > max(float(ev.max()) * 1e-4, 1e-4)
B = evec[:, keep].contiguous()
evk = ev[keep]
m_b, V_b = synth.make_variational_state(torch.diag(evk).cpu(), 2)
m_b, V_b = m_b.to(dev), V_b.to(dev)
K, Kvec = native.acosker_kvec(prob.theta["sigma_0"], X, xt, Cm, False)
K_b = native.matmul(K, B)
a = K_b / evk[None, :]
del K
Kt_inv = torch.diag(1.0 / evk)
lam_m, lam_var = native.lambda_moments(a, K_b, Kvec, m_b, V_b)
logA, lambda0 = prob.logA, prob.lambda0


def timed(fn, reps=10, warm=3):
    for _ in range(warm):
        out = fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        out = fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, out


t_rate, f = timed(lambda: native.rate(lam_m, lam_var, logA, lambda0))
t_terms, (g, G) = timed(lambda: native.estep_terms(a, r, f, logA))
W = Kt_inv + G
t_inv, (Vn, _, _) = timed(lambda: native.spd_inverse((W + W.T) * 0.5))
t_small, mn = timed(lambda: Vn @ (G @ m_b + g))
t_mom, _ = timed(lambda: native.lambda_moments(a, K_b, Kvec, mn, Vn))
t_fp, _ = timed(lambda: native.fparam_eval(lam_m, lam_var, r, logA, 0.0, want_f=False))
t_solve, _ = timed(lambda: torch.linalg.solve(torch.eye(W.shape[0], dtype=W.dtype, device=dev) + torch.diag(evk) @ G, torch.diag(evk)))
t_chol, _ = timed(lambda: torch.cholesky_inverse(torch.linalg.cholesky((W + W.T) * 0.5)))
print("rate %.3f | estep_terms (g, G) %.3f | spd_inverse %.3f | V(Gm+g) %.3f | lambda_moments %.3f | fparam %.3f  [ms]"
      % (t_rate, t_terms, t_inv, t_small, t_mom, t_fp))
print("library for comparison: torch.linalg.solve(I+KG, K) %.3f | torch cholesky + cholesky_inverse %.3f [ms]" % (t_solve, t_chol))
