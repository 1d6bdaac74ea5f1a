"""Wall time (with synchronisation) of every piece of one E-step iteration of utils.varGP at config-2 size."""
import contextlib
import io
import time

import torch

from gaussian_processes_b200 import synth

with contextlib.redirect_stdout(io.StringIO()):
    from gaussian_processes_b200 import native, utils

N, n, m = 100000, 40, 500
D = torch.device("cuda:0")
prob = synth.make_problem(N, n, m, cfg=2)
x, r, xt = prob.X.to(D), prob.r.to(D), prob.xtilde.to(D)
theta = synth.theta_tensors(prob.theta, device=D)
lims = (dict(synth.THETA_LIMS_LO), dict(synth.THETA_LIMS_HI))


def T(name, fn, reps=3):
    out = fn()
    torch.cuda.synchronize()
    t0 = time.time()
    for _ in range(reps):
        out = fn()
    torch.cuda.synchronize()
    print(f"{name:34s} {(time.time() - t0) / reps * 1e3:9.3f} ms", flush=True)
    return out


C, mask, K_tilde, K, Kvec = T("_kernels_at", lambda: utils._kernels_at(theta, *lims, n, x, xt, utils.acosker, False))
sub = T("_subspace (eigh)", lambda: utils._subspace(K_tilde))
eigvals, eigvecs, ikeep, B, K_tilde_b, K_tilde_inv_b = sub
K_b, a = T("_project", lambda: utils._project(K, B, K_tilde_inv_b, False))
q = B.shape[1]
m_b = torch.zeros(q, dtype=torch.float64, device=D)
V_b = K_tilde_b.clone()
f_params = {"logA": torch.tensor(prob.logA, device=D), "lambda0": torch.tensor(prob.lambda0, device=D)}
xm = T("x[:, mask]", lambda: x[:, mask])
lm, lv = T("lambda_moments", lambda: utils.lambda_moments(xm, K_tilde_b, a, Kvec, K_b, C, m_b, V_b, theta, kernfun=utils.acosker))
T("lambda0_given_logA", lambda: utils.lambda0_given_logA(f_params["logA"], r, lm, lv))
f = T("mean_f_given_lambda_moments", lambda: utils.mean_f_given_lambda_moments(f_params, lm, lv))
T("Estep", lambda: utils.Estep(r=r, KKtilde_inv=a, m=m_b, f_params=f_params, f_mean=f, K_tilde=K_tilde_b,
                               K_tilde_inv=K_tilde_inv_b, update_V_inv=False, alpha=1))
T("  native.estep_terms", lambda: native.estep_terms(a, r, f, float(f_params["logA"])))
W = K_tilde_inv_b + torch.eye(q, dtype=torch.float64, device=D)
T("  native.spd_inverse", lambda: native.spd_inverse(W))
T("  isfinite(V).all() -> bool", lambda: bool(torch.isfinite(W).all()))
T("mean_f(calculate_moments)", lambda: utils.mean_f(f_params=f_params, calculate_moments=True, x=xm, K_tilde=K_tilde_b,
                                                    KKtilde_inv=a, Kvec=Kvec, K=K_b, C=C, m=m_b, V=V_b, theta=theta,
                                                    kernfun=utils.acosker))
T("native.fparam_eval", lambda: native.fparam_eval(lm, lv, r, prob.logA, prob.lambda0, want_f=True))


def lbfgs():
    fp = {"logA": torch.tensor(prob.logA, device=D, requires_grad=True), "lambda0": torch.tensor(prob.lambda0, device=D)}
    opt = torch.optim.LBFGS([fp["logA"]], lr=0.1, max_iter=5, tolerance_change=1e-9, tolerance_grad=1e-7, history_size=5,
                            line_search_fn="strong_wolfe")
    calls = [0]

    def closure():
        calls[0] += 1
        opt.zero_grad()
        res, _ = native.fparam_eval(lm, lv, r, float(fp["logA"]), float(fp["lambda0"]), want_f=True)
        fp["logA"].grad = torch.tensor(-res["dlogA"], device=D)
        return torch.tensor(-res["loglik"], device=D)
    opt.step(closure)
    return calls[0]


print("closure calls per LBFGS(5)", T("LBFGS over logA (5 iters)", lbfgs))
