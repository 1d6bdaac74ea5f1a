"""TEST INFRASTRUCTURE: a CPU stand-in for ``gaussian_processes_b200.native`` built on the oracle.

``gaussian_processes_b200.utils`` reaches the CUDA library only through attributes of its ``native`` module, so
``monkeypatch.setattr(utils, "native", fake_native)`` lets the host-side orchestration (varGP's EM loop, LBFGS
closures, tracking, roll-back, the sharded reduction) run on a box without a GPU.  The fused closure is served by
the reverse-mode CPU model ``tests/adjoint_model.py`` -- the same formulation the CUDA kernels implement -- so the
host logic is exercised with gradients of the same provenance.  Never imported by the product package.
"""
import math

import torch

import adjoint_model
import oracle.spatial_gp_oracle as orc

THETA_KEYS = orc.KEYS
DC_KEYS = orc.DC_KEYS
calls = {"elbo_grad": 0, "fparam_eval": 0, "estep_terms": 0, "lambda_moments": 0}


def _th(theta):
    return {k: float(theta[k]) for k in THETA_KEYS}


def mask_host(theta, n_px):
    return orc.build_C(_th(theta), n_px)[1]


def localker(theta, n_px, device, grad=False):
    out = orc.build_C(_th(theta), n_px, grad=grad)
    if not grad:
        return out
    C, mask, dC = out
    return C, mask, torch.stack([dC[k] for k in DC_KEYS])


def acosker(sigma_0, x1, x2, Cm, symmetrize):
    K = orc.arccos_kernel({"sigma_0": sigma_0}, x1, x2, Cm)
    if x1.shape[0] == x2.shape[0] and not symmetrize:   # the oracle symmetrises on n1 == n2 like the reference
        raise AssertionError("fake acosker: n1 == n2 implies symmetrize")
    return K


def acosker_kvec(sigma_0, x1, x2, Cm, symmetrize):
    return acosker(sigma_0, x1, x2, Cm, symmetrize), acosker_diag(sigma_0, x1, Cm)


def acosker_diag(sigma_0, x1, Cm):
    return ((x1 @ Cm) * x1).sum(1) + sigma_0 * sigma_0


def lambda_moments(a, K_b, Kvec, m_b, V_b):
    calls["lambda_moments"] += 1
    return orc.moments(a, K_b, Kvec, m_b, V_b)


def fparam_eval(lam_m, lam_var, r, logA, lambda0, want_f=True):
    calls["fparam_eval"] += 1
    f = orc.rate_mean(logA, lambda0, lam_m, lam_var)
    dlogA, dl0 = orc.expected_loglik_fparam_grad(r, f, lam_m, lam_var, logA)
    res = dict(loglik=orc.expected_loglik(r, f, lam_m, logA, lambda0), dlogA=dlogA, dlambda0=dl0,
               lambda0_closed=orc.lambda0_closed_form(logA, r, lam_m, lam_var), sum_r=float(r.sum()),
               sum_f=float(f.sum()), r_lam=float(r @ lam_m), nan=float(torch.isnan(f).sum()),
               lamf=float((lam_m + math.exp(logA) * lam_var) @ f),
               sum_exp=float(torch.exp(math.exp(logA) * lam_m + 0.5 * math.exp(logA) ** 2 * lam_var).sum()))
    return res, (f if want_f else None)


def rate(lam_m, lam_var, logA, lambda0):
    return orc.rate_mean(logA, lambda0, lam_m, lam_var)


def loglik_sums(r, f, lam_m, lam_var, logA):
    A = math.exp(logA)
    lv = lam_var if lam_var is not None else torch.zeros_like(lam_m)
    return dict(r_lam=float(r @ lam_m), sum_f=float(f.sum()), lamf=float((lam_m + A * lv) @ f),
                sum_r=float(r.sum()), nan=float(torch.isnan(f).sum()))


def estep_terms(a, r, f, logA):
    calls["estep_terms"] += 1
    return orc.estep_terms(a, r, f, logA)


def cholesky(A):
    L, info = torch.linalg.cholesky_ex(A)
    flags = 0
    if int(info) != 0:
        flags |= 1
    elif bool((torch.diag(L) < 1e-10).any()):
        flags |= 2
    return L, flags


def spd_inverse(A):
    L, flags = cholesky(A)
    if flags:
        return torch.full_like(A, float("nan")), float("nan"), flags
    Ainv = torch.cholesky_inverse(L)
    return Ainv, float(2 * torch.log(torch.diag(L)).sum()), 0


def matmul(A, B):
    return A @ B


def matmul_nt(A, Bt):
    return A @ Bt.T


def matmul_tn(At, B):
    return At.T @ B


def elbo_grad(X, r, xtilde, n_px, theta, B, m_b, V_b, logA, lambda0, lims=None, logdetV=None, want_moments=False,
              kl_weight=1.0, same=None):
    calls["elbo_grad"] += 1
    out = adjoint_model.elbo_and_grad(_th(theta), X, r, xtilde, n_px, B, m_b, V_b, logA, lambda0, kl_weight=kl_weight)
    KL = out["KL"]
    if logdetV is not None:   # caller-supplied log det V_b replaces the internally computed one
        L = torch.linalg.cholesky(V_b)
        KL = KL + 0.5 * float(2 * torch.log(torch.diag(L)).sum()) - 0.5 * logdetV
    res = dict(loglik=out["loglik"], KL=KL, grad=dict(out["grad"]), flags=0, p=int(out["inter"]["mask"].sum()),
               sum_r=float(r.sum()), sum_f=float(out["f"].sum()))
    if want_moments:
        res.update(lam_m=out["lam_m"], lam_var=out["lam_var"], f=out["f"])
    return res


def nd_utility(sigma2, mu, r, want_arrays=False):
    r = r.to(torch.float64)
    U = orc.utility(sigma2, mu, r)
    if not want_arrays:
        return U
    lam, keep, p, log_p, _, _ = orc.utility_terms(sigma2, mu, r)
    return U, lam, keep, p, log_p
