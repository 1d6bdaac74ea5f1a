"""CPU ORACLE for the Spatial_GP ELBO+gradient hot path.  TEST INFRASTRUCTURE ONLY.

This file is a CPU (PyTorch fp64, no CUDA) *restatement* of the algorithm the reference
implements in ``Spatial_GP_repo/utils.py``.  It exists so that tests, ``__graft_entry__.smoke()``
and ``bench.py``'s CPU-baseline leg have something to check / time the CUDA path against on a
box where ``/root/reference`` does not exist.  Nothing under ``gaussian_processes_b200/`` may
import it: the product path is CUDA only and fails loudly without its extension.

Parity status: PINNED.  ``oracle/gen_golden.py`` (run in the build container, where the Python
reference can be imported) drives the *real* reference functions on seeded synthetic inputs and
stores their outputs in ``tests/golden/*.npz``; ``tests/test_oracle_golden.py`` checks every
function below against those fixtures and against the only known-answer vectors the reference
ships (``moments_gradients.ipynb`` cells 1-3).

The restatement deliberately follows the reference's *forward-mode* gradient algebra (dK per
hyper-parameter pushed through the moments) so that it is an independent check of the CUDA
path, which uses the reverse-mode (adjoint) contraction of the same formulas.

Citations ``U:a-b`` are lines of /root/reference/Spatial_GP_repo/utils.py.
"""
from __future__ import annotations

import math
from typing import Dict, Optional, Tuple

import torch

F64 = torch.float64

# U:25 -- the reference overwrites torch.pi with acos(0)*2 evaluated in float32.
PI32 = 3.1415927410125732
EIGVAL_TOL = 1.0e-4          # U:39
MIN_TOLERANCE = 1.0e-11      # U:37
MASK_THRESHOLD = 1.0e-3      # U:883
COS_EPS = 1.0e-7             # U:984

KEYS = ("sigma_0", "eps_0x", "eps_0y", "-2log2beta", "-log2rho2", "Amp")
DC_KEYS = ("Amp", "-2log2beta", "-log2rho2", "eps_0x", "eps_0y")   # dict order of U:910


def _f(v) -> float:
    return float(v.item()) if torch.is_tensor(v) else float(v)


def check_bounds(theta, lo, hi) -> Optional[str]:
    """U:865-867 / U:2021-2028: first key outside [lo, hi] (None if all inside)."""
    for k in theta:
        if not (lo[k] <= _f(theta[k]) <= hi[k]):
            return k
    return None


# --------------------------------------------------------------------------- a1: localker
def pixel_grid(n: int, device=None) -> Tuple[torch.Tensor, torch.Tensor]:
    """U:876-879: linspace(-1,1,n) grid, 'ij' indexing, flattened row-major: k = iy*n + ix.
    `device`: the reference builds everything on torch's default device (U:41); None = CPU.  bench.py passes the
    GPU to time the reference-shaped computation on cuBLAS / cuSOLVER / ATen (the "library GPU baseline")."""
    lin = torch.linspace(-1, 1, n, dtype=F64, device=device)
    gy, gx = torch.meshgrid(lin, lin, indexing="ij")
    return gx.reshape(-1), gy.reshape(-1)


def build_C(theta, n: int, grad: bool = False, device=None):
    """Pixel-space covariance C(theta), RF mask, and dC for 5 hyper-parameters (U:861-914)."""
    eb = math.exp(_f(theta["-2log2beta"]))
    er = math.exp(_f(theta["-log2rho2"]))
    ex, ey = _f(theta["eps_0x"]), _f(theta["eps_0y"])
    amp = _f(theta["Amp"])
    gx, gy = pixel_grid(n, device)
    log_alpha = -eb * ((gx - ex) ** 2 + (gy - ey) ** 2)          # U:880
    alpha = torch.exp(log_alpha)
    mask = alpha >= MASK_THRESHOLD                                # U:883
    alpha, log_alpha, gx, gy = alpha[mask], log_alpha[mask], gx[mask], gy[mask]
    log_smooth = -er * ((gx[None, :] - gx[:, None]) ** 2 + (gy[None, :] - gy[:, None]) ** 2)  # U:890
    C = amp * alpha[:, None] * torch.exp(log_smooth) * alpha[None, :]  # U:895
    C = (C + C.T) / 2                                             # U:898
    if not grad:
        return C, mask
    dC = {
        "Amp": C / amp,                                                       # U:902
        "-2log2beta": C * (log_alpha[:, None] + log_alpha[None, :]),          # U:907
        "-log2rho2": C * log_smooth,                                          # U:909
        "eps_0x": 2.0 * eb * C * (gx[:, None] + gx[None, :] - 2 * ex),        # U:904
        "eps_0y": 2.0 * eb * C * (gy[:, None] + gy[None, :] - 2 * ey),        # U:905
    }
    return C, mask, dC


# --------------------------------------------------------------------------- a2/a3: acosker
def arccos_kernel(theta, xa: torch.Tensor, xb: torch.Tensor, C: torch.Tensor,
                  dC: Optional[Dict[str, torch.Tensor]] = None):
    """Arc-cosine image kernel K(xa, xb) (rows = images) and its forward-mode dK (U:975-1025)."""
    s0 = _f(theta["sigma_0"])
    s02 = s0 * s0
    na = torch.sqrt(((xa @ C) * xa).sum(1) + s02)                 # U:978 (C symmetric)
    nb = torch.sqrt(((xb @ C) * xb).sum(1) + s02)                 # U:979
    nn = na[:, None] * nb[None, :]                                # U:981
    lin = xa @ C @ xb.T + s02                                     # U:982 (left-to-right)
    cosd = torch.clip(lin / (nn + COS_EPS), -1, 1)                # U:984
    delta = torch.arccos(cosd)                                    # U:986
    J = (torch.sqrt(1 - cosd ** 2) + PI32 * cosd - delta * cosd) / PI32   # U:988
    K = nn * J                                                    # U:990
    dK = None
    if dC is not None:
        dK = {}
        dnn = s02 * (nb[None, :] / na[:, None] + na[:, None] / nb[None, :])   # U:996
        dcos = (2 * s02 - cosd * dnn) / nn                                    # U:998 (no +1e-7, no clip)
        dJ = -(delta - PI32) * dcos / PI32                                    # U:1000
        dK["sigma_0"] = (nn * dJ + dnn * J) / s0                              # U:1004
        for key, dCk in dC.items():
            if key == "sigma_0":
                continue
            dna = 0.5 * ((xa @ dCk.T) * xa).sum(1) / na                       # U:1012
            dnb = 0.5 * ((xb @ dCk.T) * xb).sum(1) / nb                       # U:1013
            dnn = dna[:, None] * nb[None, :] + na[:, None] * dnb[None, :]     # U:1015
            dcos = (xa @ (dCk @ xb.T) - cosd * dnn) / nn                      # U:1017
            dJ = -(delta - PI32) * dcos / PI32                                # U:1019
            dK[key] = nn * dJ + dnn * J                                       # U:1021
    if xa.shape[0] == xb.shape[0]:
        K = (K + K.T) / 2                                         # U:1024-1025 (dK never symmetrised)
    return (K, dK) if dC is not None else K


def arccos_kernel_diag(theta, xa: torch.Tensor, C: torch.Tensor,
                       dC: Optional[Dict[str, torch.Tensor]] = None):
    """Kvec_i = xa_i^T C xa_i + sigma_0^2 and dKvec (U:1027-1044)."""
    s0 = _f(theta["sigma_0"])
    Kvec = ((xa @ C.T) * xa).sum(1) + s0 * s0                     # U:1029
    if dC is None:
        return Kvec
    dKvec = {"sigma_0": torch.full_like(Kvec, 2 * s0 * s0 / s0)}  # U:1036
    for key, dCk in dC.items():
        if key != "sigma_0":
            dKvec[key] = ((xa @ dCk.T) * xa).sum(1)               # U:1042
    return Kvec, dKvec


# --------------------------------------------------------------------------- a5: lambda_moments
def moments(a, K_b, Kvec, m_b, V_b, dK_b=None, dKt_b=None, dKvec=None, Kt_inv_b=None):
    """Predictive mean / variance of lambda at every image (+ forward-mode grads) (U:1089-1124)."""
    lam_m = a @ m_b                                               # U:1090
    VaT = V_b @ a.T
    lam_var = Kvec + (-K_b.T * a.T + a.T * VaT).sum(0)            # U:1101
    if dK_b is None:
        return lam_m, lam_var
    dlam_m, dlam_var = {}, {}
    for key in dK_b:
        da = (dK_b[key] - a @ dKt_b[key]) @ Kt_inv_b              # U:1114
        dlam_m[key] = da @ m_b                                    # U:1117
        dlam_var[key] = (dKvec[key] + 2 * (da * VaT.T).sum(1)
                         - (dK_b[key] * a).sum(1) - (K_b * da).sum(1))   # U:1120
    return lam_m, lam_var, dlam_m, dlam_var


# --------------------------------------------------------------------------- a6-a8
def rate_mean(logA: float, lambda0: float, lam_m, lam_var):
    """f = exp(A lam_m + A^2 lam_var / 2 + lambda0) (U:1133-1141)."""
    A = math.exp(logA)
    return torch.exp(A * lam_m + 0.5 * A * A * lam_var + lambda0)


def lambda0_closed_form(logA: float, r, lam_m, lam_var) -> float:
    """U:1215-1229."""
    A = math.exp(logA)
    return float(torch.log(r.sum()) - torch.log(torch.exp(A * lam_m + 0.5 * A * A * lam_var).sum()))


def expected_loglik(r, f, lam_m, logA: float, lambda0: float):
    """U:1240-1243."""
    A = math.exp(logA)
    return float(A * (r @ lam_m) + lambda0 * r.sum() - f.sum())


def expected_loglik_theta_grad(r, f, logA: float, dlam_m, dlam_var):
    """U:1261-1267."""
    A = math.exp(logA)
    return {k: float(A * (r @ dlam_m[k]) - A * (f @ dlam_m[k]) - 0.5 * A * A * (f @ dlam_var[k]))
            for k in dlam_m}


def expected_loglik_fparam_grad(r, f, lam_m, lam_var, logA: float):
    """dloglik/dlogA and dloglik/dlambda0 (U:1248-1259)."""
    A = math.exp(logA)
    dlogA = A * (r @ lam_m - (lam_m + A * lam_var) @ f)
    return float(dlogA), float(r.sum() - f.sum())


# --------------------------------------------------------------------------- a9: KL
def logdet_pd(M: torch.Tensor) -> float:
    """Cholesky log-det with the reference's fallback semantics (U:1271-1304)."""
    try:
        L = torch.linalg.cholesky(M, upper=True)
        d = torch.diag(L)
        if bool((d <= 0).any()) or bool((d < 1e-10).any()):      # safe_log, U:665-673
            raise ValueError("small pivot")
        return float(2 * torch.log(d).sum())
    except Exception:
        if bool(((M - M.T).abs() > MIN_TOLERANCE).any()):        # U:1302-1304
            return 0.0
        ev = torch.linalg.eigvalsh(M)
        keep = ev > max(float(ev.max()) * EIGVAL_TOL, EIGVAL_TOL)  # U:1287
        return float(torch.log(ev[keep]).sum())


def kl_term(m_b, V_b, Kt_b, Kt_inv_b, dKt_b=None):
    """KL(q||p) without the -q/2 constant, and dKL per key (U:1318-1337)."""
    c = V_b @ Kt_inv_b
    b = Kt_inv_b @ m_b
    KL = -0.5 * logdet_pd(V_b) + 0.5 * logdet_pd(Kt_b) + 0.5 * float(m_b @ b) + 0.5 * float(torch.trace(c))
    if dKt_b is None:
        return KL
    dKL = {}
    for key, dKk in dKt_b.items():
        Bk = dKk @ Kt_inv_b
        dKL[key] = float(0.5 * torch.trace(Bk) - 0.5 * torch.trace(c @ Bk) - 0.5 * (b @ (Bk @ m_b)))
    return KL, dKL


# --------------------------------------------------------------------------- a10: E-step
def estep_terms(a, r, f, logA: float):
    """g = A a^T (r - f), G = A^2 a^T diag(f) a (U:1420-1422)."""
    A = math.exp(logA)
    return A * (a.T @ (r - f)), A * A * (a.T @ (a * f[:, None]))


def estep_update(a, r, f, m_b, Kt_b, logA: float):
    """Newton update of (m_b, V_b), alpha = 1 branch (U:1428-1439)."""
    g, G = estep_terms(a, r, f, logA)
    q = Kt_b.shape[0]
    V_new = torch.linalg.solve(torch.eye(q, dtype=F64, device=Kt_b.device) + Kt_b @ G, Kt_b)
    m_new = V_new @ (G @ m_b + g)
    return m_new, (V_new + V_new.T) / 2


# --------------------------------------------------------------------------- a12: eigen-projection
def eig_projection(Kt: torch.Tensor):
    """eigh + threshold: B, diag K_tilde_b and its inverse (U:1682-1691)."""
    ev, evec = torch.linalg.eigh(Kt, UPLO="L")
    keep = ev > max(float(ev.max()) * EIGVAL_TOL, EIGVAL_TOL)
    return evec[:, keep], torch.diag(ev[keep]), torch.diag(1.0 / ev[keep])


# --------------------------------------------------------------------------- a11: the M-step closure body
def elbo_and_grad(theta, X, r, xtilde, n_px: int, B, m_b, V_b, logA: float, lambda0: float,
                  return_moments: bool = False):
    """Body of ``closure_hyperparams`` (U:2030-2099) for in-bounds theta.

    Returns dict with loglik, KL, elbo (= loglik - KL), grad[key] = d(loglik - KL)/dtheta_key
    (the closure hands -elbo / -grad to LBFGS), and optionally lam_m / lam_var / f.
    """
    C, mask, dC = build_C(theta, n_px, grad=True, device=X.device)                  # U:2030
    xm, um = X[:, mask], xtilde[:, mask]
    same = X.shape[0] == xtilde.shape[0]                                            # ntilde == nt, U:2032
    Kt, dKt = arccos_kernel(theta, um, um, C, dC)                                   # U:2031
    K, dK = (Kt, dKt) if same else arccos_kernel(theta, xm, um, C, dC)              # U:2032
    Kvec, dKvec = arccos_kernel_diag(theta, xm, C, dC)                              # U:2033
    Kt_b = B.T @ Kt @ B
    Kt_b = (Kt_b + Kt_b.T) * 0.5                                                    # U:2047-2048
    K_b = K @ B                                                                     # U:2049
    dKt_b = {k: B.T @ dKt[k] @ B for k in dKt}                                      # U:2061
    dK_b = {k: dK[k] @ B for k in dK}                                               # U:2062
    Kt_inv_b = torch.linalg.solve(Kt_b, torch.eye(Kt_b.shape[0], dtype=F64, device=Kt_b.device))   # U:2067
    a = B if same else K_b @ Kt_inv_b                                               # U:2068
    lam_m, lam_var, dlam_m, dlam_var = moments(a, K_b, Kvec, m_b, V_b, dK_b, dKt_b, dKvec, Kt_inv_b)
    f = rate_mean(logA, lambda0, lam_m, lam_var)                                    # U:2070 (lambda0 not re-estimated)
    loglik = expected_loglik(r, f, lam_m, logA, lambda0)                            # U:2085
    dll = expected_loglik_theta_grad(r, f, logA, dlam_m, dlam_var)
    KL, dKL = kl_term(m_b, V_b, Kt_b, Kt_inv_b, dKt_b)                              # U:2086
    out = {"loglik": loglik, "KL": KL, "elbo": loglik - KL, "dloglik": dll, "dKL": dKL,
           "grad": {k: dll[k] - dKL[k] for k in KEYS}, "p": int(mask.sum())}        # U:2095-2099
    if return_moments:
        out.update(lam_m=lam_m, lam_var=lam_var, f=f)
    return out


def elbo_and_grad_chunked(theta, X, r, xtilde, n_px, B, m_b, V_b, logA, lambda0, chunk: int = 20000):
    """the oracle on row-chunks of the images.  loglik and dloglik are plain sums
    over images; KL and dKL do not depend on the images and are taken from the first chunk.
    Exact up to summation order."""
    tot = None
    for s in range(0, X.shape[0], chunk):
        o = elbo_and_grad(theta, X[s:s + chunk], r[s:s + chunk], xtilde, n_px, B, m_b, V_b, logA, lambda0)
        if tot is None:
            tot = o
        else:
            tot["loglik"] += o["loglik"]
            for k in KEYS:
                tot["dloglik"][k] += o["dloglik"][k]
    tot["elbo"] = tot["loglik"] - tot["KL"]
    tot["grad"] = {k: tot["dloglik"][k] - tot["dKL"][k] for k in KEYS}
    return tot


# --------------------------------------------------------------------------- E-step iteration + logA closure
def estep_iteration(a, K_b, Kvec, r, m_b, V_b, Kt_b, logA: float, lambda0: float, lam_m, lam_var):
    """One pass of U:1877-1892: f from current moments, Newton update, new moments, new lambda0."""
    f = rate_mean(logA, lambda0, lam_m, lam_var)
    m_new, V_new = estep_update(a, r, f, m_b, Kt_b, logA)
    lam_m2, lam_var2 = moments(a, K_b, Kvec, m_new, V_new)
    lam0 = lambda0_closed_form(logA, r, lam_m2, lam_var2)
    return m_new, V_new, lam_m2, lam_var2, lam0


def fparam_closure(logA: float, lambda0: float, r, lam_m, lam_var):
    """closure_f_params (U:1902-1930): returns (-loglik, -dloglik/dlogA, lambda0_new)."""
    f = rate_mean(logA, lambda0, lam_m, lam_var)
    ll = expected_loglik(r, f, lam_m, logA, lambda0)
    dlogA, _ = expected_loglik_fparam_grad(r, f, lam_m, lam_var, logA)
    return -ll, -dlogA, lambda0_closed_form(logA, r, lam_m, lam_var)


# --------------------------------------------------------------------------- active-learning utility
def utility_terms(sigma2, mu, r):
    """nd_lambda_r_mean + nd_p_r_given_xD (U:438-494): returns (lam, keep, p, log_p, r2d, log_r_fact), each (nr, nstar).
    Entries whose z = exp(r sigma2 + mu) sigma2 overflows are 'masked' the reference's way: z, r sigma2, r and log r!
    are zeroed and the entry stays in every sum.  Lambert W: scipy, principal branch, tol 1e-8 (U:465)."""
    import scipy.special
    sigma2, mu, r = sigma2.to(F64), mu.to(F64), r.to(F64)
    rs = torch.outer(r, sigma2)
    z = torch.exp(rs + mu) * sigma2.unsqueeze(0)
    keep = z != math.inf
    z = torch.where(keep, z, torch.zeros((), dtype=F64))
    rs = torch.where(keep, rs, torch.zeros((), dtype=F64))
    w = torch.from_numpy(scipy.special.lambertw(z.numpy(), k=0, tol=1.0e-8).real)
    lam = rs + mu - w
    el = torch.exp(lam)
    lrf = torch.lgamma(r + 1).unsqueeze(1).repeat(1, sigma2.shape[0])
    r2d = r.unsqueeze(1).repeat(1, sigma2.shape[0])
    lrf = torch.where(keep, lrf, torch.zeros((), dtype=F64))
    r2d = torch.where(keep, r2d, torch.zeros((), dtype=F64))
    log_p = lam * r2d - el - (lam - mu) ** 2 / (2 * sigma2.unsqueeze(0)) - 0.5 * torch.log(el * sigma2 + 1) - lrf
    return lam, keep, torch.exp(log_p), log_p, r2d, lrf


def utility(sigma2, mu, r):
    """nd_utility (U:497-519) with nd_mean_noise_entropy (U:416-436): U[nstar] = H(r|x,D) - <H(r|f,x)>."""
    _, _, p, log_p, _, lrf = utility_terms(sigma2, mu, r)
    h_r = -torch.sum(p * log_p, dim=0)
    h_mean = -torch.exp(mu + 0.5 * sigma2) * (mu + sigma2 - 1) + torch.sum(p * lrf, dim=0)
    return h_r - h_mean
