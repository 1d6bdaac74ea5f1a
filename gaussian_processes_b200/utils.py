"""`utils`-compatible call surface of the Spatial_GP fit loop, backed by sm_100a CUDA kernels.

Drop-in for ``Spatial_GP_repo/utils.py`` of retinal-information-processing-lab/gaussian_processes for the
names the notebooks use: same function names, positional/keyword signatures, return
arities and dict keys.  The tensor work of the hot path (localker, acosker, lambda_moments, the Poisson
expected log-likelihood, the KL term, the E-step terms and the fused M-step closure) runs in
``libsgp_b200.so`` through ``native``; host-side orchestration (LBFGS, eigh-based subspace selection,
tracking dicts, error roll-back) stays Python/PyTorch exactly as in the reference.

There is no CPU fallback: hot functions raise on CPU tensors, and importing this module without the built
extension raises ImportError.  Citations ``U:a-b`` are lines of the reference utils.py.
"""
from __future__ import annotations

import copy
import math
import os
import pickle
import time
import warnings
from datetime import datetime

import torch

from . import native
from ._lib import (FLAG_KT_NOT_PD, FLAG_KT_SMALL_PIVOT, FLAG_NAN, FLAG_V_NOT_PD, FLAG_V_SMALL_PIVOT, SGP_E_BOUNDS,
                   SgpError)

# ---- import-time behaviour of the reference (U:2, U:25, U:31-42) -------------------------------------------
torch.set_grad_enabled(False)
torch.pi = torch.acos(torch.zeros(1)).item() * 2   # float32-valued pi, 3.1415927410125732 (parity trap 1)
TORCH_DTYPE = torch.float64
torch.set_default_dtype(TORCH_DTYPE)
MIN_TOLERANCE = 1.e-11
EIGVAL_TOL = 1.e-4
DEVICE = torch.device("cuda:0" if torch.cuda.is_available() else "cpu")
print(f'Using device: {DEVICE} (from utils.py)')

THETA_KEYS = native.THETA_KEYS
DC_KEYS = native.DC_KEYS


def _fl(v) -> float:
    return float(v.item()) if torch.is_tensor(v) else float(v)


def _theta_floats(theta) -> dict:
    return {k: _fl(theta[k]) for k in THETA_KEYS}


# ---- numerical guards (U:633-685) -------------------------------------------------------------------------
def is_simmetric(tensor, name='M'):
    difference = (tensor - tensor.T).abs()
    if bool((difference > MIN_TOLERANCE).any()):
        print(f'Matrix {name} is not symmetric, maximum difference is {difference.max()}')
        return False
    return True


def is_posdef(tensor, name='M'):
    if not is_simmetric(tensor, name=name):
        warnings.warn('The matrix is not symmetric, cannot check if it is positive definite')
        return False
    smallest = torch.linalg.eigvalsh(tensor).min()
    if smallest <= 0.:
        warnings.warn(f'Matrix {name} is simmetric but has an eigenvalue smaller than 0 ')
        return False
    if smallest <= MIN_TOLERANCE:
        warnings.warn(f'Matrix {name} is simmetric but has an eigenvalue smaller than MIN_TOLERANCE: {MIN_TOLERANCE}')
        return False
    return True


def safe_log(x):
    if bool((x <= 0).any()):
        raise ValueError("Negative or zero input to log detected")
    if bool((x < 1e-10).any()):
        raise ValueError("Very small input to log detected")
    return torch.log(x)


def safe_acos(x):
    if bool((x > 1 - 1e-6).any()) or bool((x < -1 + 1e-6).any()):
        x = torch.clamp(x, -1 + 1e-6, 1 - 1e-6)
    return torch.acos(x)


# ---- hyper-parameter initialisation and unit conversions (U:705-857, host glue) ---------------------------
def generate_xtilde(ntilde, x):
    perm = torch.randperm(ntilde, device=x.device)
    jitter = torch.finfo(TORCH_DTYPE).eps * 10 * torch.randn(ntilde, x.shape[1], dtype=x.dtype, device=x.device)
    return x[perm, :] + jitter


def logbetaexpr_to_beta(logbetaexpr):
    return torch.exp(-0.5 * torch.as_tensor(logbetaexpr)) * 0.5


def logrhoexpr_to_rho(logrhoexpr):
    return torch.exp(-0.5 * torch.as_tensor(logrhoexpr)) / math.sqrt(2.0)


def fromlogbetasam_to_logbetaexpr(logbetasam):
    return logbetasam - math.log(2.0)


def fromlogrhosam_to_logrhoexpr(logrhosam):
    return logrhosam - math.log(2.0)


def get_sta(x, r, n_px_side):
    nt = r.shape[0]
    img_mean = x.T @ torch.ones_like(r) / nt
    sta = x.T @ r / nt - img_mean
    peak = torch.argmax(sta.reshape(n_px_side, n_px_side).abs())
    return sta, torch.tensor(10), (peak // n_px_side, peak % n_px_side)


def generate_theta(x, r, n_px_side, display_hyper=False, **kwargs):
    """Default theta and its limits (U:755-857).  As in the reference, keyword overrides are only honoured
    when ``display_hyper`` is true (parity trap 11), eps_0 is reset to 0 and the RF width is sqrt(10) px."""
    dev = x.device

    def leaf(v):
        return torch.tensor(float(v), dtype=TORCH_DTYPE, device=dev, requires_grad=True)
    get_sta(x, r, n_px_side)
    beta = (math.sqrt(10.0) / n_px_side) * 2.0
    rho = beta / 2
    theta = {'sigma_0': leaf(1.0), 'eps_0x': leaf(0.0), 'eps_0y': leaf(0.0),
             '-2log2beta': leaf(-2 * math.log(2 * beta)), '-log2rho2': leaf(-math.log(2.0 * rho * rho)),
             'Amp': leaf(1.0)}
    if display_hyper:
        for key, value in kwargs.items():
            if key in theta:
                theta[key] = value
                print(f'updated {key} to {_fl(value):.4f}')
        print(f' Hyperparameters have been SET as  : beta = {beta:.4f}, rho = {rho:.4f}')
        print(' Dict of learnable hyperparameters : ' + ", ".join(f"{k} = {_fl(v):.4f}" for k, v in theta.items()))
    lo = {'sigma_0': 0, 'eps_0x': -1, 'eps_0y': -1, '-2log2beta': -float('inf'), '-log2rho2': -float('inf'),
          'Amp': 0.}
    hi = {'sigma_0': float('inf'), 'eps_0x': 1, 'eps_0y': 1, '-2log2beta': float('inf'),
          '-log2rho2': float('inf'), 'Amp': float('inf')}
    return theta, lo, hi


def print_hyp(theta):
    for key in theta.keys():
        extra = ''
        if key == '-2log2beta':
            extra = f' --> beta: {_fl(logbetaexpr_to_beta(theta[key])):>8.4f}'
        if key == '-log2rho2':
            extra = f' --> rho : {_fl(logrhoexpr_to_rho(theta[key])):>8.4f}'
        print(f' {key:<12}: {_fl(theta[key]):>8.4f}{extra}')


# ---- small host helpers -----------------------------------------------------------------------------------
def _work_device(*cands):
    """Device of the first CUDA tensor among the candidates, else torch's default device."""
    for c in cands:
        vals = c.values() if isinstance(c, dict) else (c,)
        for v in vals:
            if torch.is_tensor(v) and v.is_cuda:
                return v.device
    return torch.get_default_device() if hasattr(torch, "get_default_device") else DEVICE


def _scalar(val, like=None, device=None):
    """0-dim fp64 tensor on the device of `like` (the reference returns 0-dim tensors everywhere)."""
    dev = like.device if torch.is_tensor(like) else (device if device is not None else _work_device())
    return torch.tensor(float(val), dtype=TORCH_DTYPE, device=dev)


def _link_params(f_params):
    """(logA, lambda0) as floats; `loglambda0` is the alternative parametrisation of U:1136 / U:1238."""
    logA = _fl(f_params['logA'])
    lambda0 = math.exp(_fl(f_params['loglambda0'])) if 'loglambda0' in f_params else _fl(f_params['lambda0'])
    return logA, lambda0


# ---- image sharding hook (gaussian_processes_b200/sharded.py installs a reducer here) ---------------------
_SHARD = None


def _shard_rank() -> int:
    return 0 if _SHARD is None else _SHARD.rank


def _shard_sum(vals):
    """Sum a short list of floats over the image shards (identity when not sharded)."""
    return list(vals) if _SHARD is None or _SHARD.world == 1 else _SHARD.sum_floats(vals)


def _shard_sum_tensor_(t):
    if _SHARD is not None and _SHARD.world > 1:
        _SHARD.sum_tensor_(t)
    return t


def _out_of_bounds(theta, lo, hi):
    return [k for k, v in theta.items() if not (_fl(lo[k]) <= _fl(v) <= _fl(hi[k]))]


# ---- a1: localker (U:861-914) -----------------------------------------------------------------------------
def localker(theta, theta_higher_lims, theta_lower_lims, n_px_side, grad=False):
    """Pixel-space covariance C(theta) on the receptive-field mask; `dC` for the 5 localker hyper-parameters.
    Raises ValueError when theta leaves its limits (U:865-867).  Tensor work: sgp_localker."""
    for key in _out_of_bounds(theta, theta_lower_lims, theta_higher_lims):
        raise ValueError(f"{key} = {_fl(theta[key]):.4f} is not within the limits of {theta_lower_lims[key]} and "
                         f"{theta_higher_lims[key]}")
    dev = _work_device(theta)
    out = native.localker(_theta_floats(theta), int(n_px_side), dev, grad=grad)
    if not grad:
        return out
    C, mask, dC = out
    return C, mask, {k: dC[i] for i, k in enumerate(DC_KEYS)}       # dict order of U:910


# ---- a2/a3: acosker (U:939-1050) --------------------------------------------------------------------------
def acosker(theta, x1, x2=None, C=None, dC=None, diag=False):
    """Arc-cosine image kernel.  x1: [n1, p], x2: [n2, p] (masked images, rows = images).
    Value path: sgp_acosker / sgp_acosker_diag.  With `dC` the forward-mode `dK` dict of the reference is
    returned as well (API parity; the fit loop itself uses the fused reverse-mode closure instead)."""
    sigma_0 = _fl(theta['sigma_0'])
    if C is None:
        C = torch.eye(x1.shape[1], dtype=x1.dtype, device=x1.device)
    if diag:
        K = native.acosker_diag(sigma_0, x1, C)                                     # U:1029
        if dC is None:
            return K
        dK = {'sigma_0': torch.full_like(K, 2 * sigma_0 ** 2 / sigma_0)}           # U:1036
        for key in dC.keys():
            if key != 'sigma_0':
                dK[key] = native.acosker_diag(0.0, x1, dC[key])                     # x_i^T dC x_i, U:1042
        return K, dK
    n1, n2 = x1.shape[0], x2.shape[0]
    if dC is None:
        return native.acosker(sigma_0, x1, x2, C, symmetrize=(n1 == n2))            # U:978-990, U:1024
    s02 = sigma_0 ** 2
    X1 = torch.sqrt(native.acosker_diag(sigma_0, x1, C))                            # U:978
    X2 = torch.sqrt(native.acosker_diag(sigma_0, x2, C))                            # U:979
    X1X2 = torch.outer(X1, X2)
    x1x2 = native.matmul_nt(native.matmul(x1, C), x2) + s02                         # U:982 (left to right)
    cosdelta = torch.clip(x1x2 / (X1X2 + 1e-7), -1, 1)                              # U:984
    delta = torch.arccos(cosdelta)
    J = (torch.sqrt(1 - cosdelta ** 2) + torch.pi * cosdelta - delta * cosdelta) / torch.pi   # U:988
    K = X1X2 * J
    dK = {}
    dX1X2 = s02 * (X2 / X1[:, None] + X1[:, None] / X2)                             # U:996
    dcos = (2 * s02 - cosdelta * dX1X2) / X1X2                                      # U:998
    dJ = -(delta - torch.pi) * dcos / torch.pi                                      # U:1000
    dK['sigma_0'] = (X1X2 * dJ + dX1X2 * J) / sigma_0                               # U:1004
    for key in dC.keys():
        if key == 'sigma_0':
            continue
        dX1 = 0.5 * native.acosker_diag(0.0, x1, dC[key]) / X1                      # U:1012
        dX2 = 0.5 * native.acosker_diag(0.0, x2, dC[key]) / X2                      # U:1013
        dX1X2 = dX1[:, None] * X2 + X1[:, None] * dX2                               # U:1015
        lin = native.matmul_nt(native.matmul(x1, dC[key]), x2)
        dcos = (lin - cosdelta * dX1X2) / X1X2                                      # U:1017
        dJ = -(delta - torch.pi) * dcos / torch.pi                                  # U:1019
        dK[key] = X1X2 * dJ + dX1X2 * J                                             # U:1021
    if n1 == n2:
        K = (K + K.T) / 2                                                           # U:1024-1025
    return K, dK


def acosker_known_norms(theta, x1, Kvec1, x2, C):
    """`acosker(theta, x1, x2, C=C)` when `Kvec1 = acosker(theta, x1, C=C, diag=True)` is at hand (a fitted model keeps it
    as 'Kvec'): the x1 C product, which is all of the cost for a handful of new images x2, is not repeated.  No
    counterpart in the reference (it recomputes; one_cell_active_training.ipynb cell 17) -- same values."""
    return native.acosker_known(_fl(theta['sigma_0']), x1, Kvec1, x2, C)


# ---- a5: lambda_moments (U:1072-1124) ---------------------------------------------------------------------
@torch.no_grad()
def lambda_moments(x, K_tilde, KKtilde_inv, Kvec, K, C, m, V, theta, kernfun=None, dK=None, dK_tilde=None,
                   dK_vec=None, K_tilde_inv=None):
    """lambda_m = a m ; lambda_var = Kvec + rowsum(a o (a V^T) - K o a)  -> sgp_lambda_moments.
    With (dK, dK_tilde, dK_vec, K_tilde_inv) the forward-mode derivative dicts are returned too (API parity)."""
    a = KKtilde_inv
    if Kvec is None:
        Kvec = kernfun(theta, x, x2=None, C=C, dC=None, diag=True)                  # U:1093-1094
    lambda_m, lambda_var = native.lambda_moments(a, K, Kvec, m, V)
    if dK is None or dK_tilde is None or dK_vec is None or K_tilde_inv is None:
        return lambda_m, lambda_var
    aV = native.matmul_nt(a, V)                                                     # (V a^T)^T
    dlambda_m, dlambda_var = {}, {}
    for key in dK.keys():
        da = native.matmul(dK[key] - native.matmul(a, dK_tilde[key]), K_tilde_inv)  # U:1114
        dlambda_m[key] = da @ m                                                     # U:1117
        dlambda_var[key] = (dK_vec[key] + 2 * (da * aV).sum(1) - (dK[key] * a).sum(1) - (K * da).sum(1))  # U:1120
    return lambda_m, lambda_var, dlambda_m, dlambda_var


# ---- a6: link function (U:1126-1213) ----------------------------------------------------------------------
def mean_f_given_lambda_moments(f_params, lambda_m, lambda_var):
    """f = exp(A lambda_m + A^2 lambda_var / 2 + lambda0), no capping (U:1138-1141)."""
    logA, lambda0 = _link_params(f_params)
    return native.rate(lambda_m, lambda_var, logA, lambda0)


def lambda0_given_logA(logA, r, lambda_m, lambda_var):
    """Closed-form lambda0 = log sum r - log sum exp(A lambda_m + A^2 lambda_var / 2)  (U:1215-1229)."""
    res, _ = native.fparam_eval(lambda_m, lambda_var, r, _fl(logA), 0.0, want_f=False)
    sum_r, sum_exp = _shard_sum([res["sum_r"], res["sum_exp"]])
    return _scalar(math.log(sum_r) - math.log(sum_exp), like=r)                                 # U:1227


def mean_f(f_params, calculate_moments, lambda_m=None, lambda_var=None, x=None, K_tilde=None, KKtilde_inv=None,
           Kvec=None, K=None, C=None, m=None, V=None, V_inv=None, theta=None, kernfun=None, dK=None, dK_tilde=None,
           dK_vec=None, K_tilde_inv=None, r=None):
    """Dispatcher of U:1143-1213: (f) / (f, lambda_m, lambda_var) / (f, lambda_m, lambda_var, dlambda_m, dlambda_var)."""
    extra = ()
    computed = False
    if calculate_moments and (lambda_m is None or lambda_var is None):
        computed = True
        if dK is not None and dK_tilde is not None and dK_vec is not None and K_tilde_inv is not None:
            lambda_m, lambda_var, dlm, dlv = lambda_moments(x, K_tilde, KKtilde_inv, Kvec, K, C, m, V, theta,
                                                            kernfun=kernfun, dK=dK, dK_tilde=dK_tilde,
                                                            dK_vec=dK_vec, K_tilde_inv=K_tilde_inv)
            extra = (dlm, dlv)
        else:
            lambda_m, lambda_var = lambda_moments(x, K_tilde, KKtilde_inv, Kvec, K, C, m, V, theta, kernfun=kernfun)
    fp = f_params
    if r is not None:                                                               # "feature 2", U:1184-1186
        fp = {'logA': f_params['logA'], 'lambda0': lambda0_given_logA(f_params['logA'], r, lambda_m, lambda_var)}
    f_mean = mean_f_given_lambda_moments(fp, lambda_m, lambda_var)
    if computed:
        return (f_mean, lambda_m, lambda_var) + extra
    return f_mean


# ---- a7: expected Poisson log-likelihood (U:1231-1269) ----------------------------------------------------
def compute_loglikelihood(r, f_mean, lambda_m, lambda_var, f_params, compute_grad_for_f_params=False,
                          dlambda_m=None, dlambda_var=None):
    """loglik = A r.lambda_m + lambda0 sum r - sum f, with the three return shapes of the reference."""
    logA, lambda0 = _link_params(f_params)
    A = math.exp(logA)
    s = native.loglik_sums(r, f_mean, lambda_m, lambda_var, logA)
    s = dict(zip(s.keys(), _shard_sum(s.values())))
    loglikelihood = _scalar(A * s["r_lam"] + lambda0 * s["sum_r"] - s["sum_f"], like=r)        # U:1243
    if compute_grad_for_f_params:
        d = {'logA': _scalar(A * (s["r_lam"] - s["lamf"]), like=r)}                            # U:1253
        if 'loglambda0' in f_params:
            d['loglambda0'] = _scalar((s["sum_r"] - s["sum_f"]) * lambda0, like=r)             # U:1254
        elif 'lambda0' in f_params:
            d['lambda0'] = _scalar(s["sum_r"] - s["sum_f"], like=r)                            # U:1255
        return loglikelihood, d
    if dlambda_m is not None and dlambda_var is not None:
        d = {}
        for key in dlambda_m.keys():                                                           # U:1266
            d[key] = A * (r @ dlambda_m[key]) - A * (f_mean @ dlambda_m[key]) - 0.5 * A * A * (f_mean @ dlambda_var[key])
            _shard_sum_tensor_(d[key])
        return loglikelihood, d
    return loglikelihood, _scalar(s["r_lam"], like=r), _scalar(s["sum_r"], like=r)             # U:1269


# ---- a9: KL term (U:1271-1337) ----------------------------------------------------------------------------
def log_det(M, name='M', ignore_warning=False):
    """log det through the blocked Cholesky kernel; a failed / tiny pivot takes the reference's eigh
    fallback (safe_log raises inside the try of U:1273-1279)."""
    L, flags = native.cholesky(M.T)          # cholesky(upper=True) reads the upper triangle (U:1275)
    if flags == 0 and bool(torch.isfinite(L).all()):
        return 2 * torch.sum(torch.log(torch.diag(L)))
    if is_simmetric(M, name=name):
        eigenvalues = torch.linalg.eigvalsh(M)
        ikeep = eigenvalues > max(eigenvalues.max() * EIGVAL_TOL, EIGVAL_TOL)                  # U:1287
        if not ignore_warning:
            smallest = eigenvalues.min()
            warnings.warn(f"Matrix {name} in logdet is simmetric but not posdef, using eigendecomposition to "
                          f"calculate the log determinant")
            if smallest <= 0.:
                warnings.warn(f'Matrix {name} in logdet is simmetric but has an eigenvalue smaller than 0 ')
            elif smallest <= 1.e-10:
                warnings.warn(f'Matrix {name} in logdet is simmetric but has an eigenvalue smaller than 1e-10 ')
        return torch.sum(safe_log(eigenvalues[ikeep]))
    warnings.warn(f"Matrix {name} in logdet is not simmetric in log_det used in KL_divergence")
    return 0


def compute_KL_div(m, V, K_tilde, K_tilde_inv, dK_tilde=None, ignore_warning=False):
    """KL(q || p) up to the -q/2 constant the reference omits (U:1326) and dKL per hyper-parameter (U:1333)."""
    c = V @ K_tilde_inv
    b = K_tilde_inv @ m
    KL = (-0.5 * log_det(V, name='V', ignore_warning=ignore_warning) + 0.5 * log_det(K_tilde, name='K_tilde')
          + 0.5 * torch.dot(m, b) + 0.5 * torch.trace(c))
    if dK_tilde is None:
        return KL
    dKL = {}
    for key in dK_tilde.keys():
        Bk = dK_tilde[key] @ K_tilde_inv
        dKL[key] = 0.5 * torch.trace(Bk) - 0.5 * torch.trace(c @ Bk) - 0.5 * torch.dot(b, Bk @ m)
    return KL, dKL


# ---- a10: E-step (U:1402-1459) ----------------------------------------------------------------------------
def Estep(r, KKtilde_inv, m, f_params, f_mean, K_tilde=None, K_tilde_inv=None, V=None, update_V_inv=False, alpha=1):
    """Newton update of (m_b, V_b): g = A a^T (r-f), G = A^2 a^T diag(f) a  (sgp_estep_terms, one pass over the
    images), V_new = (I + K_tilde G)^-1 K_tilde, m_new = V_new (G m + g), V_new symmetrised.
    Only the live branch of the reference (update_V_inv=False, alpha=1, U:1881) is provided."""
    if update_V_inv or alpha != 1 or K_tilde is None:
        raise NotImplementedError('Estep: only update_V_inv=False, alpha=1 is provided (the branch varGP uses)')
    g, G = native.estep_terms(KKtilde_inv, r, f_mean, _fl(f_params['logA']))
    _shard_sum_tensor_(g)
    _shard_sum_tensor_(G)
    V_new = None
    if K_tilde_inv is not None:
        # (I + K G)^-1 K == (K^-1 + G)^-1 : symmetric positive definite -> blocked Cholesky inverse on the device
        W = K_tilde_inv + G
        Vi, _, flags = native.spd_inverse((W + W.T) * 0.5)
        if flags == 0 and bool(torch.isfinite(Vi).all()):
            V_new = Vi
    if V_new is None:
        V_new = torch.linalg.solve(torch.eye(K_tilde.shape[0], dtype=K_tilde.dtype, device=K_tilde.device)
                                   + K_tilde @ G, K_tilde)                                     # U:1430
    m_new = V_new @ (G @ m + g)                                                               # U:1431
    V_new = (V_new + V_new.T) / 2                                                             # U:1438
    return m_new, V_new


# ---- active-learning utility (U:416-637): Lambert W on the device instead of the scipy round trip (U:464-466) ----
def nd_mean_noise_entropy(p_response, log_r2d_fact, sigma2, mu):
    """<H(r|f,x)>_p(f|D) for nstar candidates (U:416-436); p_response / log_r2d_fact are (r, nstar)."""
    p_times_logr_sum = torch.sum(p_response * log_r2d_fact, dim=0)
    return -torch.exp(mu + 0.5 * sigma2) * (mu + sigma2 - 1) + p_times_logr_sum


def nd_lambda_r_mean(r, sigma2, mu):
    """argmax of the Laplace-approximated log p(r|x,D) (U:438-471): (lambW (r, nstar), sum_mask (r, nstar))."""
    _, lam, keep, _, _ = native.nd_utility(sigma2, mu, r, want_arrays=True)
    return lam, keep


def nd_p_r_given_xD(r, sigma2, mu):
    """p(r|x,D) for every count and candidate (U:473-494): (p, log p, masked r (r, nstar), masked log r! (r, nstar))."""
    _, _, keep, p, log_p = native.nd_utility(sigma2, mu, r, want_arrays=True)
    zero = torch.zeros((), dtype=p.dtype, device=p.device)
    r2d = torch.where(keep, r.to(p.dtype).unsqueeze(1).repeat(1, sigma2.shape[0]), zero)
    log_r_fact = torch.where(keep, torch.lgamma(r.to(p.dtype) + 1).unsqueeze(1).repeat(1, sigma2.shape[0]), zero)
    return p, log_p, r2d, log_r_fact


@torch.no_grad()
def nd_utility(sigma2, mu, r_masked):
    """Utility U = H(r|x,D) - <H(r|f,x)> of nstar candidates (U:496-519) in one kernel launch."""
    if sigma2.ndim == 0:
        sigma2 = sigma2[None]
        mu = mu[None]
    return native.nd_utility(sigma2, mu, r_masked)


def mean_noise_entropy(p_response, r, sigma2, mu):
    """Single-candidate <H(r|f,x)> (U:521-538)."""
    p_times_logr_sum = p_response @ torch.lgamma(r + 1)
    return -torch.exp(mu + 0.5 * sigma2) * (mu + sigma2 - 1) + p_times_logr_sum


def lambda_r_mean(r, sigma2, mu):
    """Single-candidate variant (U:540-571): overflowing counts are REMOVED, not zeroed; returns (lamb, sum_mask)."""
    _, lam, keep, _, _ = native.nd_utility(sigma2.reshape(1), mu.reshape(1), r, want_arrays=True)
    keep = keep[:, 0]
    return lam[:, 0][keep], keep


def p_r_given_xD(r, sigma2, mu):
    """Single-candidate p(r|x,D) over the kept counts (U:573-594): (p, log p, r[sum_mask])."""
    _, _, keep, p, log_p = native.nd_utility(sigma2.reshape(1), mu.reshape(1), r, want_arrays=True)
    keep = keep[:, 0]
    return p[:, 0][keep], log_p[:, 0][keep], r[keep]


@torch.no_grad()
def utility(sigma2, mu, r_masked):
    """Single-candidate utility (U:596-618)."""
    p_response, log_p_response, r_masked2 = p_r_given_xD(r=r_masked, sigma2=sigma2, mu=mu)
    H_r_xD = -torch.sum(p_response * log_p_response)
    E_H_r_f = mean_noise_entropy(p_response, r_masked2, sigma2, mu)
    return H_r_xD - E_H_r_f


def get_utility(xstar, xtilde, C, mask, theta, m, V, K_tilde, K_tilde_inv, kernfun):
    """U:620-631, kept call for call -- including the reference's own argument lists (its calls to
    lambda_moments_star / utility omit arguments the current signatures require, so it raises TypeError there too)."""
    if K_tilde_inv is None:
        K_tilde_inv = torch.linalg.solve(K_tilde, torch.eye(K_tilde.shape[0], device=K_tilde.device))
    xstar = xstar.unsqueeze(0)
    mu_star, sigma2_star = lambda_moments_star(xstar[:, mask], xtilde[:, mask], C, theta, K_tilde_inv, m, V,
                                               kernfun=kernfun)
    return utility(sigma2=sigma2_star, mu=mu_star)


# ---- inference (U:1476-1541, U:326-412) -------------------------------------------------------------------
def lambda_moments_star(xstar, xtilde, C, theta, K_tilde, K_tilde_inv, m, V, B, kernfun, _batched=False):
    """Predictive moments of lambda at test images (U:1476-1500): mu = a m, sigma2 = K** + diag(a (V - K_tilde) a^T)
    with a = K(x*, xtilde) B K_tilde^-1.  `_batched` (used by test()): xstar holds many independent test images, which
    the reference pushes through one at a time -- so K(x*, xtilde) is never symmetrised, whatever their number."""
    if kernfun == 'acosker':
        kernfun = acosker
    elif kernfun is not acosker:
        raise Exception('Kernel function not recognized')
    if _batched:
        Kstar = native.acosker(_fl(theta['sigma_0']), xstar, xtilde, C, symmetrize=False)
    else:
        Kstar = kernfun(theta, xstar, xtilde, C=C, dC=None, diag=False)
    Kvec_star = Kstar @ B
    KKtilde_inv = Kvec_star @ K_tilde_inv
    mu_star = KKtilde_inv @ m
    K_star = kernfun(theta, xstar, x2=None, C=C, dC=None, diag=True)
    sigma_star2 = K_star + ((KKtilde_inv @ (V - K_tilde)) * KKtilde_inv).sum(1)               # diag(a (V-K) a^T)
    return mu_star, torch.reshape(sigma_star2, (xstar.shape[0],))


def explained_variance(rtst, f_pred, sigma=True):
    reven = torch.mean(rtst[0::2, :], axis=0)
    rodd = torch.mean(rtst[1::2, :], axis=0)
    reliability = torch.abs(torch.corrcoef(torch.stack((reven, rodd))))[0, 1]
    accuracy_o = torch.corrcoef(torch.stack((f_pred, rodd)))[0, 1]
    accuracy_e = torch.corrcoef(torch.stack((f_pred, reven)))[0, 1]
    r2 = 0.5 * (accuracy_o + accuracy_e) / reliability
    if not sigma:
        return r2, None
    nbootstrap = 1000
    r2 = torch.zeros(nbootstrap, device=f_pred.device)
    n = rtst.shape[0]
    for i in range(nbootstrap):
        ilist = torch.randperm(n, device=rtst.device)
        reven = torch.mean(rtst[ilist[0::2], :], axis=0)
        rodd = torch.mean(rtst[ilist[1::2], :], axis=0)
        reliability = torch.abs(torch.corrcoef(torch.stack((reven, rodd)))[0, 1])
        accuracy_o = torch.corrcoef(torch.stack((f_pred, rodd)))[0, 1]
        accuracy_e = torch.corrcoef(torch.stack((f_pred, reven)))[0, 1]
        r2[i] = 0.5 * (accuracy_o + accuracy_e) / reliability
    return torch.mean(r2), torch.std(r2)


def _subspace(K_tilde):
    """eigh + threshold of U:1682-1691: (eigvals, eigvecs, ikeep, B, K_tilde_b, K_tilde_inv_b)."""
    eigvals, eigvecs = torch.linalg.eigh(K_tilde, UPLO='L')
    ikeep = eigvals > max(eigvals.max() * EIGVAL_TOL, EIGVAL_TOL)
    return (eigvals, eigvecs, ikeep, eigvecs[:, ikeep], torch.diag(eigvals[ikeep]),
            torch.diag_embed(1 / eigvals[ikeep]))


@torch.no_grad()
def test(X_test, R_test, xtilde, X_train=None, at_iteration=None, **kwargs):
    """Predicted rates for the test images and the bootstrap r2 (U:326-412).  All test images are pushed through
    the kernels in one batch instead of the reference's per-image Python loop (same arithmetic per image)."""
    fp = kwargs['fit_parameters']
    maxiter, nEstep, nMstep = fp.get('maxiter', 0), fp.get('nEstep', 0), fp.get('nMstep', 0)
    cellid, n_px_side = fp.get('cellid'), fp.get('n_px_side')
    mask = kwargs.get('mask')
    theta, theta_lower_lims, theta_higher_lims = kwargs.get('hyperparams_tuple')
    C, m, V, B = kwargs.get('C'), kwargs.get('m_b'), kwargs.get('V_b'), kwargs.get('B')
    K_tilde, K_tilde_inv = kwargs.get('K_tilde_b'), kwargs.get('K_tilde_inv_b')
    f_params = kwargs.get('f_params')
    logA = f_params['logA']
    lambda0 = torch.exp(f_params['loglambda0']) if 'loglambda0' in f_params else f_params['lambda0']
    if at_iteration is not None and X_train is not None:
        vt = kwargs['values_track']
        theta = {key: val[at_iteration] for key, val in vt['theta_track'].items()}
        m = vt['variation_par_track']['m_b'][at_iteration]
        V = vt['variation_par_track']['V_b'][at_iteration]
        logA = vt['f_par_track']['logA'][at_iteration]
        lambda0 = (vt['f_par_track']['lambda0'][at_iteration] if 'lambda0' in vt['f_par_track']
                   else torch.exp(vt['f_par_track']['loglambda0'][at_iteration]))
        C, mask = localker(theta=theta, theta_higher_lims=theta_higher_lims, theta_lower_lims=theta_lower_lims,
                           n_px_side=n_px_side, grad=False)
        Kt = acosker(theta, xtilde[:, mask], xtilde[:, mask], C=C, diag=False)
        _, _, _, B, K_tilde, K_tilde_inv = _subspace(Kt)
    A = torch.exp(logA)
    xs = torch.reshape(X_test, (X_test.shape[0], -1))
    mu_star, sigma_star2 = lambda_moments_star(xs[:, mask], xtilde[:, mask], C, theta, K_tilde, K_tilde_inv, m, V, B,
                                               'acosker', _batched=True)
    R_predicted = torch.exp(A * mu_star + 0.5 * A * A * sigma_star2 + lambda0)                # U:395
    r2, sigma_r2 = explained_variance(R_test[:, :, cellid], R_predicted, sigma=True)
    print(f"\n\n Pietro's model: R2 = {r2:.2f} ± {sigma_r2:.2f} Cell: {cellid} maxiter = {maxiter}, "
          f"nEstep = {nEstep}, nMstep = {nMstep} \n")
    return R_test[:, :, cellid], R_predicted, r2, sigma_r2


# ---- a11: the fit loop (U:1568-2316) ----------------------------------------------------------------------
def _kernels_at(theta, lo, hi, n_px_side, x, xtilde, kernfun, same):
    """C, mask, K_tilde, K, Kvec at theta (no gradients): U:1674-1680 / U:1803-1806 / U:2196-2199."""
    C, mask = localker(theta=theta, theta_higher_lims=hi, theta_lower_lims=lo, n_px_side=n_px_side, grad=False)
    xm, um = x[:, mask], xtilde[:, mask]
    K_tilde = kernfun(theta, um, um, C=C, dC=None, diag=False)
    if kernfun is acosker and not same:
        # K and Kvec share the x C product (sgp_acosker_kvec): one pass over the images instead of two
        # not `same`: globally n1 != n2, so the reference does not symmetrise (U:1024) -- a shard that happens to hold
        # ntilde rows must not either
        K, Kvec = native.acosker_kvec(_fl(theta['sigma_0']), xm, um, C, symmetrize=False)
        return C, mask, K_tilde, K, Kvec
    K = K_tilde if same else kernfun(theta, xm, um, C=C, dC=None, diag=False)
    Kvec = kernfun(theta, xm, x2=None, C=C, dC=None, diag=True)
    return C, mask, K_tilde, K, Kvec


def _project(K, B, K_tilde_inv_b, same):
    K_b = native.matmul(K, B)
    return K_b, (B if same else native.matmul(K_b, K_tilde_inv_b))


def _closure_forward_mode(theta, lo, hi, n_px_side, x, xtilde, r, kernfun, same, B, m_b, V_b, f_params):
    """The M-step closure body exactly as the reference sequences it (U:2030-2099), on the function-level entry
    points.  Used when ntilde == nt (a = B, the active-learning configuration) and when the fused call reports a
    non-positive-definite K_tilde_b (the reference then takes log_det's eigh fallback)."""
    C, mask, dC = localker(theta=theta, theta_higher_lims=hi, theta_lower_lims=lo, n_px_side=n_px_side, grad=True)
    xm, um = x[:, mask], xtilde[:, mask]
    K_tilde, dK_tilde = kernfun(theta, um, um, C=C, dC=dC, diag=False)
    K, dK = (K_tilde, dK_tilde) if same else kernfun(theta, xm, um, C=C, dC=dC, diag=False)
    Kvec, dKvec = kernfun(theta, xm, x2=None, C=C, dC=dC, diag=True)
    K_tilde_b = B.T @ K_tilde @ B
    K_tilde_b = (K_tilde_b + K_tilde_b.T) * 0.5                                               # U:2047-2048
    K_b = native.matmul(K, B)
    dK_tilde_b = {key: B.T @ dK_tilde[key] @ B for key in dK_tilde}
    dK_b = {key: native.matmul(dK[key], B) for key in dK}
    K_tilde_inv_b = torch.linalg.solve(K_tilde_b, torch.eye(K_tilde_b.shape[0], dtype=K_tilde_b.dtype,
                                                            device=K_tilde_b.device))         # U:2067
    a = B if same else native.matmul(K_b, K_tilde_inv_b)
    f_mean, lambda_m, lambda_var, dlm, dlv = mean_f(
        f_params=f_params, calculate_moments=True, x=xm, K_tilde=K_tilde_b, KKtilde_inv=a, Kvec=Kvec, K=K_b, C=C,
        m=m_b, V=V_b, theta=theta, kernfun=kernfun, dK=dK_b, dK_tilde=dK_tilde_b, dK_vec=dKvec,
        K_tilde_inv=K_tilde_inv_b)
    loglik, dloglik = compute_loglikelihood(r, f_mean, lambda_m, lambda_var, f_params, dlambda_m=dlm, dlambda_var=dlv)
    KL, dKL = compute_KL_div(m_b, V_b, K_tilde_b, K_tilde_inv=K_tilde_inv_b, dK_tilde=dK_tilde_b)
    return _fl(loglik), _fl(KL), {key: _fl(dloglik[key]) - _fl(dKL[key]) for key in theta.keys()}


def _sharded() -> bool:
    return _SHARD is not None and _SHARD.world > 1


def elbo_and_grad(theta, lo, hi, n_px_side, x, xtilde, r, B, m_b, V_b, f_params, logdetV=None, kernfun=None,
                  same=None):
    """One ELBO+gradient evaluation (the unit of BASELINE.json's metric): (loglik, KL, {key: d(loglik-KL)/dtheta}).
    ntilde != nt: ONE fused C call (sgp_elbo_grad, reverse mode).  Raises ValueError out of bounds.
    `same` (ntilde == nt, the reference's a = B branch U:2032/U:2068) is decided from the GLOBAL image count: an
    image shard that happens to hold exactly ntilde rows is not the active-learning configuration."""
    for key in _out_of_bounds(theta, lo, hi):
        raise ValueError(f"{key} = {_fl(theta[key]):.4f} is not within the limits of {lo[key]} and {hi[key]}")
    if same is None:
        same = (not _sharded()) and x.shape[0] == xtilde.shape[0]
    kernfun = kernfun or acosker
    if same and x.is_cuda and kernfun is acosker:
        # ntilde == nt (K = K_tilde, a = B: the active-learning configuration): the fused call's own routine for it
        logA, lambda0 = _link_params(f_params)
        out = native.elbo_grad(x, r, xtilde, int(n_px_side), _theta_floats(theta), B, m_b, V_b, logA, lambda0,
                               logdetV=logdetV, same=True)
        if not out["flags"] & (FLAG_KT_NOT_PD | FLAG_KT_SMALL_PIVOT):
            return out["loglik"], out["KL"], out["grad"]
    if not same:
        logA, lambda0 = _link_params(f_params)
        if _sharded() and getattr(_SHARD, "peer_exchange", False):
            # image shards with the exchange fused into the evaluation's last kernel (peer memory): already reduced
            out = native.elbo_grad(x, r, xtilde, int(n_px_side), _theta_floats(theta), B, m_b, V_b, logA, lambda0,
                                   logdetV=logdetV, kl_weight=1.0 if _shard_rank() == 0 else 0.0, same=False)
            if out["flags"] & 32:
                raise RuntimeError("image-sharded evaluation: a peer's partial result did not arrive (exchange timeout)")
            if not out["flags"] & (FLAG_KT_NOT_PD | FLAG_KT_SMALL_PIVOT):
                return out["loglik"], out["KL"], out["grad"]
            return _closure_forward_mode(theta, lo, hi, n_px_side, x, xtilde, r, kernfun, same, B, m_b, V_b, f_params)
        if _sharded() and x.is_cuda and hasattr(native, "elbo_grad_dev"):
            # image shards: asynchronous fused call, device-side all-reduce of the 9 leading result doubles, ONE read-back
            dev_out = native.elbo_grad_dev(x, r, xtilde, int(n_px_side), _theta_floats(theta), B, m_b, V_b, logA, lambda0,
                                           logdetV=logdetV, kl_weight=1.0 if _shard_rank() == 0 else 0.0)
            v = _SHARD.reduce_result_(dev_out)
            if v[8] == 0:          # count of ranks whose K_tilde_b Cholesky failed (it is replicated: all or none)
                return v[0], v[1], dict(zip(THETA_KEYS, v[2:8]))
            return _closure_forward_mode(theta, lo, hi, n_px_side, x, xtilde, r, kernfun, same, B, m_b, V_b, f_params)
        out = native.elbo_grad(x, r, xtilde, int(n_px_side), _theta_floats(theta), B, m_b, V_b, logA, lambda0,
                               logdetV=logdetV, kl_weight=1.0 if _shard_rank() == 0 else 0.0,
                               same=False if _sharded() else None)
        bad = float(bool(out["flags"] & (FLAG_KT_NOT_PD | FLAG_KT_SMALL_PIVOT)))
        vec = _shard_sum([out["loglik"]] + [out["grad"][k] for k in THETA_KEYS] + [bad])   # the one exchange step
        if vec[7] == 0:
            return vec[0], out["KL"], dict(zip(THETA_KEYS, vec[1:7]))
    return _closure_forward_mode(theta, lo, hi, n_px_side, x, xtilde, r, kernfun, same, B, m_b, V_b, f_params)


@torch.no_grad()
def varGP(x, r, **kwargs):
    """EM fit of the variational GP (U:1568-2316): same kwargs, same `(fit_model, err_dict)` result layout,
    same roll-back behaviour on errors.  Host orchestration (LBFGS, eigh-based subspace, tracking) is Python as
    in the reference; every N-sized tensor operation runs in libsgp_b200.so, the M-step closure as one fused call."""
    start_time_before_init = time.time()
    err_dict = {'is_error': False, 'error_message': None}
    nt, nx = x.shape

    fit_parameters = copy.deepcopy(kwargs['fit_parameters'])
    fit_parameters['min_tolerance'] = MIN_TOLERANCE
    fit_parameters['eigval_tol'] = EIGVAL_TOL
    ntilde = fit_parameters.get('ntilde', 100 if nt > 100 else nt)
    maxiter = fit_parameters.get('maxiter', 50)
    nEstep = fit_parameters.get('nEstep', 50)
    nMstep = fit_parameters.get('nMstep', 20)
    nFparamstep = fit_parameters.get('nFparamstep', 10)
    display_hyper = fit_parameters.get('display_hyper', True)
    n_px_side = int(round(fit_parameters.get('n_px_side', math.sqrt(nx))))
    kernfun = fit_parameters.get('kernfun', 'acosker')
    if kernfun == 'acosker':
        kernfun = acosker
    else:
        raise Exception('Kernel function not recognized')

    if _SHARD is not None and 'xtilde' not in kwargs:
        raise Exception('image-sharded varGP needs xtilde passed explicitly (it must be identical on every rank)')
    xtilde = kwargs['xtilde'] if 'xtilde' in kwargs else generate_xtilde(ntilde, x)
    if ntilde != xtilde.shape[0]:
        raise Exception('Number of inducing points does not match ntilde')
    hyperparams_tuple = (copy.deepcopy(kwargs['hyperparams_tuple']) if 'hyperparams_tuple' in kwargs
                         else generate_theta(x, r, n_px_side, display_hyper))
    theta = copy.deepcopy(kwargs.get('theta', hyperparams_tuple[0]))
    theta_lower_lims = copy.deepcopy(kwargs.get('theta_lower_lims', hyperparams_tuple[1]))
    theta_higher_lims = copy.deepcopy(kwargs.get('theta_higher_lims', hyperparams_tuple[2]))
    if 'f_params' not in kwargs:
        raise Exception('f_params not provided')
    f_params = copy.deepcopy(kwargs['f_params'])
    for key in f_params.keys():
        f_params[key] = f_params[key].requires_grad_(True)
    # ntilde == nt (a = B, U:1677-1678 / U:2032) is a statement about the whole image set, not about one shard
    nt_global = int(round(_shard_sum([float(nt)])[0])) if _sharded() else nt
    same = ntilde == nt_global
    if same and _sharded():
        raise Exception('image-sharded varGP does not support ntilde == nt (K = K_tilde needs every image on every rank)')
    lims = (theta_lower_lims, theta_higher_lims)
    if x.is_cuda:   # a new fit may bring new images at a recycled address: never reuse cached copies / slices of them
        native.new_image_epoch(x.device.index if x.device.index is not None else torch.cuda.current_device())

    # ---- initial kernels and subspace (U:1674-1694); every piece can be handed over via init_kernel ----------
    eigvecs = None
    if 'init_kernel' not in kwargs:
        C, mask, K_tilde, K, Kvec = _kernels_at(theta, *lims, n_px_side, x, xtilde, kernfun, same)
        eigvals, eigvecs, ikeep, B, K_tilde_b, K_tilde_inv_b = _subspace(K_tilde)
        K_b, KKtilde_inv_b = _project(K, B, K_tilde_inv_b, same)
    else:
        ik = kwargs['init_kernel']
        C, mask, K_tilde, Kvec, B = ik['C'], ik['mask'], ik['K_tilde'], ik['Kvec'], ik['B']
        K = K_tilde if same else ik['K']
        K_tilde_b, K_b, K_tilde_inv_b = ik['K_tilde_b'], ik['K_b'], ik['K_tilde_inv_b']
        KKtilde_inv_b = B if same else ik['KKtilde_inv_b']

    m = copy.deepcopy(kwargs.get('m', torch.zeros((ntilde), dtype=x.dtype, device=x.device)).detach())
    V = copy.deepcopy(kwargs.get('V', K_tilde).detach())
    V_b = B.T @ V @ B if 'V' in kwargs else K_tilde_b
    m_b = B.T @ m

    lambda_m, lambda_var = lambda_moments(x[:, mask], K_tilde_b, KKtilde_inv_b, Kvec, K_b, C, m_b, V_b, theta,
                                          kernfun=kernfun)
    f_mean = mean_f_given_lambda_moments(f_params, lambda_m, lambda_var)
    loglikelihood, _, __ = compute_loglikelihood(r, f_mean, lambda_m, lambda_var, f_params)
    KL_div = compute_KL_div(m_b, V_b, K_tilde_b, K_tilde_inv_b, dK_tilde=None, ignore_warning=True)
    logmarginal = loglikelihood - KL_div

    dev = x.device
    zeros = lambda: torch.zeros((maxiter), dtype=TORCH_DTYPE, device=dev)   # noqa: E731
    lam0_key = 'lambda0' if 'lambda0' in f_params else 'loglambda0'
    values_track = {'loss_track': {'logmarginal': zeros(), 'loglikelihood': zeros(), 'KL': zeros()},
                    'theta_track': {key: zeros() for key in theta.keys()},
                    'f_par_track': {'logA': zeros(), lam0_key: zeros()},
                    'variation_par_track': {'V_b': (), 'm_b': ()}}

    def track(it):
        values_track['loss_track']['loglikelihood'][it].copy_(loglikelihood)
        values_track['loss_track']['KL'][it].copy_(KL_div)
        values_track['loss_track']['logmarginal'][it].copy_(loglikelihood - KL_div)
        for key in theta.keys():
            values_track['theta_track'][key][it].copy_(theta[key])
        values_track['f_par_track']['logA'][it].copy_(f_params['logA'])
        values_track['f_par_track'][lam0_key][it].copy_(f_params[lam0_key])
        values_track['variation_par_track']['V_b'] += (V_b.clone(),)
        values_track['variation_par_track']['m_b'] += (m_b.clone(),)

    def restore(it):
        """Roll theta / f_params / (m_b, V_b) back to the values tracked at iteration it-1 (U:2140-2153)."""
        th = {key: values_track['theta_track'][key][it - 1] for key in theta.keys()}
        f_params['logA'] = values_track['f_par_track']['logA'][it - 1]
        f_params[lam0_key] = values_track['f_par_track'][lam0_key][it - 1]
        return (th, values_track['variation_par_track']['V_b'][it - 1],
                values_track['variation_par_track']['m_b'][it - 1])

    time_estep_total = time_f_params_total = time_mstep_total = 0.0
    time_computing_kernels = time_computing_loss = 0.0
    start_time_loop = time.time()
    iteration = 0
    try:
        track(0)
        print(f'Initial Loss: {-(loglikelihood - KL_div):.4f}')

        for iteration in range(1, maxiter):
            # ---- kernels + subspace at the theta of the last M-step (U:1801-1841) ------------------------------
            t0 = time.time()
            if nMstep > 0 and iteration > 1:
                C, mask, K_tilde, K, Kvec = _kernels_at(theta, *lims, n_px_side, x, xtilde, kernfun, same)
                B_old = B
                eigvals, eigvecs, ikeep, B, K_tilde_b, K_tilde_inv_b = _subspace(K_tilde)
                K_b, KKtilde_inv_b = _project(K, B, K_tilde_inv_b, same)
                V_b = B.T @ (B_old @ V_b @ B_old.T) @ B                                       # U:1833
                m_b = B.T @ B_old @ m_b                                                       # U:1840
            time_computing_kernels += time.time() - t0

            # ---- E-step (U:1861-1937) ---------------------------------------------------------------------
            t0 = time.time()
            if nEstep > 0:
                # lambda_moments only looks at the images when Kvec is missing (U:1093-1094): no per-step gather
                xm_e = x[:, mask] if Kvec is None else None
                for i_estep in range(nEstep):
                    if i_estep == 0 and nMstep > 0:
                        lambda_m, lambda_var = lambda_moments(xm_e, K_tilde_b, KKtilde_inv_b, Kvec, K_b, C, m_b,
                                                              V_b, theta, kernfun=kernfun)
                        f_params['lambda0'] = lambda0_given_logA(f_params['logA'], r, lambda_m, lambda_var)
                    f_mean = mean_f_given_lambda_moments(f_params, lambda_m, lambda_var)
                    m_b, V_b = Estep(r=r, KKtilde_inv=KKtilde_inv_b, m=m_b, f_params=f_params, f_mean=f_mean,
                                     K_tilde=K_tilde_b, K_tilde_inv=K_tilde_inv_b, update_V_inv=False, alpha=1)
                    f_mean, lambda_m, lambda_var = mean_f(f_params=f_params, calculate_moments=True, x=xm_e,
                                                          K_tilde=K_tilde_b, KKtilde_inv=KKtilde_inv_b, Kvec=Kvec,
                                                          K=K_b, C=C, m=m_b, V=V_b, theta=theta, kernfun=kernfun)
                    t1 = time.time()
                    f_params['lambda0'] = lambda0_given_logA(f_params['logA'], r, lambda_m, lambda_var)
                    # the optimiser's own arithmetic is a handful of scalar operations per step: it runs on a host
                    # copy of logA (same IEEE operations, no device launches / read-backs), written back afterwards
                    logA_dev = f_params['logA']
                    f_params['logA'] = logA_dev.detach().cpu().clone()
                    optimizer_f_params = torch.optim.LBFGS([f_params['logA']], lr=0.1, max_iter=nFparamstep,
                                                           tolerance_change=1.e-9, tolerance_grad=1.e-7,
                                                           history_size=nFparamstep, line_search_fn='strong_wolfe')
                    calls = [0]

                    @torch.no_grad()
                    def closure_f_params():
                        # ONE streaming pass over (lambda_m, lambda_var, r): f, loglik, dloglik/dlogA and the
                        # closed-form lambda0 for the new logA (U:1902-1930)
                        nonlocal f_mean
                        calls[0] += 1
                        optimizer_f_params.zero_grad()
                        logA_now, lambda0_now = _link_params(f_params)
                        res, f_mean = native.fparam_eval(lambda_m, lambda_var, r, logA_now, lambda0_now, want_f=True)
                        if _SHARD is not None and _SHARD.world > 1:
                            A_now = math.exp(logA_now)
                            names = ("r_lam", "sum_f", "lamf", "sum_r", "sum_exp", "nan")
                            res.update(zip(names, _shard_sum([res[k] for k in names])))
                            res["loglik"] = A_now * res["r_lam"] + lambda0_now * res["sum_r"] - res["sum_f"]
                            res["dlogA"] = A_now * (res["r_lam"] - res["lamf"])
                            res["lambda0_closed"] = math.log(res["sum_r"]) - math.log(res["sum_exp"])
                        f_params['logA'].grad = _scalar(-res["dlogA"], like=f_params['logA'])
                        f_params['lambda0'] = res["lambda0_closed"]          # float until the optimiser returns
                        if res["nan"] > 0:
                            raise ValueError(f'Nan in f_mean during f param update in Estep, closure has been called '
                                             f'{calls[0]} times in estep {i_estep} iteration. Try substituting them '
                                             f'with inf.')
                        if _SHARD is None and res["sum_f"] / lambda_m.shape[0] > 100:
                            print(f'f_mean mean is {res["sum_f"] / lambda_m.shape[0]} at i_step {i_estep} iteration '
                                  f'{iteration} at closure call {calls[0]}')
                        return _scalar(-res["loglik"], like=f_params['logA'])

                    try:
                        optimizer_f_params.step(closure_f_params)
                    finally:
                        logA_dev.data.copy_(f_params['logA'])
                        logA_dev.grad = (f_params['logA'].grad.to(logA_dev.device)
                                         if f_params['logA'].grad is not None else None)
                        f_params['logA'] = logA_dev
                    f_params['lambda0'] = lambda0_given_logA(f_params['logA'], r, lambda_m, lambda_var)
                    time_f_params_total += time.time() - t1
            else:
                print('No E-step')
            time_estep_total += time.time() - t0

            # ---- loss tracking (U:1953-1991) --------------------------------------------------------------
            t0 = time.time()
            f_mean = mean_f_given_lambda_moments(f_params, lambda_m, lambda_var)
            loglikelihood, _, __ = compute_loglikelihood(r, f_mean, lambda_m, lambda_var, f_params)
            KL_div = compute_KL_div(m_b, V_b, K_tilde_b, K_tilde_inv_b, dK_tilde=None, ignore_warning=True)
            logmarginal = loglikelihood - KL_div
            time_computing_loss += time.time() - t0
            track(iteration)
            print(f'Loss iter {iteration}: {-(loglikelihood - KL_div):.4f}')

            # ---- M-step (U:1999-2114): LBFGS over theta, B frozen, one fused ELBO+grad call per closure -----
            t0 = time.time()
            if nMstep > 0 and iteration < maxiter - 1:
                optimizer_hyperparams = torch.optim.LBFGS(theta.values(), lr=0.1, max_iter=nMstep,
                                                          line_search_fn='strong_wolfe', tolerance_change=1.e-9,
                                                          tolerance_grad=1.e-7, history_size=100)
                mcalls = [0]
                ldV = log_det(V_b, name='V')          # V_b is fixed during the M-step: one log det for all closures
                logdetV = _fl(ldV)

                @torch.no_grad()
                def closure_hyperparams():
                    mcalls[0] += 1
                    optimizer_hyperparams.zero_grad()
                    bad = _out_of_bounds(theta, theta_lower_lims, theta_higher_lims)
                    for key in bad:                                                           # U:2021-2028
                        print(f"{key} = {_fl(theta[key]):.4f} is not within the limits of {theta_lower_lims[key]} "
                              f"and {theta_higher_lims[key]}, returning inifinite loss in closure call {mcalls[0]}")
                        if theta[key].requires_grad:
                            theta[key].grad = _scalar(float('inf'), like=theta[key])
                    if bad:
                        return _scalar(float('inf'), like=r)
                    ll, kl, grad = elbo_and_grad(theta, theta_lower_lims, theta_higher_lims, n_px_side, x, xtilde, r,
                                                 B, m_b, V_b, f_params, logdetV=logdetV, kernfun=kernfun, same=same)
                    for key in theta.keys():                                                  # U:2095-2099
                        if theta[key].requires_grad:
                            theta[key].grad = _scalar(-grad[key], like=theta[key])
                    return _scalar(-(ll - kl), like=r)

                optimizer_hyperparams.step(closure_hyperparams)
            elif iteration < maxiter - 1:
                print(' No M-step')
            time_mstep_total += time.time() - t0

    except (KeyboardInterrupt, Exception) as e:
        if isinstance(e, KeyboardInterrupt):
            print(' ===================  Interrupted  ===================\n')
        else:
            print(f' ===================  Error During iteration: {iteration} =================== \n')
        print(f'During iteration: {iteration}, there should be {iteration} completed iterations')
        fit_parameters['maxiter'] = iteration
        err_dict['is_error'] = True
        err_dict['error'] = e
        if fit_parameters['maxiter'] <= 1:
            # the reference re-raises here, but its `return` inside `finally` swallows the exception again
            # (U:2134-2138 vs U:2316): the caller always gets (fit_model, err_dict) and re-raises err_dict['error']
            print('Too few iterations iterations were done to save')
            fit_parameters['maxiter'] = max(iteration, 1)
        else:
            theta, V_b, m_b = restore(iteration)

    finally:
        if err_dict['is_error']:
            # kernels of the restored theta have yet to be rebuilt (U:2194-2231)
            C, mask, K_tilde, K, Kvec = _kernels_at(theta, *lims, n_px_side, x, xtilde, kernfun, same)
            eigvals, eigvecs, ikeep, B, K_tilde_b, K_tilde_inv_b = _subspace(K_tilde)
            K_b, KKtilde_inv_b = _project(K, B, K_tilde_inv_b, same)
            f_mean, lambda_m, lambda_var = mean_f(f_params=f_params, calculate_moments=True, x=x[:, mask],
                                                  K_tilde=K_tilde_b, KKtilde_inv=KKtilde_inv_b, Kvec=Kvec, K=K_b, C=C,
                                                  m=m_b, V=V_b, theta=theta, kernfun=kernfun)
            loglikelihood = compute_loglikelihood(r, f_mean, lambda_m, lambda_var, f_params)[0]
            KL = compute_KL_div(m_b, V_b, K_tilde_b, K_tilde_inv_b, dK_tilde=None)
            logmarginal = loglikelihood - KL
            last = fit_parameters['maxiter'] - 1
            values_track['loss_track']['loglikelihood'][last] = loglikelihood
            values_track['loss_track']['KL'][last] = KL
            values_track['loss_track']['logmarginal'][last] = logmarginal

        final_kernel = {'C': C, 'mask': mask, 'K_tilde': K_tilde, 'K': K, 'Kvec': Kvec, 'eigvecs': eigvecs}
        if not is_simmetric(V_b, 'V_b'):
            print('Final V_b is not simmetric, maximum difference: ', torch.max(torch.abs(V_b - V_b.T)))
            V_b = (V_b + V_b.T) / 2
        if not is_posdef(V_b, 'V_b'):
            print('Final V_b is not posdef, this should not be possible if you are skipping the last M-step')
            V_b = V_b + torch.eye(V_b.shape[0], dtype=V_b.dtype, device=V_b.device) * EIGVAL_TOL

        print(f'\nTime spent for E-steps:       {time_estep_total:.3f}s,')
        print(f'Time spent for f params:      {time_f_params_total:.3f}s')
        print(f'Time spent for m / V update:  {time_estep_total - time_f_params_total:.3f}s')
        print(f'Time spent for M-steps:       {time_mstep_total:.3f}s')
        print(f'Time spent for All-steps:     {time_estep_total + time_mstep_total:.3f}s')
        print(f'Time spent computing Kernels: {time_computing_kernels:.3f}s')
        print(f'Time spent computing Loss:    {time_computing_loss:.3f}s')
        print(f'\nTime total after init:        {time.time() - start_time_loop:.3f}s')
        print(f"Time total before init:       {time.time() - start_time_before_init:.3f}s")
        print(f'Final Loss: {-_fl(logmarginal):.4f}')

        for key in values_track.keys():
            for subkey in values_track[key].keys():
                values_track[key][subkey] = values_track[key][subkey][:fit_parameters['maxiter']]
        fit_model = {
            'fit_parameters': fit_parameters, 'final_kernel': final_kernel, 'err_dict': err_dict,
            'xtilde': xtilde, 'hyperparams_tuple': (theta, theta_lower_lims, theta_higher_lims),
            'f_params': f_params, 'm_b': m_b, 'V_b': V_b, 'C': C, 'mask': mask, 'K_tilde_b': K_tilde_b,
            'K_tilde_inv_b': K_tilde_inv_b, 'K_b': K_b, 'Kvec': Kvec, 'B': B, 'values_track': values_track,
        }
        return fit_model, err_dict


# ---- persistence (U:46-109, U:312-324, U:689-703): host I/O, same on-disk layout ---------------------------
def save_model(model, directory, additional_description=None):
    """Pickle of the result dict + a text `metadata` file, in the reference's on-disk layout (U:46-109): refuses an
    existing directory, stores the description under model['description'], writes the same description text as
    metadata.  tests/test_gpu_utils.py compares the text with a model written by the reference's own save_model."""
    if os.path.exists(directory):
        raise ValueError(f"Directory {directory} already exists")
    os.makedirs(directory)
    fp, vt = model['fit_parameters'], model['values_track']
    pad = " " * 8

    def row(label, first, last, width=12):
        return f"{pad}{label + ':':<{width}} {_fl(first):>8.4f} -> {_fl(last):>8.4f}"

    lines = ["", f"{pad}Model Description:"]
    for label, key in (("Cell ID", 'cellid'), ("ntilde", 'ntilde'), ("maxiter", 'maxiter'), ("nMstep", 'nMstep'),
                       ("nEstep", 'nEstep')):
        lines.append(f"{pad}{label + ':':<14} {fp[key]:>8}")
    lines.append(f"{pad}MIN_TOLERANCE: {fp['min_tolerance']:>8.12f}")
    lines.append(f"{pad}EIGVAL_TOL:    {fp['eigval_tol']:>8.4f}")
    lines += [pad, f"{pad}Hyperparameters results:", f"{pad}Start                 ->   End"]
    for key in ('sigma_0', 'eps_0x', 'eps_0y', 'Amp', '-2log2beta', '-log2rho2'):
        lines.append(row(key, vt['theta_track'][key][0], vt['theta_track'][key][-1]))
    bt, rt = vt['theta_track']['-2log2beta'], vt['theta_track']['-log2rho2']
    lines += ["", row("beta", logbetaexpr_to_beta(bt[0]), logbetaexpr_to_beta(bt[-1])),
              row("rho", logrhoexpr_to_rho(rt[0]), logrhoexpr_to_rho(rt[-1])), "",
              f"{pad}Link function results [f_params]:", ""]
    la = vt['f_par_track']['logA']
    lines += [row("logA", la[0], la[-1]), "", row("A", math.exp(_fl(la[0])), math.exp(_fl(la[-1])))]
    if 'lambda0' in vt['f_par_track']:
        l0 = (_fl(vt['f_par_track']['lambda0'][0]), _fl(vt['f_par_track']['lambda0'][-1]))
    else:   # the reference raises KeyError here (U:91); the loglambda0 parametrisation is reported as lambda0
        l0 = tuple(math.exp(_fl(vt['f_par_track']['loglambda0'][i])) for i in (0, -1))
    lines += [f"{pad}{'lambda0:':<12} {l0[0]:>8.4f} ->...", f"{pad}  {l0[1]:>8.4f}",
              f"{pad}We are optimising a negative lambda0", pad]
    description = "\n".join(lines)
    if additional_description is not None:
        description += f"\n\n{additional_description}"
    model['description'] = description
    with open(os.path.join(directory, 'model'), 'wb') as f:
        pickle.dump(model, f)
    with open(os.path.join(directory, 'metadata'), 'w') as f:
        f.write(description)


def load_model(directory):
    with open(f'{directory}/model', 'rb') as f:
        return pickle.load(f)


# ---- plotting entry points of the call surface (U:111-310, U:1543-1563) --------------------------------------
# Visualisation is outside the hot path; the notebooks call these after every fit, so the names
# exist with the reference's signatures.  matplotlib is imported lazily: a headless box without it gets a warning
# and None instead of an ImportError at `import utils` (the reference imports it at module level, U:21-22).
def _pyplot():
    try:
        import matplotlib.pyplot as plt
        return plt
    except Exception as e:                                   # pragma: no cover - depends on the host
        warnings.warn(f'matplotlib is not available ({type(e).__name__}): nothing plotted')
        return None


def plot_loss_and_theta_notebook(model, linestyle='-', marker='o', figsize=(10, 10), xlim=None, ylim_logmarg=None,
                                 ylim_lambda0=None, ylim_eigvals=None):
    """Tracks of one fit: -ELBO / loglik / KL, the six hyper-parameters, logA and lambda0, eigenvalues of V_b."""
    plt = _pyplot()
    if plt is None:
        return None
    vt = model['values_track']

    def arr(t):
        return t.detach().cpu().numpy() if torch.is_tensor(t) else t

    fig, axes = plt.subplots(3, 3, figsize=figsize)
    axes = axes.ravel()
    lt = vt['loss_track']
    axes[0].plot(-arr(lt['logmarginal']), linestyle=linestyle, marker=marker)
    axes[0].set_title('-ELBO')
    if ylim_logmarg is not None:
        axes[0].set_ylim(ylim_logmarg)
    axes[1].plot(arr(lt['loglikelihood']), linestyle=linestyle, marker=marker, label='loglikelihood')
    axes[1].plot(arr(lt['KL']), linestyle=linestyle, marker=marker, label='KL')
    axes[1].legend()
    for ax, key in zip(axes[2:8], vt['theta_track'].keys()):
        ax.plot(arr(vt['theta_track'][key]), linestyle=linestyle, marker=marker)
        ax.set_title(key)
    for key, tr in vt['f_par_track'].items():
        axes[8].plot(arr(tr), linestyle=linestyle, marker=marker, label=key)
    if ylim_lambda0 is not None:
        axes[8].set_ylim(ylim_lambda0)
    axes[8].legend()
    if xlim is not None:
        for ax in axes:
            ax.set_xlim(xlim)
    fp = model['fit_parameters']
    fig.suptitle(f"cell {fp.get('cellid')}  ntilde={fp.get('ntilde')}  maxiter={fp.get('maxiter')}  "
                 f"nEstep={fp.get('nEstep')}  nMstep={fp.get('nMstep')}")
    return fig


def plot_fit(R_predicted, rtst, r2, sigma_r2, cellid):
    """Predicted rate over the trial-averaged test response (U:1543-1563)."""
    plt = _pyplot()
    if plt is None:
        return None

    def arr(t):
        return t.detach().cpu().numpy() if torch.is_tensor(t) else t

    fig = plt.figure(figsize=(12, 8))
    ax = fig.add_subplot(1, 1, 1)
    R_predicted, rtst = arr(R_predicted), arr(rtst)
    tvals = 0.05 * torch.arange(len(R_predicted)).numpy()
    ax.plot(tvals, rtst.mean(axis=0), 'k', linewidth=1)
    ax.plot(tvals, R_predicted, color='red', label='GP')
    ax.set_title(f'Pietro adjusted r^2 = {float(r2):.2f} ± {float(sigma_r2):.2f} Cell: {cellid}')
    return fig


def save_pickle(filename, **kwargs):
    path = os.path.join(os.path.dirname(os.path.realpath(__file__)), f'data/{filename}.pkl')
    os.makedirs(os.path.dirname(path), exist_ok=True)
    with open(path, 'wb') as f:
        pickle.dump(kwargs, f)
    return 0
