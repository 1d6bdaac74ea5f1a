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
