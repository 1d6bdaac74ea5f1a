"""Synthetic stimuli / spikes / variational state for the Spatial_GP hot path.

Implements the input recipe of SURVEY.md §8(d) so the oracle, the CUDA path, the tests
and bench.py all see bit-identical inputs.  Everything is generated on the CPU in fp64
from seeded generators (``1000+cfg`` images, ``2000+cfg`` spikes, ``3000+cfg`` variational
state) and moved to the device by the caller.

Nothing here is on the hot path; it is plain PyTorch glue.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, Optional

import torch

THETA_KEYS = ("sigma_0", "eps_0x", "eps_0y", "-2log2beta", "-log2rho2", "Amp")

# theta presets of SURVEY.md §8(d)
THETA_DENSE = {"sigma_0": 1.0, "eps_0x": 0.05, "eps_0y": -0.03,
               "-2log2beta": 0.9, "-log2rho2": 3.0, "Amp": 1.0}
THETA_NOTEBOOK = {"sigma_0": 1.0, "eps_0x": 1e-4, "eps_0y": 1e-4,
                  "-2log2beta": 4.8069, "-log2rho2": 4.3069, "Amp": 1.0}


def theta_localised(n_px_side: int) -> Dict[str, float]:
    """`generate_theta` defaults (reference utils.py:793-824): RF width sqrt(10) px, eps0 = 0."""
    beta = (math.sqrt(10.0) / n_px_side) * 2.0
    rho = beta / 2.0
    return {"sigma_0": 1.0, "eps_0x": 0.0, "eps_0y": 0.0,
            "-2log2beta": -2.0 * math.log(2.0 * beta),
            "-log2rho2": -math.log(2.0 * rho * rho), "Amp": 1.0}


THETA_LIMS_LO = {"sigma_0": 0.0, "eps_0x": -1.0, "eps_0y": -1.0,
                 "-2log2beta": -float("inf"), "-log2rho2": -float("inf"), "Amp": 0.0}
THETA_LIMS_HI = {"sigma_0": float("inf"), "eps_0x": 1.0, "eps_0y": 1.0,
                 "-2log2beta": float("inf"), "-log2rho2": float("inf"), "Amp": float("inf")}


@dataclass
class Problem:
    """One synthetic single-cell problem (CPU, fp64)."""
    n_px_side: int
    X: torch.Tensor            # (N, n^2) images
    r: torch.Tensor            # (N,) spike counts (fp64)
    xtilde: torch.Tensor       # (m, n^2) inducing images
    theta: Dict[str, float]
    logA: float = math.log(0.05)
    lambda0: float = -0.5
    cfg: int = 0
    extras: dict = field(default_factory=dict)

    @property
    def N(self) -> int:
        return self.X.shape[0]

    @property
    def m(self) -> int:
        return self.xtilde.shape[0]


def make_images(N: int, n_px_side: int, cfg: int, smooth: bool = False,
                dtype=torch.float64) -> torch.Tensor:
    g = torch.Generator(device="cpu").manual_seed(1000 + cfg)
    X = torch.randn(N, n_px_side * n_px_side, generator=g, dtype=torch.float64, device="cpu")
    if smooth:  # 3x3 box filter, zero padded (mimics natural-image correlations)
        img = X.view(N, 1, n_px_side, n_px_side)
        k = torch.full((1, 1, 3, 3), 1.0 / 9.0, dtype=torch.float64)
        X = torch.nn.functional.conv2d(img, k, padding=1).reshape(N, -1).contiguous()
    return X.to(dtype)


def make_spikes(X: torch.Tensor, n_px_side: int, cfg: int) -> torch.Tensor:
    """Ground-truth cell of SURVEY §8(d): Gaussian RF at (0.2,-0.1), width 0.15, mean count ~0.8."""
    lin = torch.linspace(-1, 1, n_px_side, dtype=torch.float64)
    yc, xc = torch.meshgrid(lin, lin, indexing="ij")
    w = torch.exp(-((xc.flatten() - 0.2) ** 2 + (yc.flatten() + 0.1) ** 2) / (2 * 0.15 ** 2))
    drive = X.to(torch.float64) @ w
    rate = torch.exp(0.8 * drive / drive.std() - 0.5)
    g = torch.Generator(device="cpu").manual_seed(2000 + cfg)
    return torch.poisson(rate, generator=g).to(torch.float64)


def make_problem(N: int, n_px_side: int, m: int, cfg: int = 0,
                 theta: Optional[Dict[str, float]] = None, smooth: bool = False) -> Problem:
    X = make_images(N, n_px_side, cfg, smooth=smooth)
    r = make_spikes(X, n_px_side, cfg)
    xtilde = X[:m].clone()
    return Problem(n_px_side=n_px_side, X=X, r=r, xtilde=xtilde,
                   theta=dict(theta if theta is not None else THETA_DENSE), cfg=cfg)


def make_variational_state(K_tilde_b: torch.Tensor, cfg: int):
    """m_b = 0.1*randn(q), V_b = K_tilde_b/2 + 0.01*W W^T  (deliberately V != K_tilde)."""
    q = K_tilde_b.shape[0]
    g = torch.Generator(device="cpu").manual_seed(3000 + cfg)
    m_b = 0.1 * torch.randn(q, generator=g, dtype=torch.float64)
    W = torch.randn(q, q, generator=g, dtype=torch.float64)
    V_b = 0.5 * K_tilde_b.cpu().to(torch.float64) + 0.01 * (W @ W.T)
    V_b = 0.5 * (V_b + V_b.T)
    return m_b, V_b


def theta_tensors(theta: Dict[str, float], device=None, requires_grad: bool = False):
    """theta as the reference wants it: dict of 0-dim fp64 tensors in THETA_KEYS order... the
    reference dict order is sigma_0, eps_0x, eps_0y, -2log2beta, -log2rho2, Amp (utils.py:824)."""
    return {k: torch.tensor(float(theta[k]), dtype=torch.float64, device=device,
                            requires_grad=requires_grad) for k in THETA_KEYS}
