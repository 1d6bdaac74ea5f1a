"""Active-learning cycle: add one image, grow the kernels by one row / column, re-factorise, warm-start the fit.

Restates the driver-level step of ``one_cell_active_training.ipynb`` cell 17
as a function, so that it can be tested and timed; the notebook itself runs unchanged on ``utils`` and needs none of
this.  Every tensor operation goes through the module ``U`` handed in -- ``gaussian_processes_b200.utils`` on the GPU,
or the reference's own ``utils`` (oracle/gen_golden.py drives the REAL reference through this very function to make
the fixtures the GPU test compares with).

What the cell does, per added image (names as in the notebook):
  * lift the variational state to the un-projected space with the old basis, V = B V_b B^T, m = B m_b, and pad it:
    identity row / column for V, mean(m) for m;
  * ONE new column of K_tilde: ``acosker(xtilde_updated, last image)``, appended as row and column -- the rest of
    K_tilde is reused (theta has not changed since the last fit);
  * ``K = K_tilde`` when every image in use is an inducing image (ntilde == nt; the only case the notebook implements --
    it raises NotImplementedError otherwise); for a FIXED training set with a growing inducing set (BASELINE.json
    configs[2]: N = 50 000, m up to 1 000) the same idea gives one new column of K: ``acosker(x, last image)``;
  * ``eigh(K_tilde)`` + threshold -> B, diagonal K_tilde_b and its inverse, K_b = K B, a = K_b K_tilde_b^-1 (a = B when
    ntilde == nt): the ``init_kernel`` dict;
  * ``varGP(X_in_use, R_in_use, **model)`` with ``m``, ``V`` and ``init_kernel`` set: the warm start.
"""
from __future__ import annotations

import time
from typing import Optional

import torch


def _matmul(A: torch.Tensor, B: torch.Tensor) -> torch.Tensor:
    if A.is_cuda:
        from . import native
        return native.matmul(A.contiguous(), B.contiguous())
    return A @ B


def grow_model(U, model: dict, X_in_use: torch.Tensor, xtilde_updated: torch.Tensor, timings: Optional[dict] = None) -> dict:
    """kwargs for ``U.varGP(X_in_use, R_in_use, **kwargs)`` after appending ONE inducing image (the last row of
    ``xtilde_updated``) to the fitted ``model`` (a ``fit_model`` dict as returned by varGP).

    ``X_in_use`` is the training set of the next fit: the old one plus the new image when ntilde == nt (the notebook's
    case), or the unchanged training set when only the inducing set grows.  ``model`` is not modified."""
    def tick(name, t0):
        if timings is not None:
            if xtilde_updated.is_cuda:
                torch.cuda.synchronize(xtilde_updated.device)
            timings[name] = timings.get(name, 0.0) + (time.perf_counter() - t0)

    theta = model['hyperparams_tuple'][0]
    fk = model['final_kernel']
    C, mask = fk['C'], fk['mask']
    B_old, V_b, m_b = model['B'], model['V_b'], model['m_b']
    ntilde, nt = xtilde_updated.shape[0], X_in_use.shape[0]
    same = ntilde == nt
    if fk['K_tilde'].shape[0] != ntilde - 1:
        raise ValueError('grow_model appends exactly one inducing image to the fitted model')
    tol = model['fit_parameters'].get('eigval_tol', getattr(U, 'EIGVAL_TOL', 1.e-4))

    # ---- variational state: through the un-projected space, padded for the new inducing image
    t0 = time.perf_counter()
    V = B_old @ V_b @ B_old.T
    V = 0.5 * (V + V.T)
    m = B_old @ m_b
    V_new = torch.eye(ntilde, dtype=V_b.dtype, device=V_b.device)
    V_new[:ntilde - 1, :ntilde - 1] = V
    m_new = torch.cat((m, m.mean()[None]))
    tick('lift_state', t0)

    # ---- kernels: only the new column(s)
    t0 = time.perf_counter()
    um = xtilde_updated[:, mask]
    x_new = um[-1][None]
    col = U.acosker(theta, um, x_new, C=C, dC=None, diag=False)                  # (ntilde, 1); n1 != n2: not symmetrised
    K_tilde = torch.cat((fk['K_tilde'], col[:-1]), dim=1)
    K_tilde = torch.cat((K_tilde, col.T), dim=0)
    if same:
        K = K_tilde
        Kvec = U.acosker(theta, X_in_use[:, mask], x2=None, C=C, dC=None, diag=True)
    else:
        if fk['K'].shape[0] != nt:
            raise ValueError('the training set changed: the cached K does not match it')
        Xm = X_in_use if bool(mask.all()) else X_in_use[:, mask]
        known = getattr(U, 'acosker_known_norms', None)       # ours: the images' norms are the model's Kvec (same theta)
        if known is not None and Xm.is_cuda:
            kcol = known(theta, Xm, fk['Kvec'], x_new, C)                              # (nt, 1), one pass over X
        else:
            kcol = U.acosker(theta, Xm, x_new, C=C, dC=None, diag=False)
        K = torch.cat((fk['K'], kcol), dim=1)
        Kvec = fk['Kvec']                                                         # same images, same theta
    tick('new_kernel_columns', t0)

    # ---- subspace of the grown K_tilde
    t0 = time.perf_counter()
    eigvals, eigvecs = torch.linalg.eigh(K_tilde, UPLO='L')
    ikeep = eigvals > max(eigvals.max() * tol, tol)
    B = eigvecs[:, ikeep]
    K_tilde_b = torch.diag(eigvals[ikeep])
    K_tilde_inv_b = torch.diag_embed(1 / eigvals[ikeep])
    tick('eigh', t0)
    t0 = time.perf_counter()
    K_b = _matmul(K, B)
    KKtilde_inv_b = B if same else K_b * (1 / eigvals[ikeep])[None, :]            # K_b @ diag(1/eig)
    tick('project', t0)

    out = dict(model)
    out['fit_parameters'] = dict(model['fit_parameters'], ntilde=ntilde)
    out['xtilde'] = xtilde_updated
    out['V'], out['m'] = V_new, m_new
    out['init_kernel'] = {'C': C, 'mask': mask, 'K_tilde': K_tilde, 'K': K, 'Kvec': Kvec, 'B': B,
                          'K_tilde_b': K_tilde_b, 'K_b': K_b, 'K_tilde_inv_b': K_tilde_inv_b,
                          'KKtilde_inv_b': KKtilde_inv_b}
    return out


def add_image_and_refit(U, model: dict, X_in_use: torch.Tensor, R_in_use: torch.Tensor, xtilde_updated: torch.Tensor,
                        timings: Optional[dict] = None):
    """One full cycle: grow_model + the warm-started fit.  Returns ``(fit_model, err_dict)`` of ``U.varGP``."""
    kwargs = grow_model(U, model, X_in_use, xtilde_updated, timings)
    t0 = time.perf_counter()
    fit, err = U.varGP(X_in_use, R_in_use, **kwargs)
    if timings is not None:
        if X_in_use.is_cuda:
            torch.cuda.synchronize(X_in_use.device)
        timings['refit'] = timings.get('refit', 0.0) + (time.perf_counter() - t0)
    return fit, err


def score_candidates(U, model: dict, xtilde: torch.Tensor, xstar: torch.Tensor, r_max: int = 100) -> torch.Tensor:
    """Utility of every candidate image (cell 17, 'Calculate the utility of each remaining image'): predictive moments
    of log f at the candidates, then ``nd_utility`` over the counts 0..r_max-1."""
    theta = model['hyperparams_tuple'][0]
    C, mask, B = model['C'], model['mask'], model['B']
    f_params = model['f_params']
    A = torch.exp(f_params['logA'])
    lambda0 = torch.exp(f_params['loglambda0']) if 'loglambda0' in f_params else f_params['lambda0']
    xs = xstar[:, mask]
    Kvec_star = U.acosker(theta, xs, x2=None, C=C, dC=None, diag=True)
    K_star_b = _matmul(U.acosker(theta, xs, x2=xtilde[:, mask], C=C, dC=None, diag=False), B)
    lam_m, lam_var = U.lambda_moments(xs, model['K_tilde_b'], _matmul(K_star_b, model['K_tilde_inv_b']), Kvec_star,
                                      K_star_b, C, model['m_b'], model['V_b'], theta)
    r_masked = torch.arange(0, r_max, dtype=lam_m.dtype, device=lam_m.device)
    return U.nd_utility(A ** 2 * lam_var, A * lam_m + lambda0, r_masked)
