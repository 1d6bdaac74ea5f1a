"""Thin torch-tensor front end of the C ABI (include/sgp_b200.h).

Every function takes CUDA fp64 tensors, passes ``data_ptr()`` + the current CUDA stream to
libsgp_b200.so and returns fresh torch tensors.  CPU tensors raise: there is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional, Sequence

import torch

from . import _lib
from ._lib import lib

THETA_KEYS = ("sigma_0", "eps_0x", "eps_0y", "-2log2beta", "-log2rho2", "Amp")   # utils.py:824
DC_KEYS = ("Amp", "-2log2beta", "-log2rho2", "eps_0x", "eps_0y")                  # utils.py:910


def _need_cuda(*ts):
    for t in ts:
        if t is None:
            continue
        if not t.is_cuda:
            raise RuntimeError("gaussian_processes_b200: CUDA tensors required (no CPU fallback on this path)")
        if t.dtype != torch.float64:
            raise TypeError("gaussian_processes_b200: fp64 tensors required")


def _c(t: torch.Tensor) -> torch.Tensor:
    """row-major view with unit inner stride (copy only if needed)"""
    if t.dim() == 2 and t.stride(1) == 1 and t.stride(0) >= t.shape[1]:
        return t
    return t.contiguous()


def _even(t: torch.Tensor) -> torch.Tensor:
    """row-major view whose leading dimension is even and whose base is 16-byte aligned (GEMM operand rule)"""
    t = _c(t)
    if t.stride(0) % 2 == 0 and t.data_ptr() % 16 == 0:
        return t
    w = torch.zeros((t.shape[0], t.shape[1] + (t.shape[1] & 1)), dtype=t.dtype, device=t.device)
    w[:, :t.shape[1]] = t
    return w[:, :t.shape[1]]


def _ld(t: torch.Tensor) -> int:
    return t.stride(0) if t.dim() == 2 else t.shape[-1]


def _ptr(t: Optional[torch.Tensor]):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def _stream(dev) -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


def _h(t: torch.Tensor) -> _lib.Handle:
    return _lib.handle_for(t.device.index if t.device.index is not None else torch.cuda.current_device())


def theta_array(theta) -> "C.Array":
    vals = [float(theta[k]) for k in THETA_KEYS] if isinstance(theta, dict) else [float(v) for v in theta]
    return (C.c_double * 6)(*vals)


def mask_host(theta, n_px: int) -> torch.Tensor:
    """RF mask of localker (utils.py:880-883) as a CPU bool tensor; pure host arithmetic."""
    buf = (C.c_uint8 * (n_px * n_px))()
    p = C.c_int(0)
    rc = lib.sgp_mask_host(theta_array(theta), int(n_px), buf, C.byref(p))
    if rc != _lib.SGP_OK:
        raise _lib.SgpError(rc, "sgp_mask_host")
    return torch.frombuffer(bytearray(buf), dtype=torch.uint8).bool()


def localker(theta, n_px: int, device, grad: bool = False):
    mask = mask_host(theta, n_px)
    p = int(mask.sum())
    dev = torch.device(device)
    if dev.type != "cuda":
        raise RuntimeError("gaussian_processes_b200: CUDA device required (no CPU fallback on this path)")
    Cm = torch.empty((p, p), dtype=torch.float64, device=dev)
    dC = torch.empty((5, p, p), dtype=torch.float64, device=dev) if grad else None
    h = _h(Cm)
    h.check(lib.sgp_localker(h.ptr, theta_array(theta), int(n_px), _ptr(Cm), p, _ptr(dC), p * p, _stream(dev)))
    return (Cm, mask.to(dev), dC) if grad else (Cm, mask.to(dev))


def acosker(sigma_0: float, x1: torch.Tensor, x2: torch.Tensor, Cm: torch.Tensor, symmetrize: bool) -> torch.Tensor:
    _need_cuda(x1, x2, Cm)
    x1, x2, Cm = _c(x1), _c(x2), _c(Cm)
    n1, p = x1.shape
    n2 = x2.shape[0]
    K = torch.empty((n1, n2), dtype=torch.float64, device=x1.device)
    h = _h(x1)
    h.check(lib.sgp_acosker(h.ptr, _ptr(x1), n1, _ld(x1), _ptr(x2), n2, _ld(x2), int(p), _ptr(Cm), _ld(Cm),
                            float(sigma_0), _ptr(K), n2, 1 if symmetrize else 0, _stream(x1.device)))
    return K


def acosker_kvec(sigma_0: float, x1: torch.Tensor, x2: torch.Tensor, Cm: torch.Tensor, symmetrize: bool):
    """(K, Kvec) in one pass over x1: acosker(x1, x2) and acosker(x1, diag=True) share the x1 C product."""
    _need_cuda(x1, x2, Cm)
    x1, x2, Cm = _c(x1), _c(x2), _c(Cm)
    n1, p = x1.shape
    n2 = x2.shape[0]
    K = torch.empty((n1, n2), dtype=torch.float64, device=x1.device)
    Kvec = torch.empty((n1,), dtype=torch.float64, device=x1.device)
    h = _h(x1)
    h.check(lib.sgp_acosker_kvec(h.ptr, _ptr(x1), n1, _ld(x1), _ptr(x2), n2, _ld(x2), int(p), _ptr(Cm), _ld(Cm),
                                 float(sigma_0), _ptr(K), n2, 1 if symmetrize else 0, _ptr(Kvec), _stream(x1.device)))
    return K, Kvec


def acosker_known(sigma_0: float, x1: torch.Tensor, kvec1: torch.Tensor, x2: torch.Tensor, Cm: torch.Tensor) -> torch.Tensor:
    """acosker(x1, x2) for images whose Kvec at this theta is known (sgp_acosker_known): one pass over x1."""
    _need_cuda(x1, kvec1, x2, Cm)
    x1, x2, Cm, kvec1 = _c(x1), _c(x2), _c(Cm), kvec1.contiguous()
    n1, p = x1.shape
    n2 = x2.shape[0]
    if kvec1.shape != (n1,) or kvec1.dtype != torch.float64:
        raise ValueError("acosker_known: kvec1 must be a float64 vector with one entry per row of x1")
    K = torch.empty((n1, n2), dtype=torch.float64, device=x1.device)
    h = _h(x1)
    h.check(lib.sgp_acosker_known(h.ptr, _ptr(x1), n1, _ld(x1), _ptr(kvec1), _ptr(x2), n2, _ld(x2), int(p), _ptr(Cm),
                                  _ld(Cm), float(sigma_0), _ptr(K), n2, _stream(x1.device)))
    return K


def acosker_diag(sigma_0: float, x1: torch.Tensor, Cm: torch.Tensor) -> torch.Tensor:
    _need_cuda(x1, Cm)
    x1, Cm = _c(x1), _c(Cm)
    n1, p = x1.shape
    Kvec = torch.empty((n1,), dtype=torch.float64, device=x1.device)
    h = _h(x1)
    h.check(lib.sgp_acosker_diag(h.ptr, _ptr(x1), n1, _ld(x1), int(p), _ptr(Cm), _ld(Cm), float(sigma_0),
                                 _ptr(Kvec), _stream(x1.device)))
    return Kvec


def lambda_moments(a: torch.Tensor, K_b: torch.Tensor, Kvec: torch.Tensor, m_b: torch.Tensor, V_b: torch.Tensor):
    _need_cuda(a, K_b, Kvec, m_b, V_b)
    a, K_b, V_b = _c(a), _c(K_b), _c(V_b)
    Kvec, m_b = Kvec.contiguous(), m_b.contiguous()
    N, q = a.shape
    lam_m = torch.empty((N,), dtype=torch.float64, device=a.device)
    lam_var = torch.empty_like(lam_m)
    h = _h(a)
    h.check(lib.sgp_lambda_moments(h.ptr, _ptr(a), _ld(a), _ptr(K_b), _ld(K_b), _ptr(Kvec), N, int(q), _ptr(m_b),
                                   _ptr(V_b), _ld(V_b), _ptr(lam_m), _ptr(lam_var), _stream(a.device)))
    return lam_m, lam_var


def fparam_eval(lam_m: torch.Tensor, lam_var: torch.Tensor, r: torch.Tensor, logA: float, lambda0: float,
                want_f: bool = True):
    """One streaming pass: dict(loglik, dlogA, dlambda0, lambda0_closed, sum_r, sum_f, r_lam, nan) and f."""
    _need_cuda(lam_m, lam_var, r)
    lam_m, lam_var, r = lam_m.contiguous(), lam_var.contiguous(), r.contiguous()
    N = lam_m.shape[0]
    f = torch.empty_like(lam_m) if want_f else None
    out = (C.c_double * 10)()
    h = _h(lam_m)
    h.check(lib.sgp_fparam_eval(h.ptr, _ptr(lam_m), _ptr(lam_var), _ptr(r), N, float(logA), float(lambda0), _ptr(f),
                                out, _stream(lam_m.device)))
    keys = ("loglik", "dlogA", "dlambda0", "lambda0_closed", "sum_r", "sum_f", "r_lam", "nan", "lamf", "sum_exp")
    return dict(zip(keys, list(out))), f


def loglik_sums(r: torch.Tensor, f: torch.Tensor, lam_m: torch.Tensor, lam_var: Optional[torch.Tensor], logA: float):
    """dict(r_lam, sum_f, lamf, sum_r, nan): the N-sized reductions of compute_loglikelihood (utils.py:1240-1255)
    for a caller-supplied f; lamf = sum (lam_m + A lam_var) f."""
    _need_cuda(r, f, lam_m, lam_var)
    r, f, lam_m = r.contiguous(), f.contiguous(), lam_m.contiguous()
    lam_var = lam_var.contiguous() if lam_var is not None else None
    out = (C.c_double * 8)()
    h = _h(r)
    h.check(lib.sgp_loglik_sums(h.ptr, _ptr(r), _ptr(f), _ptr(lam_m), _ptr(lam_var), r.shape[0], float(logA), out,
                                _stream(r.device)))
    return dict(zip(("r_lam", "sum_f", "lamf", "sum_r", "nan"), list(out)[:5]))


def estep_terms(a: torch.Tensor, r: torch.Tensor, f: torch.Tensor, logA: float):
    _need_cuda(a, r, f)
    a, r, f = _c(a), r.contiguous(), f.contiguous()
    N, q = a.shape
    g = torch.empty((q,), dtype=torch.float64, device=a.device)
    G = torch.empty((q, q), dtype=torch.float64, device=a.device)
    h = _h(a)
    h.check(lib.sgp_estep_terms(h.ptr, _ptr(a), _ld(a), _ptr(r), _ptr(f), N, int(q), float(logA), _ptr(g), _ptr(G),
                                q, _stream(a.device)))
    return g, G


def cholesky(A: torch.Tensor):
    """Lower Cholesky factor (new tensor) and status flags."""
    _need_cuda(A)
    n = A.shape[0]
    ld = (n + 1) // 2 * 2
    W = torch.zeros((n, ld), dtype=torch.float64, device=A.device)
    W[:, :n] = A
    fl = C.c_int(0)
    h = _h(A)
    h.check(lib.sgp_cholesky(h.ptr, _ptr(W), ld, int(n), C.byref(fl), _stream(A.device)))
    return torch.tril(W[:, :n]), int(fl.value)


def spd_inverse(A: torch.Tensor):
    """(A^-1, logdet A, flags) through the blocked Cholesky + triangular inverse kernels."""
    _need_cuda(A)
    A = _c(A)
    n = A.shape[0]
    Ainv = torch.empty((n, n), dtype=torch.float64, device=A.device)
    ld_, fl = C.c_double(0.0), C.c_int(0)
    h = _h(A)
    h.check(lib.sgp_spd_inverse(h.ptr, _ptr(A), _ld(A), int(n), _ptr(Ainv), n, C.byref(ld_), C.byref(fl),
                                _stream(A.device)))
    return Ainv, float(ld_.value), int(fl.value)


def gemm_f64(A: torch.Tensor, B: torch.Tensor, a_layout: int = 0, b_layout: int = 0, alpha: float = 1.0,
             out: Optional[torch.Tensor] = None, accumulate: bool = False, small_tile: bool = False):
    """C[M,N] = alpha * sum_k A(m,k) B(n,k); layout 0: tensor is [rows, K]; layout 1: tensor is [K, rows]."""
    _need_cuda(A, B)
    A, B = _even(A), _even(B)
    M, K = (A.shape if a_layout == 0 else (A.shape[1], A.shape[0]))
    N = B.shape[0] if b_layout == 0 else B.shape[1]
    ldc = (N + 1) // 2 * 2
    if out is None:
        out = torch.zeros((M, ldc), dtype=torch.float64, device=A.device)
    h = _h(A)
    h.check(lib.sgp_gemm_f64(h.ptr, int(a_layout), int(b_layout), _ptr(A), _ld(A), _ptr(B), _ld(B), _ptr(out),
                             _ld(out), int(M), int(N), int(K), float(alpha), 1 if accumulate else 0,
                             1 if small_tile else 0, _stream(A.device)))
    return out[:, :N]


def nd_utility(sigma2: torch.Tensor, mu: torch.Tensor, r: torch.Tensor, want_arrays: bool = False):
    """utils.py:416-525 on the device: U[nstar]; with want_arrays also (lam, mask, p, log p), each (nr, nstar)."""
    _need_cuda(sigma2, mu, r)
    sigma2, mu, r = sigma2.contiguous(), mu.contiguous(), r.to(torch.float64).contiguous()
    nstar, nr = sigma2.numel(), r.numel()
    U = torch.empty(nstar, dtype=torch.float64, device=sigma2.device)
    lam = mask = p = logp = None
    if want_arrays:
        lam, p, logp = (torch.empty((nr, nstar), dtype=torch.float64, device=sigma2.device) for _ in range(3))
        mask = torch.empty((nr, nstar), dtype=torch.uint8, device=sigma2.device)
    h = _h(sigma2)
    h.check(lib.sgp_nd_utility(h.ptr, _ptr(sigma2), _ptr(mu), int(nstar), _ptr(r), int(nr), _ptr(U), _ptr(lam),
                               _ptr(mask), _ptr(p), _ptr(logp), _stream(sigma2.device)))
    if want_arrays:
        return U, lam, mask.bool(), p, logp
    return U


def gemm_i8emu(A: torch.Tensor, B: torch.Tensor, slices: int = 8, transposed: bool = False) -> torch.Tensor:
    """A @ B.T (A: [M,K], B: [N,K]) -- or A.T @ B (A: [K,M], B: [K,N]) with ``transposed`` -- on
    tcgen05.mma.kind::i8 from int8 digit slices, recombined in fp64."""
    _need_cuda(A, B)
    A, B = _c(A), _c(B)
    if transposed:
        K, M = A.shape
        N = B.shape[1]
    else:
        M, K = A.shape
        N = B.shape[0]
    out = torch.empty((M, N), dtype=torch.float64, device=A.device)
    h = _h(A)
    h.check(lib.sgp_gemm_i8emu(h.ptr, 1 if transposed else 0, _ptr(A), int(M), _ld(A), _ptr(B), int(N), _ld(B),
                               int(K), int(slices), _ptr(out), _ld(out), _stream(A.device)))
    return out


def matmul(A: torch.Tensor, B: torch.Tensor) -> torch.Tensor:
    """A @ B on the DMMA GEMM (A: [M,K], B: [K,N]); contiguous result."""
    small = max(A.shape[0], B.shape[1]) <= 1024
    return gemm_f64(A, B, 0, 1, small_tile=small).contiguous()


def matmul_nt(A: torch.Tensor, Bt: torch.Tensor) -> torch.Tensor:
    """A @ Bt.T on the DMMA GEMM (A: [M,K], Bt: [N,K])."""
    small = max(A.shape[0], Bt.shape[0]) <= 1024
    return gemm_f64(A, Bt, 0, 0, small_tile=small).contiguous()


def matmul_tn(At: torch.Tensor, B: torch.Tensor) -> torch.Tensor:
    """At.T @ B on the DMMA GEMM (At: [K,M], B: [K,N])."""
    small = max(At.shape[1], B.shape[1]) <= 1024
    return gemm_f64(At, B, 1, 1, small_tile=small).contiguous()


def rate(lam_m: torch.Tensor, lam_var: torch.Tensor, logA: float, lambda0: float) -> torch.Tensor:
    """f = exp(A lam_m + A^2 lam_var / 2 + lambda0)  (utils.py:1138), one streaming pass."""
    _, f = fparam_eval(lam_m, lam_var, lam_m, logA, lambda0, want_f=True)
    return f


OUT_KEYS = ("loglik", "KL", "g_sigma_0", "g_eps_0x", "g_eps_0y", "g_-2log2beta", "g_-log2rho2", "g_Amp",
            "flags", "p", "sum_r", "sum_f", "r_lam", "logdetK", "logdetV", "projected")


def _check_closure_shapes(X, r, xtilde, n_px, B, m_b, V_b):
    """The C ABI sees pointers and leading dimensions only: the shapes are checked here, where they are known (the
    reference fails on these mismatches inside x[:, mask] / its matmuls)."""
    P = int(n_px) * int(n_px)
    if X.dim() != 2 or xtilde.dim() != 2 or X.shape[1] != P or xtilde.shape[1] != P:
        raise ValueError(f"images must be (N, {P}) and (ntilde, {P}) for n_px_side = {n_px}: got {tuple(X.shape)} and "
                         f"{tuple(xtilde.shape)}")
    N, m = X.shape[0], xtilde.shape[0]
    if r.shape != (N,):
        raise ValueError(f"r must have one spike count per image: got {tuple(r.shape)} for {N} images")
    if B.dim() != 2 or B.shape[0] != m or B.shape[1] < 1 or B.shape[1] > m:
        raise ValueError(f"B must be (ntilde, q) with 1 <= q <= ntilde = {m}: got {tuple(B.shape)}")
    q = B.shape[1]
    if m_b.shape != (q,) or V_b.shape != (q, q):
        raise ValueError(f"m_b must be ({q},) and V_b ({q}, {q}): got {tuple(m_b.shape)} and {tuple(V_b.shape)}")


def elbo_grad(X: torch.Tensor, r: torch.Tensor, xtilde: torch.Tensor, n_px: int, theta, B: torch.Tensor,
              m_b: torch.Tensor, V_b: torch.Tensor, logA: float, lambda0: float, lims=None,
              logdetV: Optional[float] = None, want_moments: bool = False, kl_weight: float = 1.0,
              same: Optional[bool] = None) -> Dict:
    """Fused M-step closure body (utils.py:2030-2099): one C call, one 16-double read-back.
    kl_weight scales the KL part of the returned gradient (image shards other than the first pass 0).
    same: None = as the reference decides (ntilde == nt by the row counts: K = K_tilde, a = B); False = never (callers
    whose X is a shard of the images)."""
    _need_cuda(X, r, xtilde, B, m_b, V_b)
    _dev = X.device.index if X.device.index is not None else torch.cuda.current_device()
    set_option(_dev, "kl_weight", kl_weight)
    set_option(_dev, "square_is_same", 0.0 if same is False else 1.0)
    X, xtilde, B, V_b = _c(X), _c(xtilde), _c(B), _c(V_b)
    r, m_b = r.contiguous(), m_b.contiguous()
    _check_closure_shapes(X, r, xtilde, n_px, B, m_b, V_b)
    N, m, q = X.shape[0], xtilde.shape[0], B.shape[1]
    lam_m = lam_var = f = None
    if want_moments:
        lam_m = torch.empty((N,), dtype=torch.float64, device=X.device)
        lam_var, f = torch.empty_like(lam_m), torch.empty_like(lam_m)
    lo = hi = None
    if lims is not None:
        lo = (C.c_double * 6)(*[float(lims[0][k]) for k in THETA_KEYS])
        hi = (C.c_double * 6)(*[float(lims[1][k]) for k in THETA_KEYS])
    ldv_val = C.c_double(logdetV if logdetV is not None else 0.0)
    out = (C.c_double * 16)()
    h = _h(X)
    h.bind_images(X, xtilde)     # the library's cached copies / int8 slices of the images are keyed on this token
    h.check(lib.sgp_elbo_grad(h.ptr, _ptr(X), _ld(X), _ptr(r), N, _ptr(xtilde), _ld(xtilde), int(m), int(n_px),
                              theta_array(theta), lo, hi, _ptr(B), _ld(B), int(q), _ptr(m_b), _ptr(V_b), _ld(V_b),
                              float(logA), float(lambda0), C.byref(ldv_val) if logdetV is not None else None,
                              _ptr(lam_m), _ptr(lam_var), _ptr(f), out, _stream(X.device)))
    res = dict(zip(OUT_KEYS, list(out)))
    res["grad"] = {k: res["g_" + k] for k in THETA_KEYS}
    res["flags"] = int(res["flags"])
    res["p"] = int(res["p"])
    if want_moments:
        res.update(lam_m=lam_m, lam_var=lam_var, f=f)
    return res


def elbo_grad_f32(X: torch.Tensor, r: torch.Tensor, xtilde: torch.Tensor, n_px: int, theta, B: torch.Tensor,
                  m_b: torch.Tensor, V_b: torch.Tensor, logA: float, lambda0: float, lims=None,
                  logdetV: Optional[float] = None, want_moments: bool = False, kl_weight: float = 1.0,
                  same: Optional[bool] = None) -> Dict:
    """fp32 mode (sgp_elbo_grad_f32): the images X / xtilde are float32 CUDA tensors, everything else fp64.  The library
    keeps its own fp64 working copy + int8 slices of the images; products run on the fp32-class engine (5 int8 slices)
    unless option "i8_slices" says otherwise.  Same result dict as elbo_grad."""
    for t in (X, xtilde):
        if not t.is_cuda or t.dtype != torch.float32:
            raise TypeError("elbo_grad_f32 expects float32 CUDA images")
    _need_cuda(r, B, m_b, V_b)
    _dev = X.device.index if X.device.index is not None else torch.cuda.current_device()
    set_option(_dev, "kl_weight", kl_weight)
    set_option(_dev, "square_is_same", 0.0 if same is False else 1.0)
    X, xtilde, B, V_b = _c(X), _c(xtilde), _c(B), _c(V_b)
    r, m_b = r.contiguous(), m_b.contiguous()
    _check_closure_shapes(X, r, xtilde, n_px, B, m_b, V_b)
    N, m, q = X.shape[0], xtilde.shape[0], B.shape[1]
    lam_m = lam_var = f = None
    if want_moments:
        lam_m = torch.empty((N,), dtype=torch.float64, device=X.device)
        lam_var, f = torch.empty_like(lam_m), torch.empty_like(lam_m)
    lo = hi = None
    if lims is not None:
        lo = (C.c_double * 6)(*[float(lims[0][k]) for k in THETA_KEYS])
        hi = (C.c_double * 6)(*[float(lims[1][k]) for k in THETA_KEYS])
    ldv_val = C.c_double(logdetV if logdetV is not None else 0.0)
    out = (C.c_double * 16)()
    h = _h(X)
    h.bind_images(X, xtilde)
    h.check(lib.sgp_elbo_grad_f32(h.ptr, _ptr(X), _ld(X), _ptr(r), N, _ptr(xtilde), _ld(xtilde), int(m), int(n_px),
                                  theta_array(theta), lo, hi, _ptr(B), _ld(B), int(q), _ptr(m_b), _ptr(V_b), _ld(V_b),
                                  float(logA), float(lambda0), C.byref(ldv_val) if logdetV is not None else None,
                                  _ptr(lam_m), _ptr(lam_var), _ptr(f), out, _stream(X.device)))
    res = result_dict(list(out))
    if want_moments:
        res.update(lam_m=lam_m, lam_var=lam_var, f=f)
    return res


def elbo_grad_dev(X: torch.Tensor, r: torch.Tensor, xtilde: torch.Tensor, n_px: int, theta, B: torch.Tensor,
                  m_b: torch.Tensor, V_b: torch.Tensor, logA: float, lambda0: float,
                  logdetV: Optional[float] = None, kl_weight: float = 1.0,
                  out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Asynchronous fused closure (sgp_elbo_grad_dev): returns the 16 result doubles as a DEVICE tensor (layout
    OUT_KEYS) without synchronising the host; the caller reduces / reads them in stream order."""
    _need_cuda(X, r, xtilde, B, m_b, V_b)
    _dev = X.device.index if X.device.index is not None else torch.cuda.current_device()
    set_option(_dev, "kl_weight", kl_weight)
    set_option(_dev, "square_is_same", 0.0)       # image shards: the local row count says nothing about ntilde == nt
    X, xtilde, B, V_b = _c(X), _c(xtilde), _c(B), _c(V_b)
    r, m_b = r.contiguous(), m_b.contiguous()
    _check_closure_shapes(X, r, xtilde, n_px, B, m_b, V_b)
    N, m, q = X.shape[0], xtilde.shape[0], B.shape[1]
    if out is None:
        out = torch.empty(16, dtype=torch.float64, device=X.device)
    ldv_val = C.c_double(logdetV if logdetV is not None else 0.0)
    h = _h(X)
    h.bind_images(X, xtilde)
    h.check(lib.sgp_elbo_grad_dev(h.ptr, _ptr(X), _ld(X), _ptr(r), N, _ptr(xtilde), _ld(xtilde), int(m), int(n_px),
                                  theta_array(theta), _ptr(B), _ld(B), int(q), _ptr(m_b), _ptr(V_b), _ld(V_b),
                                  float(logA), float(lambda0), C.byref(ldv_val) if logdetV is not None else None,
                                  None, None, None, _ptr(out), _stream(X.device)))
    return out


def result_dict(vals) -> Dict:
    """OUT_KEYS dict (+ grad / int flags) from 16 doubles."""
    res = dict(zip(OUT_KEYS, [float(v) for v in vals]))
    res["grad"] = {k: res["g_" + k] for k in THETA_KEYS}
    res["flags"] = int(res["flags"])
    res["p"] = int(res["p"])
    return res


def elbo_grad_cells(X: torch.Tensor, r: Sequence[torch.Tensor], xtilde: torch.Tensor, n_px: int, thetas: Sequence,
                    B: Sequence[torch.Tensor], m_b: Sequence[torch.Tensor], V_b: Sequence[torch.Tensor],
                    logA: Sequence[float], lambda0: Sequence[float],
                    logdetV: Optional[Sequence[float]] = None) -> list:
    """Population fit (sgp_elbo_grad_cells): one fused closure per cell, back to back on the current stream, all cells
    on the same resident images and the same int8 slices of them.  Returns one result dict per cell."""
    nc = len(thetas)
    if not (len(r) == len(B) == len(m_b) == len(V_b) == len(logA) == len(lambda0) == nc):
        raise ValueError("elbo_grad_cells: per-cell argument lists differ in length")
    _need_cuda(X, xtilde, *r, *B, *m_b, *V_b)
    X, xtilde = _c(X), _c(xtilde)
    r = [t.contiguous() for t in r]
    B = [_c(t) for t in B]
    V_b = [_c(t) for t in V_b]
    m_b = [t.contiguous() for t in m_b]
    N, m = X.shape[0], xtilde.shape[0]
    th = (C.c_double * (6 * nc))(*[float(t[k]) if isinstance(t, dict) else float(t[i])
                                    for t in thetas for i, k in enumerate(THETA_KEYS)])

    def parr(ts):
        return (C.c_void_p * nc)(*[t.data_ptr() for t in ts])
    ldb = (C.c_long * nc)(*[_ld(t) for t in B])
    ldv = (C.c_long * nc)(*[_ld(t) for t in V_b])
    qs = (C.c_int * nc)(*[t.shape[1] for t in B])
    la = (C.c_double * nc)(*[float(v) for v in logA])
    l0 = (C.c_double * nc)(*[float(v) for v in lambda0])
    ldvv = (C.c_double * nc)(*[float(v) for v in logdetV]) if logdetV is not None else None
    out = (C.c_double * (16 * nc))()
    h = _h(X)
    h.bind_images(X, xtilde)
    h.check(lib.sgp_elbo_grad_cells(h.ptr, _ptr(X), _ld(X), N, _ptr(xtilde), _ld(xtilde), int(m), int(n_px), nc, parr(r),
                                    th, parr(B), ldb, qs, parr(m_b), parr(V_b), ldv, la, l0, ldvv, out,
                                    _stream(X.device)))
    vals = list(out)
    return [result_dict(vals[16 * c:16 * c + 16]) for c in range(nc)]


def elbo_grad_host(X: torch.Tensor, r: torch.Tensor, xtilde: torch.Tensor, n_px: int, theta, B: torch.Tensor,
                   m_b: torch.Tensor, V_b: torch.Tensor, logA: float, lambda0: float, device: int = 0) -> Dict:
    """Same evaluation with every operand in HOST memory (pinned preferred); H2D streamed slab by slab."""
    for t in (X, r, xtilde, B, m_b, V_b):
        if t.is_cuda or t.dtype != torch.float64:
            raise TypeError("elbo_grad_host expects CPU fp64 tensors")
    X, xtilde, B, V_b = _c(X), _c(xtilde), _c(B), _c(V_b)
    r, m_b = r.contiguous(), m_b.contiguous()
    _check_closure_shapes(X, r, xtilde, n_px, B, m_b, V_b)
    N, m, q = X.shape[0], xtilde.shape[0], B.shape[1]
    out = (C.c_double * 16)()
    h = _lib.handle_for(device)
    h.check(lib.sgp_elbo_grad_host(h.ptr, _ptr(X), _ld(X), _ptr(r), N, _ptr(xtilde), _ld(xtilde), int(m), int(n_px),
                                   theta_array(theta), _ptr(B), _ld(B), int(q), _ptr(m_b), _ptr(V_b), _ld(V_b),
                                   float(logA), float(lambda0), out))
    res = dict(zip(OUT_KEYS, list(out)))
    res["grad"] = {k: res["g_" + k] for k in THETA_KEYS}
    res["flags"] = int(res["flags"])
    return res


def comm_export(device: int, world: int) -> bytes:
    """Allocate this rank's exchange buffer and return its 64-byte CUDA IPC handle (sgp_comm_export)."""
    h = _lib.handle_for(device)
    buf = (C.c_ubyte * 64)()
    h.check(lib.sgp_comm_export(h.ptr, int(world), buf))
    return bytes(buf)


def comm_import(device: int, rank: int, world: int, handles: bytes) -> None:
    """Map every peer's exchange buffer (handles: world x 64 bytes, in rank order); from now on sgp_elbo_grad /
    sgp_elbo_grad_dev on this thread's handle return results already reduced over the image shards."""
    if len(handles) != 64 * world:
        raise ValueError("comm_import: need world x 64 bytes of IPC handles")
    h = _lib.handle_for(device)
    buf = (C.c_ubyte * len(handles)).from_buffer_copy(handles)
    h.check(lib.sgp_comm_import(h.ptr, int(rank), int(world), buf))


def comm_close(device: int) -> None:
    h = _lib.handle_for(device)
    h.check(lib.sgp_comm_close(h.ptr))


def upload_images(X: torch.Tensor, r: torch.Tensor, xtilde: torch.Tensor, n_px: int, device: int = 0) -> None:
    """Resident-image protocol, step 1 (once per fit): host images / spike counts / inducing images -> device buffers owned
    by this thread's handle (sgp_upload_images)."""
    for t in (X, r, xtilde):
        if t.is_cuda or t.dtype != torch.float64:
            raise TypeError("upload_images expects CPU fp64 tensors")
    X, xtilde, r = _c(X), _c(xtilde), r.contiguous()
    h = _lib.handle_for(device)
    h.check(lib.sgp_upload_images(h.ptr, _ptr(X), _ld(X), _ptr(r), X.shape[0], _ptr(xtilde), _ld(xtilde),
                                  int(xtilde.shape[0]), int(n_px)))
    h._bound = None


def elbo_grad_resident(theta, B: torch.Tensor, m_b: torch.Tensor, V_b: torch.Tensor, logA: float, lambda0: float,
                       lims=None, device: int = 0) -> Dict:
    """Resident-image protocol, step 2 (per closure evaluation): theta by value, B / m_b / V_b from HOST memory (pinned
    preferred); the images uploaded by upload_images are reused (sgp_elbo_grad_resident)."""
    for t in (B, m_b, V_b):
        if t.is_cuda or t.dtype != torch.float64:
            raise TypeError("elbo_grad_resident expects CPU fp64 tensors")
    B, V_b, m_b = _c(B), _c(V_b), m_b.contiguous()
    lo = hi = None
    if lims is not None:
        lo = (C.c_double * 6)(*[float(lims[0][k]) for k in THETA_KEYS])
        hi = (C.c_double * 6)(*[float(lims[1][k]) for k in THETA_KEYS])
    out = (C.c_double * 16)()
    h = _lib.handle_for(device)
    h.check(lib.sgp_elbo_grad_resident(h.ptr, theta_array(theta), lo, hi, _ptr(B), _ld(B), int(B.shape[1]), _ptr(m_b),
                                       _ptr(V_b), _ld(V_b), float(logA), float(lambda0), out))
    return result_dict(list(out))


def debug_get(device, name: str) -> torch.Tensor:
    h = _lib.handle_for(device.index if isinstance(device, torch.device) else int(device))
    rows, cols, ld = C.c_long(0), C.c_long(0), C.c_long(0)
    h.check(lib.sgp_debug_get(h.ptr, name.encode(), None, 0, C.byref(rows), C.byref(cols), C.byref(ld)))
    buf = torch.empty((rows.value, ld.value), dtype=torch.float64, device=f"cuda:{h.device}")
    h.check(lib.sgp_debug_get(h.ptr, name.encode(), _ptr(buf), buf.numel(), None, None, None))
    return buf[:, :cols.value]


def set_slab_rows(device: int, rows: int):
    h = _lib.handle_for(device)
    h.check(lib.sgp_set_slab_rows(h.ptr, int(rows)))


def launch_count(device: int = 0, reset: bool = False) -> int:
    return _lib.handle_for(device).launch_count(reset)


# with int8 slices stage 6 is the transposed slicing of R (and of X when not cached) and stage 10 the product itself
STAGES = ("setup+inducing_fwd", "quad_forms", "gemm_X_UC", "gemm_moments", "rows+acos_adjoint", "wq_reduce", "gemm_XtR",
          "gemm_KtvK", "tail_qxq+inducing_bwd+contract", "acos_map", "gemm_XtR_i8")


def set_option(device: int, name: str, value: float):
    h = _lib.handle_for(device)
    h.check(lib.sgp_set_option(h.ptr, name.encode(), float(value)))


def invalidate_cache(device: int = 0):
    """Forget the cached gathered copies / int8 slices of the images.  Not needed for torch tensors: elbo_grad binds
    an image token from (tensor identity, version counter, pointer, shape) on every call."""
    h = _lib.handle_for(device)
    h.check(lib.sgp_invalidate_cache(h.ptr))
    h.new_image_epoch()


def new_image_epoch(device: int = 0):
    """Start a new image token on this thread's handle (utils.varGP calls it on entry)."""
    _lib.handle_for(device).new_image_epoch()


def release_cache(device: int = 0):
    """invalidate_cache + give the memory of the cached copies, slices, staging buffers and scratch arena back."""
    h = _lib.handle_for(device)
    h.check(lib.sgp_release_cache(h.ptr))
    h.new_image_epoch()


def get_profile(device: int = 0, reset: bool = True):
    """Per-stage device milliseconds accumulated by sgp_elbo_grad while option "profile" is on."""
    h = _lib.handle_for(device)
    ms = (C.c_double * 16)()
    calls = C.c_long(0)
    h.check(lib.sgp_get_profile(h.ptr, ms, 16, C.byref(calls), 1 if reset else 0))
    return dict(zip(STAGES, list(ms)[:len(STAGES)])), int(calls.value)
