"""Plain-PyTorch (CPU, fp64) statement of the *reverse-mode* ELBO+gradient evaluation that the CUDA
path implements.  Test helper only: it (a) proves that the adjoint contraction
reproduces the reference's forward-mode analytic gradient to round-off, and (b) exposes every
intermediate the CUDA kernels produce (Z, K, cos, T, R, Y, GC, ...) under the same names, so GPU
tests can localise a failing stage.

The local partials are the reference's own (SURVEY §8a parity trap 2): d cos/d lin := 1/(X1 X2),
d cos/d(X1X2) := -cos/(X1 X2), dJ/dcos := (pi~ - delta)/pi~ -- the +1e-7 and the clip are ignored
by the derivative exactly as in U:998/U:1017.
"""
import math

import torch

PI32 = 3.1415927410125732
KEYS = ("sigma_0", "eps_0x", "eps_0y", "-2log2beta", "-log2rho2", "Amp")


def grid(n):
    lin = torch.linspace(-1, 1, n, dtype=torch.float64)
    gy, gx = torch.meshgrid(lin, lin, indexing="ij")
    return gx.reshape(-1), gy.reshape(-1)


def build_C(theta, n):
    eb, er = math.exp(theta["-2log2beta"]), math.exp(theta["-log2rho2"])
    gx, gy = grid(n)
    la = -eb * ((gx - theta["eps_0x"]) ** 2 + (gy - theta["eps_0y"]) ** 2)
    al = torch.exp(la)
    mask = al >= 1e-3
    gx, gy, la, al = gx[mask], gy[mask], la[mask], al[mask]
    ls = -er * ((gx[None] - gx[:, None]) ** 2 + (gy[None] - gy[:, None]) ** 2)
    C = theta["Amp"] * al[:, None] * torch.exp(ls) * al[None]
    C = 0.5 * (C + C.T)
    return C, mask, dict(gx=gx, gy=gy, la=la, ls=ls, eb=eb)


def acos_forward(xa, C, s02, xb=None, nb=None):
    """Z = xa C, s = rowdot(Z, xa) + s0^2, and (if xb given) cos / P / Q / Kraw against xb."""
    Z = xa @ C
    s = (Z * xa).sum(1) + s02
    na = torch.sqrt(s)
    if xb is None:
        return Z, s, na
    nn = na[:, None] * nb[None, :]
    lin = Z @ xb.T + s02
    cos = torch.clip(lin / (nn + 1e-7), -1, 1)
    delta = torch.arccos(cos)
    P = (PI32 - delta) / PI32
    Q = torch.sqrt(1 - cos * cos) / PI32
    K = nn * (Q + P * cos)          # == nn * (sqrt(1-c^2) + pi c - delta c)/pi  (U:988-990)
    return Z, s, na, cos, P, Q, K


def acos_backward(GK, P, Q, na, nb):
    """Adjoint of K wrt (lin, na^2-quadratic forms): R, row weights, column weights, sum(R)."""
    R = GK * P
    Rq = GK * Q
    rho = Rq @ nb                   # sum_j GK_ij Q_ij X2_j
    kap = Rq.T @ na                 # sum_i GK_ij Q_ij X1_i
    return R, 0.5 * rho / na, 0.5 * kap / nb


def elbo_and_grad(theta, X, r, xtilde, n, B, m_b, V_b, logA, lambda0, kl_weight=1.0):
    """Returns dict(loglik, KL, grad[key], lam_m, lam_var, f, inter=dict of intermediates)."""
    th = {k: float(v) for k, v in theta.items()}
    s0 = th["sigma_0"]
    s02 = s0 * s0
    A = math.exp(logA)
    C, mask, aux = build_C(th, n)
    Xm, Um = X[:, mask], xtilde[:, mask]
    N, m = Xm.shape[0], Um.shape[0]
    if N == m:
        raise NotImplementedError("ntilde == nt (a = B) uses a separate small-problem path")

    # ---- inducing-point path (m x m) ------------------------------------------------------
    ZU, tU, X2 = acos_forward(Um, C, s02)
    _, _, _, cosU, PU, QU, KtRaw = acos_forward(Um, C, s02, Um, X2)
    Kt = 0.5 * (KtRaw + KtRaw.T)
    Kt_b = B.T @ Kt @ B
    Kt_b = 0.5 * (Kt_b + Kt_b.T)
    q = Kt_b.shape[0]
    Kinv = torch.linalg.solve(Kt_b, torch.eye(q, dtype=torch.float64))
    b = Kinv @ m_b
    bb = B @ b
    KVK = Kinv @ V_b @ Kinv.T
    Mg = 2 * KVK - Kinv - Kinv.T
    Mfull = B @ Mg @ B.T

    # ---- image pass, forward ----------------------------------------------------------------
    Z, s, X1, cos, P, Q, K = acos_forward(Xm, C, s02, Um, X2)
    T = K @ Mfull.T
    lam_m = K @ bb
    lam_var = s + 0.5 * (T * K).sum(1)
    f = torch.exp(A * lam_m + 0.5 * A * A * lam_var + lambda0)
    loglik = A * (r @ lam_m) + lambda0 * r.sum() - f.sum()

    # ---- image pass, backward ---------------------------------------------------------------
    u = A * (r - f)
    v = -0.5 * A * A * f
    GK = u[:, None] * bb[None, :] + v[:, None] * T
    R, wrow, beta = acos_backward(GK, P, Q, X1, X2)
    w = v + wrow
    Y = R @ Um + w[:, None] * Xm
    GC = Xm.T @ Y
    y = K.T @ u
    Sig2 = K.T @ (v[:, None] * K)
    sig_acc = R.sum() + w.sum() + beta.sum()

    # ---- q x q algebra: adjoint of K_tilde_b ---------------------------------------------------
    y_b = B.T @ y
    Sig2_b = B.T @ Sig2 @ B
    S = Kinv.T @ (torch.outer(y_b, m_b) + Sig2_b @ (2 * Kinv @ V_b.T - torch.eye(q, dtype=torch.float64)))
    # kl_weight: image shards other than the first leave the (N-independent) KL adjoint out, so that the
    # all-reduced gradient counts it once
    GKt_b = -S @ Kinv.T + kl_weight * (-0.5 * Kinv.T + 0.5 * (Kinv @ V_b @ Kinv).T + 0.5 * torch.outer(b, b))
    GKt = B @ GKt_b @ B.T

    # ---- inducing-point path, backward ------------------------------------------------------
    RU, wr, wc = acos_backward(GKt, PU, QU, X2, X2)
    wU = wr + wc + beta
    YU = RU @ Um + wU[:, None] * Um
    GC = GC + Um.T @ YU
    sig_acc = sig_acc + RU.sum() + (wr + wc).sum()

    # ---- contraction with dC_k (never materialised on the GPU) ---------------------------------
    gx, gy, la, ls, eb = aux["gx"], aux["gy"], aux["la"], aux["ls"], aux["eb"]
    GCC = GC * C
    grad = {
        "sigma_0": float(2 * s0 * sig_acc),
        "Amp": float(GCC.sum() / th["Amp"]),
        "-2log2beta": float((GCC * (la[:, None] + la[None, :])).sum()),
        "-log2rho2": float((GCC * ls).sum()),
        "eps_0x": float(2 * eb * (GCC * (gx[:, None] + gx[None, :] - 2 * th["eps_0x"])).sum()),
        "eps_0y": float(2 * eb * (GCC * (gy[:, None] + gy[None, :] - 2 * th["eps_0y"])).sum()),
    }

    # ---- KL ------------------------------------------------------------------------------------
    def logdet(M):
        L = torch.linalg.cholesky(M)
        return 2 * torch.log(torch.diag(L)).sum()
    KL = -0.5 * logdet(V_b) + 0.5 * logdet(Kt_b) + 0.5 * (m_b @ b) + 0.5 * torch.trace(V_b @ Kinv)

    inter = dict(C=C, mask=mask, Z=Z, s=s, X1=X1, X2=X2, cos=cos, K=K, T=T, u=u, v=v, R=R, w=w, beta=beta, Y=Y,
                 GC=GC, y=y, Sig2=Sig2, Kt=Kt, Kt_b=Kt_b, Kinv=Kinv, b=b, bb=bb, Mg=Mg, Mfull=Mfull, GKt_b=GKt_b,
                 GKt=GKt, cosU=cosU, KtRaw=KtRaw, sig_acc=sig_acc)
    return dict(loglik=float(loglik), KL=float(KL), elbo=float(loglik - KL), grad=grad, lam_m=lam_m,
                lam_var=lam_var, f=f, inter=inter)
