"""GPU: the DMMA GEMM family, the blocked Cholesky / inverse and the function-level entry points,
each against a plain torch fp64 computation of the same op (kernel-level numerics) and against
fixtures produced by the real reference (function-level parity).  All calls go through the C ABI."""
import numpy as np
import pytest
import torch

from helpers import case_inputs, load, rel_err

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def nat():
    from gaussian_processes_b200 import native
    return native


def dev():
    return torch.device("cuda:0")


@pytest.mark.parametrize("small", [False, True])
@pytest.mark.parametrize("al,bl", [(0, 0), (0, 1), (1, 0), (1, 1)])
@pytest.mark.parametrize("M,N,K", [(128, 128, 16), (130, 70, 37), (257, 129, 200), (64, 400, 1000), (1, 1, 1),
                                   (300, 300, 3)])
def test_gemm_layouts(nat, small, al, bl, M, N, K):
    g = torch.Generator(device="cpu").manual_seed(M * 7 + N * 3 + K)
    A = torch.randn(M, K, generator=g, dtype=torch.float64)
    B = torch.randn(N, K, generator=g, dtype=torch.float64)
    ref = A @ B.T

    def pad(t):  # even leading dimension, as the ABI requires
        r, c = t.shape
        w = torch.zeros(r, c + (c & 1), dtype=torch.float64)
        w[:, :c] = t
        return w.to(dev())[:, :c]
    Ad = pad(A if al == 0 else A.T.contiguous())
    Bd = pad(B if bl == 0 else B.T.contiguous())
    out = nat.gemm_f64(Ad, Bd, al, bl, alpha=1.0, small_tile=small)
    torch.cuda.synchronize()
    err = (out.cpu() - ref).abs().max().item()
    assert err < 1e-11 * max(1.0, K ** 0.5) * 10, err


@pytest.mark.parametrize("shape", [(128, 64, 64), (300, 70, 100), (1000, 500, 1600), (129, 65, 33)])
@pytest.mark.parametrize("slices", [8, 7])
def test_gemm_int8_slices(nat, shape, slices):
    """fp64-accurate GEMM on tcgen05.mma.kind::i8 (ozaki_i8.cu): error relative to sum_k |a||b| (the DGEMM bound)."""
    M, N, K = shape
    g = torch.Generator().manual_seed(M + N + K)
    A = torch.randn(M, K, generator=g, dtype=torch.float64)
    A *= torch.exp(4 * torch.randn(M, 1, generator=g, dtype=torch.float64))       # rows of very different scale
    B = torch.randn(N, K, generator=g, dtype=torch.float64)
    C = nat.gemm_i8emu(A.to(dev()), B.to(dev()), slices).cpu()
    bound = A.abs() @ B.abs().T
    tol = 4e-15 if slices == 8 else 4e-13
    assert float(((C - A @ B.T).abs() / bound).max()) < tol


@pytest.mark.parametrize("shape", [(300, 70, 50), (5000, 200, 130), (70001, 160, 100)])
def test_gemm_int8_slices_over_rows(nat, shape):
    """A^T B with the contraction over the rows (the X^T R shape): transposed slicing, split-K, ragged K."""
    K, M, N = shape
    g = torch.Generator().manual_seed(K)
    A = torch.randn(K, M, generator=g, dtype=torch.float64)
    B = torch.randn(K, N, generator=g, dtype=torch.float64) * torch.exp(3 * torch.randn(1, N, generator=g, dtype=torch.float64))
    C = nat.gemm_i8emu(A.to(dev()), B.to(dev()), 8, transposed=True).cpu()
    bound = A.abs().T @ B.abs()
    assert float(((C - A.T @ B).abs() / bound).max()) < 4e-15


def test_gemm_alpha_accumulate(nat):
    g = torch.Generator(device="cpu").manual_seed(5)
    A = torch.randn(200, 64, generator=g, dtype=torch.float64).to(dev())
    B = torch.randn(150, 64, generator=g, dtype=torch.float64).to(dev())
    C0 = torch.randn(200, 150, generator=g, dtype=torch.float64).to(dev())
    out = C0.clone()
    nat.gemm_f64(A, B, 0, 0, alpha=-1.0, out=out, accumulate=True, small_tile=True)
    assert (out - (C0 - A @ B.T)).abs().max().item() < 1e-11


@pytest.mark.parametrize("fused", [1024, 0])
@pytest.mark.parametrize("n", [1, 5, 64, 65, 100, 128, 257, 500, 513, 1000, 1024, 1100])
def test_cholesky_and_spd_inverse(nat, n, fused):
    """Option "chol_fused" = largest order that takes chol_fused.cu (one launch: a cluster of 8 CTAs; at most 1024, so 1100
    falls back; the default is 320); 0: always the launch-per-step path."""
    nat.set_option(0, "chol_fused", fused)
    try:
        _cholesky_and_spd_inverse(nat, n)
    finally:
        nat.set_option(0, "chol_fused", 320)


def _cholesky_and_spd_inverse(nat, n):
    g = torch.Generator(device="cpu").manual_seed(n)
    W = torch.randn(n, n + 3, generator=g, dtype=torch.float64)
    A = (W @ W.T + 0.5 * torch.eye(n, dtype=torch.float64))
    Ad = A.to(dev())
    L, fl = nat.cholesky(Ad)
    assert fl == 0
    Lref = torch.linalg.cholesky(A)
    assert (L.cpu() - Lref).abs().max().item() < 1e-10 * Lref.abs().max().item()
    Ainv, logdet, fl = nat.spd_inverse(Ad)
    assert fl == 0
    ref = torch.linalg.inv(A)
    assert (Ainv.cpu() - ref).abs().max().item() < 1e-9 * ref.abs().max().item()
    assert abs(logdet - torch.logdet(A).item()) < 1e-10 * max(1.0, abs(torch.logdet(A).item()))


@pytest.mark.parametrize("fused", [1024, 0])
def test_cholesky_flags_indefinite(nat, fused):
    nat.set_option(0, "chol_fused", fused)
    try:
        A = torch.eye(70, dtype=torch.float64)
        A[40, 40] = -1.0
        _, fl = nat.cholesky(A.to(dev()))
        assert fl & 1
        A = torch.eye(300, dtype=torch.float64)
        A[250, 250] = 1e-22                           # pivot 1e-11 < 1e-10: the reference's safe_log raises (U:670)
        _, fl = nat.cholesky(A.to(dev()))
        assert fl == 2
    finally:
        nat.set_option(0, "chol_fused", 320)


def test_fused_and_stepwise_factorisations_agree(nat):
    """Same blocked algorithm, same tile products: the factor, the inverse and log det agree to rounding (the trailing
    update sums in a different tile order), on a matrix shaped like K_tilde_b + G (E-step, utils.py:1430)."""
    g = torch.Generator(device="cpu").manual_seed(9)
    n = 500
    W = torch.randn(n, 2 * n, generator=g, dtype=torch.float64)
    A = (W @ W.T / n + torch.diag(torch.logspace(-3, 2, n, dtype=torch.float64))).to(dev())
    res = {}
    for fused in (1024, 0):
        nat.set_option(0, "chol_fused", fused)
        try:
            res[fused] = (nat.cholesky(A)[0], *nat.spd_inverse(A))
        finally:
            nat.set_option(0, "chol_fused", 320)
    L1, I1, ld1, f1 = res[1024]
    L0, I0, ld0, f0 = res[0]
    assert f1 == f0 == 0
    assert (torch.tril(L1) - torch.tril(L0)).abs().max().item() < 1e-12 * L0.abs().max().item()
    assert (I1 - I0).abs().max().item() < 1e-11 * I0.abs().max().item()
    assert abs(ld1 - ld0) < 1e-12 * abs(ld0)
    assert torch.equal(I1, I1.T)


@pytest.mark.parametrize("name", ["closure_dense_n8", "closure_local_n12"])
def test_localker_acosker_against_reference(nat, name):
    fx = load(name)
    prob, st = case_inputs(fx)
    n = prob.n_px_side
    C, mask, dC = nat.localker(prob.theta, n, dev(), grad=True)
    assert np.array_equal(mask.cpu().numpy(), fx["mask"])
    assert rel_err(C.cpu().numpy(), fx["C"]) < 1e-13
    for i, k in enumerate([str(s) for s in fx["dC_keys"]]):
        j = nat.DC_KEYS.index(k)
        assert rel_err(dC[j].cpu().numpy(), fx["dC"][i]) < 1e-12, k
    mk = mask.cpu()
    xm = prob.X[:, mk].to(dev())
    um = prob.xtilde[:, mk].to(dev())
    s0 = prob.theta["sigma_0"]
    K = nat.acosker(s0, xm, um, C, False)
    Kt = nat.acosker(s0, um, um, C, True)
    Kvec = nat.acosker_diag(s0, xm, C)
    assert rel_err(K.cpu().numpy(), fx["K"]) < 1e-11
    assert rel_err(Kt.cpu().numpy(), fx["K_tilde"]) < 1e-11
    assert rel_err(Kvec.cpu().numpy(), fx["Kvec"]) < 1e-12


def test_acosker_many_rows_odd_leading_dimension(nat):
    """More rows than one slab (65536) and more than gridDim.y allows, with an odd p so the operands go through the
    staging copies: standalone acosker / acosker_diag against the oracle (utils.py:986-1034)."""
    import oracle.spatial_gp_oracle as orc
    g = torch.Generator().manual_seed(11)
    n1, n2, p = 70001, 9, 15
    x1 = torch.randn(n1, p, generator=g, dtype=torch.float64)
    x2 = torch.randn(n2, p, generator=g, dtype=torch.float64)
    A = torch.randn(p, p, generator=g, dtype=torch.float64)
    C = A @ A.T / p + 0.1 * torch.eye(p, dtype=torch.float64)
    K = nat.acosker(0.7, x1.to(dev()), x2.to(dev()), C.to(dev()), False)
    Kvec = nat.acosker_diag(0.7, x1.to(dev()), C.to(dev()))
    assert rel_err(K.cpu().numpy(), orc.arccos_kernel({"sigma_0": 0.7}, x1, x2, C).numpy()) < 1e-11
    assert rel_err(Kvec.cpu().numpy(), (((x1 @ C) * x1).sum(1) + 0.49).numpy()) < 1e-12
    K2, Kvec2 = nat.acosker_kvec(0.7, x1.to(dev()), x2.to(dev()), C.to(dev()), False)
    assert torch.equal(K2, K) and rel_err(Kvec2.cpu().numpy(), Kvec.cpu().numpy()) < 1e-14


@pytest.mark.parametrize("n1,n2,p,sym_c", [(6000, 70, 100, True), (9000, 96, 131, False)])
def test_acosker_image_sized_rows_int8_path(nat, n1, n2, p, sym_c):
    """Standalone acosker / acosker_diag above the int8 threshold (x1 C with the row-dot epilogue and x1 (x2 C^T)^T on
    the tcgen05 int8-slice path), including a non-symmetric C, against the oracle (utils.py:978-990, 1029)."""
    import oracle.spatial_gp_oracle as orc
    g = torch.Generator().manual_seed(n1 + p)
    x1 = torch.randn(n1, p, generator=g, dtype=torch.float64)
    x2 = torch.randn(n2, p, generator=g, dtype=torch.float64)
    A = torch.randn(p, p, generator=g, dtype=torch.float64)
    C = A @ A.T / p + 0.1 * torch.eye(p, dtype=torch.float64)
    if not sym_c:
        C = C + 0.05 * torch.triu(torch.randn(p, p, generator=g, dtype=torch.float64), 1)
    K = nat.acosker(0.7, x1.to(dev()), x2.to(dev()), C.to(dev()), False)
    Kvec = nat.acosker_diag(0.7, x1.to(dev()), C.to(dev()))
    assert rel_err(K.cpu().numpy(), orc.arccos_kernel({"sigma_0": 0.7}, x1, x2, C).numpy()) < 1e-11
    assert rel_err(Kvec.cpu().numpy(), (((x1 @ C) * x1).sum(1) + 0.49).numpy()) < 1e-12
    K2, Kvec2 = nat.acosker_kvec(0.7, x1.to(dev()), x2.to(dev()), C.to(dev()), False)
    assert torch.equal(K2, K) and rel_err(Kvec2.cpu().numpy(), Kvec.cpu().numpy()) < 1e-14


@pytest.mark.parametrize("n1,n2,p", [(50, 1, 37), (3000, 1, 400), (9000, 3, 131), (70000, 2, 64)])
def test_acosker_known_norms_new_columns(nat, n1, n2, p):
    """sgp_acosker_known: new columns of K for images whose Kvec is at hand (the active-learning cycle's one new column):
    the oracle's acosker on the same inputs, and sgp_acosker itself; odd leading dimensions and several slabs included."""
    import oracle.spatial_gp_oracle as orc
    g = torch.Generator().manual_seed(n1 + p)
    x1 = torch.randn(n1, p, generator=g, dtype=torch.float64)
    x2 = torch.randn(n2, p, generator=g, dtype=torch.float64)
    A = torch.randn(p, p, generator=g, dtype=torch.float64)
    C = A @ A.T / p + 0.1 * torch.eye(p, dtype=torch.float64)
    d = dev()
    Kvec = nat.acosker_diag(0.7, x1.to(d), C.to(d))
    K = nat.acosker_known(0.7, x1.to(d), Kvec, x2.to(d), C.to(d))
    assert K.shape == (n1, n2)
    if n1 <= 9000:
        assert rel_err(K.cpu().numpy(), orc.arccos_kernel({"sigma_0": 0.7}, x1, x2, C).numpy()) < 1e-11
    Kfull = nat.acosker(0.7, x1.to(d), x2.to(d), C.to(d), False)
    assert rel_err(K.cpu().numpy(), Kfull.cpu().numpy()) < 1e-12
    with pytest.raises(ValueError):
        nat.acosker_known(0.7, x1.to(d), Kvec[:-1], x2.to(d), C.to(d))


@pytest.mark.parametrize("name", ["closure_dense_n8", "closure_local_n12", "closure_rankdef_n10", "closure_c1_dense"])
def test_moments_estep_fparams_against_reference(nat, name):
    import oracle.spatial_gp_oracle as orc
    fx = load(name)
    prob, st = case_inputs(fx)
    # rebuild a / K_b on the CPU with the oracle (inputs of the functions under test)
    C, mask = orc.build_C(prob.theta, prob.n_px_side)
    xm, um = prob.X[:, mask], prob.xtilde[:, mask]
    Kt = orc.arccos_kernel(prob.theta, um, um, C)
    K = orc.arccos_kernel(prob.theta, xm, um, C)
    B = st["B"]
    Kt_b = B.T @ Kt @ B
    Kt_b = 0.5 * (Kt_b + Kt_b.T)
    Kinv = torch.linalg.solve(Kt_b, torch.eye(Kt_b.shape[0], dtype=torch.float64))
    K_b = K @ B
    a = K_b @ Kinv
    Kvec = orc.arccos_kernel_diag(prob.theta, xm, C)
    d = dev()
    lam_m, lam_var = nat.lambda_moments(a.to(d), K_b.to(d), Kvec.to(d), st["m_b"].to(d), st["V_b"].to(d))
    assert np.allclose(lam_m.cpu().numpy(), fx["lam_m"], rtol=1e-9, atol=1e-12)
    assert np.allclose(lam_var.cpu().numpy(), fx["lam_var"], rtol=1e-9, atol=1e-10)
    res, f = nat.fparam_eval(torch.from_numpy(fx["lam_m"]).to(d), torch.from_numpy(fx["lam_var"]).to(d),
                             prob.r.to(d), st["logA"], st["lambda0"])
    assert np.allclose(f.cpu().numpy(), fx["f"], rtol=1e-12)
    assert abs(res["loglik"] - float(fx["fpar_loglik"])) <= 1e-11 * abs(float(fx["fpar_loglik"]))
    assert abs(res["dlogA"] - float(fx["fpar_dlogA"])) <= 1e-10 * abs(float(fx["fpar_dlogA"]))
    assert abs(res["dlambda0"] - float(fx["fpar_dlambda0"])) <= 1e-10 * max(1.0, abs(float(fx["fpar_dlambda0"])))
    assert abs(res["lambda0_closed"] - float(fx["lambda0_closed"])) < 1e-12
    g, G = nat.estep_terms(a.to(d), prob.r.to(d), torch.from_numpy(fx["f"]).to(d), st["logA"])
    assert rel_err(g.cpu().numpy(), fx["estep_g"]) < 1e-9
    assert rel_err(G.cpu().numpy(), fx["estep_G"]) < 1e-9


@pytest.mark.parametrize("N,q", [(6000, 96), (9001, 131), (20000, 500)])
def test_moments_and_estep_terms_many_images(nat, N, q):
    """Image counts above the int8 threshold: lambda_moments (a V^T with the row-sum epilogue) and estep_terms
    (a^T diag(f) a over lower tiles) on the tcgen05 int8-slice path, against torch fp64 (U:1090-1101, U:1421-1422)."""
    g = torch.Generator().manual_seed(N + q)
    a = torch.randn(N, q, generator=g, dtype=torch.float64) * torch.exp(torch.randn(N, 1, generator=g, dtype=torch.float64))
    K_b = torch.randn(N, q, generator=g, dtype=torch.float64)
    W = torch.randn(q, q, generator=g, dtype=torch.float64)
    V = W @ W.T / q + 0.1 * torch.eye(q, dtype=torch.float64)
    m_b = torch.randn(q, generator=g, dtype=torch.float64)
    Kvec = torch.rand(N, generator=g, dtype=torch.float64) + 1.0
    r = torch.poisson(torch.full((N,), 0.8, dtype=torch.float64), generator=g)
    f = torch.exp(0.3 * torch.randn(N, generator=g, dtype=torch.float64))
    d = dev()
    lam_m, lam_var = nat.lambda_moments(a.to(d), K_b.to(d), Kvec.to(d), m_b.to(d), V.to(d))
    ref_m = a @ m_b
    ref_v = Kvec + (-K_b * a + a * (a @ V.T)).sum(1)
    scale_v = Kvec + (K_b * a).abs().sum(1) + (a.abs() * (a.abs() @ V.abs().T)).sum(1)
    assert float(((lam_m.cpu() - ref_m).abs() / (a.abs() @ m_b.abs())).max()) < 1e-14
    assert float(((lam_var.cpu() - ref_v).abs() / scale_v).max()) < 1e-14
    logA = -0.7
    gv, G = nat.estep_terms(a.to(d), r.to(d), f.to(d), logA)
    A = float(np.exp(logA))
    G_ref = A * A * (a.T * f) @ a
    g_ref = A * (a.T @ (r - f))
    # the digit slices are relative to one scale per COLUMN of sqrt(f) a, so the error is bounded through the column
    # maxima (include/sgp_b200.h, sgp_gemm_i8emu), not through sum |a||a| f; rows here span ~3 decades on purpose
    cm = (a.abs() * f.sqrt()[:, None]).max(0).values
    err = (G.cpu() - G_ref).abs()
    assert float((err / (A * A * N * cm[:, None] * cm[None, :])).max()) < 2e-16
    assert float((err / (A * A * (a.abs().T * f) @ a.abs())).max()) < 1e-12
    assert float((G.cpu() - G.cpu().T).abs().max()) == 0.0
    assert rel_err(gv.cpu().numpy(), g_ref.numpy()) < 1e-12


def test_matmul_image_sized_goes_through_int8_slices(nat):
    """native.matmul (K_b = K B, a = K_b K_tilde_b^-1: U:1693-1694) at image-sized row counts."""
    g = torch.Generator().manual_seed(77)
    K = torch.randn(5000, 130, generator=g, dtype=torch.float64)
    B = torch.randn(130, 90, generator=g, dtype=torch.float64)
    C = nat.matmul(K.to(dev()), B.to(dev())).cpu()
    assert float(((C - K @ B).abs() / (K.abs() @ B.abs())).max()) < 4e-15
    C2 = nat.matmul_nt(K.to(dev()), B.T.contiguous().to(dev())).cpu()
    assert float(((C2 - K @ B).abs() / (K.abs() @ B.abs())).max()) < 4e-15


def test_nd_utility_against_reference(nat):
    """sgp_nd_utility (Lambert W on the device) against the reference's scipy-based nd_utility family, U:416-519."""
    fx = load("utility_n96")
    s2, mu, r = (torch.from_numpy(fx[k]).to(dev()) for k in ("sigma2", "mu", "r"))
    U, lam, keep, p, log_p = nat.nd_utility(s2, mu, r, want_arrays=True)
    assert np.array_equal(keep.cpu().numpy(), fx["keep"])
    assert np.allclose(lam.cpu().numpy(), fx["lam"], rtol=1e-9, atol=1e-12)
    assert np.allclose(log_p.cpu().numpy(), fx["log_p"], rtol=1e-9, atol=1e-11)
    assert np.allclose(p.cpu().numpy(), fx["p"], rtol=1e-9, atol=1e-300)
    assert np.allclose(U.cpu().numpy(), fx["U"], rtol=1e-9, atol=1e-12)
    assert np.array_equal(nat.nd_utility(s2, mu, r).cpu().numpy(), U.cpu().numpy())


def test_nd_utility_many_candidates(nat):
    """Every candidate is independent: a large batch equals the small batches it is made of (bitwise)."""
    g = torch.Generator().manual_seed(5)
    n = 200_003
    mu = (-1.0 + torch.randn(n, generator=g, dtype=torch.float64)).to(dev())
    s2 = torch.exp(-1.0 + torch.randn(n, generator=g, dtype=torch.float64)).to(dev())
    r = torch.arange(0, 100, dtype=torch.float64, device=dev())
    U = nat.nd_utility(s2, mu, r)
    assert torch.isfinite(U).all()
    assert torch.equal(U[1000:1064], nat.nd_utility(s2[1000:1064].clone(), mu[1000:1064].clone(), r))
