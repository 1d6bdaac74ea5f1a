"""The oracle restatement against (a) the reference's only known-answer vectors
(moments_gradients.ipynb cells 1-3) and (b) fixtures produced by the real reference
(oracle/gen_golden.py).  CPU only."""
import math

import numpy as np
import pytest
import torch

import oracle.spatial_gp_oracle as orc
from helpers import BIG_CASES, CLOSURE_CASES, case_inputs, load, rel_err

TOL = 1e-10  # oracle vs reference: same library, same formulas -> far below the 1e-9 product tolerance


def test_pi_is_float32_valued():
    assert orc.PI32 == float(np.float32(np.arccos(np.float32(0.0))) * 2)  # U:25


def test_kat_moments_gradients_notebook():
    # literals of moments_gradients.ipynb cell 1-3; expected values are the notebook's stored output
    a_mat = torch.tensor([[1.0, 2, 3, 4], [5, 6, 7, 8], [9, 10, 11, 12]], dtype=torch.float64)
    Sigma = torch.tensor([[1.0, 2, 3], [5, 6, 7], [8, 7, 6]], dtype=torch.float64); Sigma = Sigma + Sigma.T
    dSigma = torch.tensor([[4.0, 5, 6], [5, 6, 7], [6, 7, 8]], dtype=torch.float64); dSigma = dSigma + dSigma.T
    m = torch.tensor([1.0, 2, 3], dtype=torch.float64)
    dki = torch.tensor([[6.0, 7, 8, 9], [2, 3, 4, 5], [9, 8, 7, 6]], dtype=torch.float64)
    dkstar = torch.tensor([3.0, 4, 5, 6], dtype=torch.float64)
    V = torch.linalg.inv(Sigma * 3)
    Kinv = torch.linalg.pinv(Sigma)
    a, K = a_mat.T, dki.T
    _, _, dlm, dlv = orc.moments(a, K, dkstar, m, V, {"k": K}, {"k": dSigma}, {"k": dkstar}, Kinv)
    assert np.allclose(dlm["k"].numpy(), [13.75, 19.9167, 26.0833, 32.25], atol=5e-5)
    assert np.allclose(dlv["k"].numpy(), [-2199.1944, -3161.6944, -4291.75, -5589.3611], atol=5e-5)
    f = torch.tensor([55.0, 4, 22, 5], dtype=torch.float64)
    r = torch.tensor([23.0, 47, 2, 1], dtype=torch.float64)
    lam_m = torch.tensor([1.0, 2, 3, 4], dtype=torch.float64)
    assert orc.expected_loglik(r, f, lam_m, logA=0.0, lambda0=0.0) == 41.0
    dll = orc.expected_loglik_theta_grad(r, f, 0.0, dlm, dlv)["k"]
    assert abs(dll - 127749.63888888987) / 127749.63888888987 < 1e-12
    I10 = 10 * torch.eye(3, dtype=torch.float64)
    c = V @ Kinv
    b = Kinv @ m
    KLD = -0.5 * (orc.logdet_pd(V + I10) - orc.logdet_pd(Sigma + I10)) + 0.5 * float(m @ b) + 0.5 * float(torch.trace(c))
    assert abs(KLD - 34.91913283450331) < 1e-10
    # dKL through kl_term's gradient branch (log-dets are irrelevant for dKL)
    _, dKL = orc.kl_term(m, V + 0 * I10, Sigma, Kinv, {"k": dSigma})
    assert abs(dKL["k"] - 110.47222222222383) < 1e-9


@pytest.mark.parametrize("name", CLOSURE_CASES)
def test_closure_against_reference(name):
    fx = load(name)
    prob, st = case_inputs(fx)
    out = orc.elbo_and_grad(prob.theta, prob.X, prob.r, prob.xtilde, prob.n_px_side, st["B"], st["m_b"], st["V_b"],
                            st["logA"], st["lambda0"], return_moments=True)
    assert out["p"] == int(fx["p"])
    assert abs(out["loglik"] - float(fx["loglik"])) <= TOL * abs(float(fx["loglik"]))
    assert abs(out["KL"] - float(fx["KL"])) <= TOL * abs(float(fx["KL"]))
    keys = [str(k) for k in fx["theta_keys"]]
    dll = np.array([out["dloglik"][k] for k in keys])
    dkl = np.array([out["dKL"][k] for k in keys])
    # dloglik is identically zero (pure round-off) when ntilde == nt and q == m (a = B frozen), so the
    # gradient pieces are compared on the scale of the full gradient
    scale = max(np.abs(fx["dloglik"]).max(), np.abs(fx["dKL"]).max())
    assert np.abs(dll - fx["dloglik"]).max() < 1e-9 * scale
    assert np.abs(dkl - fx["dKL"]).max() < 1e-9 * scale
    assert rel_err(out["lam_m"].numpy(), fx["lam_m"]) < TOL
    assert rel_err(out["lam_var"].numpy(), fx["lam_var"]) < 1e-9
    assert rel_err(out["f"].numpy(), fx["f"]) < TOL


@pytest.mark.parametrize("name", BIG_CASES)
def test_closure_against_reference_benchmark_sizes(name):
    """The port at n = 40, m = 500 / 1000 against the real reference's outputs (chunked over the images, SURVEY 8c-iv:
    the same chunking bench.py's cpu_baseline leg and its in-run parity figure use)."""
    fx = load(name)
    prob, st = case_inputs(fx)
    full = orc.elbo_and_grad_chunked(prob.theta, prob.X, prob.r, prob.xtilde, prob.n_px_side, st["B"], st["m_b"],
                                     st["V_b"], st["logA"], st["lambda0"], chunk=4096)
    assert full["p"] == int(fx["p"])
    assert abs(full["loglik"] - float(fx["loglik"])) <= TOL * abs(float(fx["loglik"]))
    assert abs(full["KL"] - float(fx["KL"])) <= TOL * abs(float(fx["KL"]))
    keys = [str(k) for k in fx["theta_keys"]]
    for i, k in enumerate(keys):   # per component, on the cancellation scale of that component
        scale = max(abs(fx["dloglik"][i]), abs(fx["dKL"][i]))
        assert abs(full["dloglik"][k] - fx["dloglik"][i]) <= 1e-9 * scale, k
        assert abs(full["dKL"][k] - fx["dKL"][i]) <= 1e-9 * scale, k


@pytest.mark.parametrize("name", ["closure_dense_n8", "closure_local_n12"])
def test_kernel_pieces_against_reference(name):
    fx = load(name)
    prob, st = case_inputs(fx)
    C, mask, dC = orc.build_C(prob.theta, prob.n_px_side, grad=True)
    assert np.array_equal(mask.numpy(), fx["mask"])
    assert rel_err(C.numpy(), fx["C"]) < 1e-13
    for i, k in enumerate([str(s) for s in fx["dC_keys"]]):
        assert rel_err(dC[k].numpy(), fx["dC"][i]) < 1e-12, k
    xm, um = prob.X[:, mask], prob.xtilde[:, mask]
    Kt, dKt = orc.arccos_kernel(prob.theta, um, um, C, dC)
    K, dK = orc.arccos_kernel(prob.theta, xm, um, C, dC)
    Kvec, dKvec = orc.arccos_kernel_diag(prob.theta, xm, C, dC)
    assert rel_err(Kt.numpy(), fx["K_tilde"]) < 1e-12
    assert rel_err(K.numpy(), fx["K"]) < 1e-12
    assert rel_err(Kvec.numpy(), fx["Kvec"]) < 1e-12
    for i, k in enumerate([str(s) for s in fx["dK_keys"]]):
        assert rel_err(dK[k].numpy(), fx["dK"][i]) < 1e-10, k
        assert rel_err(dKt[k].numpy(), fx["dK_tilde"][i]) < 1e-10, k
        assert rel_err(dKvec[k].numpy(), fx["dKvec"][i]) < 1e-12, k


@pytest.mark.parametrize("name", CLOSURE_CASES)
def test_estep_and_fparams_against_reference(name):
    fx = load(name)
    prob, st = case_inputs(fx)
    out = orc.elbo_and_grad(prob.theta, prob.X, prob.r, prob.xtilde, prob.n_px_side, st["B"], st["m_b"], st["V_b"],
                            st["logA"], st["lambda0"], return_moments=True)
    # rebuild a / K_b as the closure does
    C, mask = orc.build_C(prob.theta, prob.n_px_side)
    xm, um = prob.X[:, mask], prob.xtilde[:, mask]
    Kt = orc.arccos_kernel(prob.theta, um, um, C)
    K = Kt if prob.N == prob.m else orc.arccos_kernel(prob.theta, xm, um, C)
    B = st["B"]
    Kt_b = B.T @ Kt @ B
    Kt_b = 0.5 * (Kt_b + Kt_b.T)
    Kinv = torch.linalg.solve(Kt_b, torch.eye(Kt_b.shape[0], dtype=torch.float64))
    K_b = K @ B
    a = B if prob.N == prob.m else K_b @ Kinv
    g, G = orc.estep_terms(a, prob.r, out["f"], st["logA"])
    assert rel_err(g.numpy(), fx["estep_g"]) < 1e-9
    assert rel_err(G.numpy(), fx["estep_G"]) < 1e-9
    m_new, V_new = orc.estep_update(a, prob.r, out["f"], st["m_b"], Kt_b, st["logA"])
    assert rel_err(m_new.numpy(), fx["estep_m"]) < 1e-9
    assert rel_err(V_new.numpy(), fx["estep_V"]) < 1e-9
    lam0 = orc.lambda0_closed_form(st["logA"], prob.r, out["lam_m"], out["lam_var"])
    assert abs(lam0 - float(fx["lambda0_closed"])) < 1e-12
    nll, ndlogA, _ = orc.fparam_closure(st["logA"], st["lambda0"], prob.r, out["lam_m"], out["lam_var"])
    assert abs(-nll - float(fx["fpar_loglik"])) <= 1e-11 * abs(float(fx["fpar_loglik"]))
    assert abs(-ndlogA - float(fx["fpar_dlogA"])) <= 1e-10 * abs(float(fx["fpar_dlogA"]))


def test_chunked_oracle_matches_full():
    fx = load("closure_c1_dense")
    prob, st = case_inputs(fx)
    full = orc.elbo_and_grad(prob.theta, prob.X, prob.r, prob.xtilde, prob.n_px_side, st["B"], st["m_b"], st["V_b"],
                             st["logA"], st["lambda0"])
    ch = orc.elbo_and_grad_chunked(prob.theta, prob.X, prob.r, prob.xtilde, prob.n_px_side, st["B"], st["m_b"],
                                   st["V_b"], st["logA"], st["lambda0"], chunk=700)
    assert abs(full["elbo"] - ch["elbo"]) < 1e-11 * abs(full["elbo"])
    for k in orc.KEYS:
        assert abs(full["grad"][k] - ch["grad"][k]) <= 1e-10 * max(abs(full["grad"][k]), 1e-3)


def test_utility_family_against_reference():
    """nd_lambda_r_mean / nd_p_r_given_xD / nd_utility (U:416-519), including the overflow-masked entries."""
    fx = load("utility_n96")
    s2, mu, r = (torch.from_numpy(fx[k]) for k in ("sigma2", "mu", "r"))
    lam, keep, p, log_p, r2d, lrf = orc.utility_terms(s2, mu, r)
    assert np.array_equal(keep.numpy(), fx["keep"]) and int((~keep).sum()) > 0
    assert np.allclose(lam.numpy(), fx["lam"], rtol=1e-13, atol=1e-13)
    assert np.allclose(log_p.numpy(), fx["log_p"], rtol=1e-12, atol=1e-12)
    assert np.allclose(p.numpy(), fx["p"], rtol=1e-11, atol=1e-300)
    assert np.array_equal(r2d.numpy(), fx["r2d"]) and np.allclose(lrf.numpy(), fx["lrf"], rtol=1e-15)
    assert np.allclose(orc.utility(s2, mu, r).numpy(), fx["U"], rtol=1e-10, atol=1e-12)
