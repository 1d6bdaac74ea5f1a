"""GPU: the `utils`-compatible call surface end to end on the CUDA library (no stand-ins): loop-level parity of
varGP against the reference's own varGP tracks, the function-level surface against reference fixtures, the
ntilde == nt closure, and the no-CPU-fallback rule."""
import contextlib
import io

import numpy as np
import pytest
import torch

from gaussian_processes_b200 import synth
from helpers import case_inputs, load

pytestmark = pytest.mark.gpu
D = "cuda:0"


@pytest.fixture(scope="module")
def U():
    with contextlib.redirect_stdout(io.StringIO()):
        from gaussian_processes_b200 import utils
    return utils


def theta_on(theta, dev, requires_grad=False):
    return synth.theta_tensors(theta, device=dev, requires_grad=requires_grad)


def fit_args(fx, prob, dev):
    n, m = int(fx["n"]), int(fx["m"])
    return dict(
        fit_parameters=dict(ntilde=m, maxiter=int(fx["maxiter"]), nEstep=int(fx["nEstep"]), nMstep=int(fx["nMstep"]),
                            nFparamstep=int(fx["nFparamstep"]), kernfun="acosker", cellid=0, n_px_side=n,
                            display_hyper=False),
        xtilde=prob.xtilde.to(dev),
        hyperparams_tuple=(theta_on(prob.theta, dev, True), dict(synth.THETA_LIMS_LO), dict(synth.THETA_LIMS_HI)),
        f_params={"logA": torch.tensor(prob.logA, device=dev), "lambda0": torch.tensor(prob.lambda0, device=dev)})


def test_vargp_tracks_match_reference_on_gpu(U):
    from gaussian_processes_b200 import native
    fx = load("fit_small_n10")
    keys = [str(k) for k in fx["theta_keys"]]
    prob = synth.make_problem(int(fx["N"]), int(fx["n"]), int(fx["m"]), cfg=int(fx["cfg"]),
                              theta=dict(zip(keys, fx["theta0"].tolist())))
    native.launch_count(0, reset=True)
    with contextlib.redirect_stdout(io.StringIO()):
        fit, err = U.varGP(prob.X.to(D), prob.r.to(D), **fit_args(fx, prob, D))
    assert not err["is_error"], err
    assert native.launch_count(0) > 100          # the CUDA library did the work
    vt = fit["values_track"]
    tol = 1e-6
    assert np.allclose(vt["loss_track"]["logmarginal"].cpu().numpy(), fx["logmarginal"], rtol=tol)
    assert np.allclose(vt["loss_track"]["loglikelihood"].cpu().numpy(), fx["loglikelihood"], rtol=tol)
    assert np.allclose(vt["loss_track"]["KL"].cpu().numpy(), fx["KL"], rtol=tol, atol=1e-8)
    got = np.stack([vt["theta_track"][k].cpu().numpy() for k in keys])
    assert np.allclose(got, fx["theta_track"], rtol=tol, atol=1e-8)
    assert np.allclose(vt["f_par_track"]["logA"].cpu().numpy(), fx["logA_track"], rtol=tol)
    assert np.allclose(vt["f_par_track"]["lambda0"].cpu().numpy(), fx["lambda0_track"], rtol=tol)
    assert fit["V_b"].is_cuda and fit["K_b"].is_cuda
    # eigh fixes eigenvector signs differently on cuSOLVER and LAPACK: V_b is compared up to the sign of each basis vector
    assert np.allclose(np.abs(fit["V_b"].cpu().numpy()), np.abs(fx["V_b"]), rtol=1e-5, atol=1e-8)
    assert np.allclose(np.linalg.eigvalsh(fit["V_b"].cpu().numpy()), np.linalg.eigvalsh(fx["V_b"]), rtol=1e-6)


@pytest.mark.parametrize("name", ["closure_square_n8", "closure_local_n12", "closure_c1_dense"])
def test_elbo_and_grad_surface(U, name):
    """utils.elbo_and_grad: fused reverse-mode call when ntilde != nt, reference-sequenced closure when ntilde == nt."""
    fx = load(name)
    prob, st = case_inputs(fx)
    th = theta_on(prob.theta, D)
    fp = {"logA": torch.tensor(st["logA"], device=D), "lambda0": torch.tensor(st["lambda0"], device=D)}
    ll, kl, g = U.elbo_and_grad(th, synth.THETA_LIMS_LO, synth.THETA_LIMS_HI, prob.n_px_side, prob.X.to(D),
                                prob.xtilde.to(D), prob.r.to(D), st["B"].to(D), st["m_b"].to(D), st["V_b"].to(D), fp)
    assert abs(ll - float(fx["loglik"])) <= 1e-9 * abs(float(fx["loglik"]))
    assert abs(kl - float(fx["KL"])) <= 1e-9 * abs(float(fx["KL"]))
    keys = [str(k) for k in fx["theta_keys"]]
    ref = fx["dloglik"] - fx["dKL"]
    scale = max(np.abs(fx["dloglik"]).max(), np.abs(fx["dKL"]).max())
    assert np.abs(np.array([g[k] for k in keys]) - ref).max() <= 1e-9 * scale


def test_function_level_surface_on_gpu(U):
    fx = load("closure_local_n12")
    prob, st = case_inputs(fx)
    th = theta_on(prob.theta, D)
    C, mask, dC = U.localker(th, synth.THETA_LIMS_HI, synth.THETA_LIMS_LO, prob.n_px_side, grad=True)
    assert C.is_cuda and mask.is_cuda and list(dC.keys()) == ["Amp", "-2log2beta", "-log2rho2", "eps_0x", "eps_0y"]
    X, Ut, r = prob.X.to(D), prob.xtilde.to(D), prob.r.to(D)
    xm, um = X[:, mask], Ut[:, mask]
    K, dK = U.acosker(th, xm, um, C=C, dC=dC)
    assert np.allclose(K.cpu().numpy(), fx["K"], rtol=1e-10, atol=1e-13)
    for i, k in enumerate([str(s) for s in fx["dK_keys"]]):
        assert np.allclose(dK[k].cpu().numpy(), fx["dK"][i], rtol=1e-9, atol=1e-11), k
    Kt, dKt = U.acosker(th, um, um, C=C, dC=dC)
    assert np.allclose(Kt.cpu().numpy(), fx["K_tilde"], rtol=1e-10, atol=1e-13)
    Kvec, dKvec = U.acosker(th, xm, C=C, dC=dC, diag=True)
    assert np.allclose(Kvec.cpu().numpy(), fx["Kvec"], rtol=1e-12)
    for i, k in enumerate([str(s) for s in fx["dK_keys"]]):
        assert np.allclose(dKvec[k].cpu().numpy(), fx["dKvec"][i], rtol=1e-9, atol=1e-11), k
    fp = {"logA": torch.tensor(st["logA"], device=D), "lambda0": torch.tensor(st["lambda0"], device=D)}
    lam_m, lam_var = torch.from_numpy(fx["lam_m"]).to(D), torch.from_numpy(fx["lam_var"]).to(D)
    f = U.mean_f_given_lambda_moments(fp, lam_m, lam_var)
    assert np.allclose(f.cpu().numpy(), fx["f"], rtol=1e-12)
    ll, d = U.compute_loglikelihood(r, f, lam_m, lam_var, fp, compute_grad_for_f_params=True)
    assert abs(float(ll) - float(fx["fpar_loglik"])) <= 1e-11 * abs(float(fx["fpar_loglik"]))
    assert abs(float(d["logA"]) - float(fx["fpar_dlogA"])) <= 1e-10 * abs(float(fx["fpar_dlogA"]))
    assert abs(float(d["lambda0"]) - float(fx["fpar_dlambda0"])) <= 1e-10 * max(1.0, abs(float(fx["fpar_dlambda0"])))
    assert abs(float(U.lambda0_given_logA(fp["logA"], r, lam_m, lam_var)) - float(fx["lambda0_closed"])) < 1e-12
    Kt_b, Kinv = torch.from_numpy(fx["K_tilde_b"]).to(D), torch.from_numpy(fx["K_tilde_inv_b"]).to(D)
    B = st["B"].to(D)
    K_b = K @ B
    a = K_b @ Kinv
    lm, lv = U.lambda_moments(xm, Kt_b, a, Kvec, K_b, C, st["m_b"].to(D), st["V_b"].to(D), th)
    assert np.allclose(lm.cpu().numpy(), fx["lam_m"], rtol=1e-9, atol=1e-12)
    assert np.allclose(lv.cpu().numpy(), fx["lam_var"], rtol=1e-9, atol=1e-10)
    m_new, V_new = U.Estep(r=r, KKtilde_inv=a, m=st["m_b"].to(D), f_params=fp, f_mean=torch.from_numpy(fx["f"]).to(D),
                           K_tilde=Kt_b, K_tilde_inv=Kinv)
    assert np.allclose(m_new.cpu().numpy(), fx["estep_m"], rtol=1e-8, atol=1e-11)
    assert np.allclose(V_new.cpu().numpy(), fx["estep_V"], rtol=1e-8, atol=1e-11)
    KL, dKL = U.compute_KL_div(st["m_b"].to(D), st["V_b"].to(D), Kt_b, Kinv,
                               dK_tilde={k: B.T @ dKt[k] @ B for k in dKt})
    assert abs(float(KL) - float(fx["KL"])) <= 1e-9 * abs(float(fx["KL"]))
    keys = [str(k) for k in fx["theta_keys"]]
    assert np.allclose(np.array([float(dKL[k]) for k in keys]), fx["dKL"], rtol=1e-8, atol=1e-10)


def test_cpu_tensors_raise(U):
    fx = load("closure_dense_n8")
    prob, st = case_inputs(fx)
    th = synth.theta_tensors(prob.theta)
    fp = {"logA": torch.tensor(st["logA"]), "lambda0": torch.tensor(st["lambda0"])}
    with pytest.raises((RuntimeError, TypeError)):
        U.elbo_and_grad(th, synth.THETA_LIMS_LO, synth.THETA_LIMS_HI, prob.n_px_side, prob.X, prob.xtilde, prob.r,
                        st["B"], st["m_b"], st["V_b"], fp)


def test_utility_family_on_gpu(U):
    fx = load("utility_n96")
    s2, mu, r = (torch.from_numpy(fx[k]).to(D) for k in ("sigma2", "mu", "r"))
    Uv = U.nd_utility(s2, mu, r)
    assert Uv.is_cuda and np.allclose(Uv.cpu().numpy(), fx["U"], rtol=1e-9, atol=1e-12)
    p, log_p, r2d, lrf = U.nd_p_r_given_xD(r, s2, mu)
    assert np.allclose(log_p.cpu().numpy(), fx["log_p"], rtol=1e-9, atol=1e-11)
    assert np.array_equal(r2d.cpu().numpy(), fx["r2d"])
    j = int(fx["single_index"])
    assert abs(float(U.utility(s2[j], mu[j], r)) - float(fx["U1"])) <= 1e-9 * abs(float(fx["U1"]))
