"""GPU: the `utils`-compatible call surface end to end on the CUDA library (no stand-ins): loop-level parity of
varGP against the reference's own varGP tracks, the function-level surface against reference fixtures, the
ntilde == nt closure, and the no-CPU-fallback rule."""
import contextlib
import io

import numpy as np
import pytest
import torch

from gaussian_processes_b200 import synth
from helpers import case_inputs, load, record, rel_err

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


@pytest.mark.parametrize("name", ["fit_small_n10", "fit_c1_n20", "fit_n40_N8192_m200", "fit_c2_n40_N100000_m500"])
def test_vargp_tracks_match_reference_on_gpu(U, name):
    """The whole EM loop against the REAL reference's varGP: the small case, BASELINE config 1 itself (separable quadratic
    forms, cluster Cholesky), N = 8192 / 40 x 40 / m = 200, whose image-sized products run on the int8-slice engine, and
    BASELINE config 2 itself (N = 100 000, 40 x 40, m = 500: the size bench.py times; ten minutes of the reference on 8 host
    cores, oracle/gen_golden.py fitc2)."""
    from gaussian_processes_b200 import native
    fx = load(name)
    keys = [str(k) for k in fx["theta_keys"]]
    prob = synth.make_problem(int(fx["N"]), int(fx["n"]), int(fx["m"]), cfg=int(fx["cfg"]),
                              theta=dict(zip(keys, fx["theta0"].tolist())))
    native.launch_count(0, reset=True)
    with contextlib.redirect_stdout(io.StringIO()):
        fit, err = U.varGP(prob.X.to(D), prob.r.to(D), **fit_args(fx, prob, D))
    assert not err["is_error"], err
    assert native.launch_count(0) > 100          # the CUDA library did the work
    vt = fit["values_track"]
    got_t = np.stack([vt["theta_track"][k].cpu().numpy() for k in keys])
    record("vargp_tracks/" + name, logmarginal=rel_err(vt["loss_track"]["logmarginal"].cpu().numpy(), fx["logmarginal"]),
           loglik=rel_err(vt["loss_track"]["loglikelihood"].cpu().numpy(), fx["loglikelihood"]),
           KL=rel_err(vt["loss_track"]["KL"].cpu().numpy(), fx["KL"]), theta=rel_err(got_t, fx["theta_track"]),
           logA=rel_err(vt["f_par_track"]["logA"].cpu().numpy(), fx["logA_track"]))
    # 3 EM iterations x (3 E-steps with an LBFGS on logA + 3 LBFGS M-step iterations with strong-Wolfe line searches);
    # measured on the B200: 2e-13 (loss tracks) .. 4e-12 (theta tracks), profiles/r2_parity_report_v1.jsonl
    tol = 1e-9
    assert np.allclose(vt["loss_track"]["logmarginal"].cpu().numpy(), fx["logmarginal"], rtol=tol)
    assert np.allclose(vt["loss_track"]["loglikelihood"].cpu().numpy(), fx["loglikelihood"], rtol=tol)
    assert np.allclose(vt["loss_track"]["KL"].cpu().numpy(), fx["KL"], rtol=tol, atol=1e-10)
    got = np.stack([vt["theta_track"][k].cpu().numpy() for k in keys])
    assert np.allclose(got, fx["theta_track"], rtol=tol, atol=1e-10)
    assert np.allclose(vt["f_par_track"]["logA"].cpu().numpy(), fx["logA_track"], rtol=tol)
    assert np.allclose(vt["f_par_track"]["lambda0"].cpu().numpy(), fx["lambda0_track"], rtol=tol)
    assert fit["V_b"].is_cuda and fit["K_b"].is_cuda
    # eigh fixes eigenvector signs differently on cuSOLVER and LAPACK: V_b is compared up to the sign of each basis vector
    assert np.allclose(np.abs(fit["V_b"].cpu().numpy()), np.abs(fx["V_b"]), rtol=1e-5, atol=1e-8)
    assert np.allclose(np.linalg.eigvalsh(fit["V_b"].cpu().numpy()), np.linalg.eigvalsh(fx["V_b"]), rtol=1e-6)


@pytest.mark.parametrize("name", ["closure_square_n8", "closure_square_rankdef_n10", "closure_square_n20_m300",
                                  "closure_local_n12", "closure_c1_dense"])
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
    for i, k in enumerate(keys):       # per component, on its own cancellation scale
        assert abs(g[k] - ref[i]) <= 1e-9 * max(abs(fx["dloglik"][i]), abs(fx["dKL"][i])), k


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
    # (K^-1 + G)^-1 by blocked Cholesky instead of solve(I + K G, K) (U:1430): with the exact G the two routes agree
    # to 2e-15 on every fixture (cond(K^-1 + G) <= 6e3 even on the rank-deficient one), so the 1e-9
    # bar applies
    record("estep", m=rel_err(m_new.cpu().numpy(), fx["estep_m"]), V=rel_err(V_new.cpu().numpy(), fx["estep_V"]))
    assert np.allclose(m_new.cpu().numpy(), fx["estep_m"], rtol=1e-9, atol=1e-9 * np.abs(fx["estep_m"]).max())
    assert np.allclose(V_new.cpu().numpy(), fx["estep_V"], rtol=1e-9, atol=1e-9 * np.abs(fx["estep_V"]).max())
    KL, dKL = U.compute_KL_div(st["m_b"].to(D), st["V_b"].to(D), Kt_b, Kinv,
                               dK_tilde={k: B.T @ dKt[k] @ B for k in dKt})
    assert abs(float(KL) - float(fx["KL"])) <= 1e-9 * abs(float(fx["KL"]))
    keys = [str(k) for k in fx["theta_keys"]]
    record("dKL", err=rel_err(np.array([float(dKL[k]) for k in keys]), fx["dKL"]))
    assert np.allclose(np.array([float(dKL[k]) for k in keys]), fx["dKL"], rtol=1e-9, atol=1e-9 * np.abs(fx["dKL"]).max())


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


# ---- SURVEY 8(f) ranks 3 and 4 on the CUDA library -------------------------------------------------------------
def test_test_function_matches_reference_on_gpu(U):
    """utils.test / lambda_moments_star on a model fitted and pickled by the REFERENCE (moved to the GPU): all test
    images in one batch through sgp_acosker / sgp_acosker_diag; predicted rates and the deterministic r2 at 1e-9."""
    from helpers import reference_saved_model_dir, to_device
    fx = load("test_small_n10")
    model = to_device(U.load_model(reference_saved_model_dir()), D)
    X_test, R_test = torch.from_numpy(fx["X_test"]).to(D), torch.from_numpy(fx["R_test"]).to(D)
    with contextlib.redirect_stdout(io.StringIO()):
        R_cell, R_pred, r2, s2 = U.test(X_test, R_test, X_train=None, at_iteration=None, **model)
    assert R_pred.is_cuda
    record("test_fn", R_pred=rel_err(R_pred.cpu().numpy(), fx["R_pred"]))
    assert np.allclose(R_pred.cpu().numpy(), fx["R_pred"], rtol=1e-9, atol=0)
    assert np.isfinite(float(r2)) and np.isfinite(float(s2))
    r2_ns, _ = U.explained_variance(R_test[:, :, 0], R_pred, sigma=False)
    assert abs(float(r2_ns) - float(fx["r2_nosigma"])) <= 1e-9 * abs(float(fx["r2_nosigma"]))
    theta, mask = model["hyperparams_tuple"][0], model["mask"]
    xs = X_test.reshape(X_test.shape[0], -1)
    mu, s2s = U.lambda_moments_star(xs[:, mask], model["xtilde"][:, mask], model["C"], theta, model["K_tilde_b"],
                                    model["K_tilde_inv_b"], model["m_b"], model["V_b"], model["B"], "acosker")
    assert np.allclose(mu.cpu().numpy(), fx["mu_star"], rtol=1e-9, atol=1e-12)
    assert np.allclose(s2s.cpu().numpy(), fx["sigma_star2"], rtol=1e-9, atol=1e-12)


def test_test_function_at_iteration_on_gpu(U):
    """test(at_iteration=k, X_train=X) rebuilds C / K_tilde / the subspace at the tracked theta of iteration k
    (U:358-386); at the last iteration it must reproduce the plain call on the same fit."""
    fx = load("fit_small_n10")
    keys = [str(k) for k in fx["theta_keys"]]
    prob = synth.make_problem(int(fx["N"]), int(fx["n"]), int(fx["m"]), cfg=int(fx["cfg"]),
                              theta=dict(zip(keys, fx["theta0"].tolist())))
    tx = load("test_small_n10")
    X_test, R_test = torch.from_numpy(tx["X_test"]).to(D), torch.from_numpy(tx["R_test"]).to(D)
    with contextlib.redirect_stdout(io.StringIO()):
        fit, err = U.varGP(prob.X.to(D), prob.r.to(D), **fit_args(fx, prob, D))
        _, plain, _, _ = U.test(X_test, R_test, X_train=prob.X.to(D), at_iteration=None, **fit)
        last = int(fit["fit_parameters"]["maxiter"]) - 1
        _, at_last, _, _ = U.test(X_test, R_test, X_train=prob.X.to(D), at_iteration=last, **fit)
    assert not err["is_error"]
    assert np.allclose(at_last.cpu().numpy(), plain.cpu().numpy(), rtol=1e-8, atol=0)
    # and the GPU fit predicts what the reference's fit predicts (tracks agree to ~1e-8, see the track test)
    assert np.allclose(plain.cpu().numpy(), tx["R_pred"], rtol=1e-6, atol=0)


def test_checkpoint_round_trip_and_warm_start_on_gpu(U, tmp_path):
    """save_model -> load_model -> varGP(**model) on the GPU: a CUDA model pickles and reloads onto the device, the
    warm start (m, V lifted with B as notebook cell 17 does) begins at the loss the saved fit ended at; a checkpoint
    written by the reference's save_model warm-starts the same way."""
    from helpers import reference_saved_model_dir, to_device
    fx = load("fit_small_n10")
    keys = [str(k) for k in fx["theta_keys"]]
    prob = synth.make_problem(int(fx["N"]), int(fx["n"]), int(fx["m"]), cfg=int(fx["cfg"]),
                              theta=dict(zip(keys, fx["theta0"].tolist())))
    X, r = prob.X.to(D), prob.r.to(D)
    with contextlib.redirect_stdout(io.StringIO()):
        fit, err = U.varGP(X, r, **fit_args(fx, prob, D))
    assert not err["is_error"]
    U.save_model(fit, str(tmp_path / "ckpt"))
    for src in (str(tmp_path / "ckpt"), reference_saved_model_dir()):
        model = to_device(U.load_model(src), D)
        B = model["B"]
        model["m"] = B @ model["m_b"]
        model["V"] = B @ model["V_b"] @ B.T
        last = float(model["values_track"]["loss_track"]["logmarginal"][-1])
        model["fit_parameters"] = dict(model["fit_parameters"], maxiter=2, nEstep=1, nMstep=0)
        with contextlib.redirect_stdout(io.StringIO()):
            fit2, err2 = U.varGP(X, r, **model)
        assert not err2["is_error"], err2
        first = float(fit2["values_track"]["loss_track"]["logmarginal"][0])
        record("warm_start", rel=abs(first - last) / abs(last))
        assert abs(first - last) <= 1e-8 * abs(last)


def test_plot_entry_points_exist(U):
    """The notebooks call utils.plot_loss_and_theta_notebook after every fit (11 call sites): the names exist and
    degrade to a warning + None on a box without matplotlib instead of an AttributeError."""
    import warnings
    model = U.load_model(__import__("helpers").reference_saved_model_dir())
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        fig = U.plot_loss_and_theta_notebook(model, xlim=None, ylim_logmarg=None)
        fig2 = U.plot_fit(torch.ones(5), torch.ones(3, 5), 0.5, 0.1, 0)
    try:
        import matplotlib  # noqa: F401
        assert fig is not None and fig2 is not None
    except ImportError:
        assert fig is None and fig2 is None


@pytest.mark.parametrize("name", ["active_same_n10", "active_fixedN_n10", "active_fixedN_n20_N5000", "active_same_n20_m150"])
def test_active_learning_cycles_match_reference_on_gpu(U, name):
    """The active-learning cycle on the CUDA library: incremental K_tilde / K columns (sgp_acosker on one image; sgp_acosker_known for the column of K),
    eigh, warm start through init_kernel; ntilde == nt runs the fused a = B closure.  Tracks of every warm-started fit
    against the REAL reference's (driven through the same gaussian_processes_b200.active.grow_model)."""
    from active_flow import run_active_fixture
    from gaussian_processes_b200 import native
    native.launch_count(0, reset=True)
    worst = run_active_fixture(U, name, dev=D)
    assert native.launch_count(0) > 100
    record("active/" + name, **worst)
    assert max(worst.values()) <= 1e-9, worst      # measured 2e-15 (ntilde == nt) and 5e-12 (fixed N)
