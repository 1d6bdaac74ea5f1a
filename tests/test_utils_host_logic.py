"""CPU: host-side logic of the `utils`-compatible call surface (varGP's EM loop, closures, result layout,
error semantics) with the CUDA library replaced by the oracle-backed stand-in tests/fake_native.py.
Reference values: fixtures produced by the real reference (oracle/gen_golden.py)."""
import contextlib
import io
import math

import numpy as np
import pytest
import torch

import fake_native
from gaussian_processes_b200 import synth
from helpers import case_inputs, load


@pytest.fixture()
def U(monkeypatch):
    with contextlib.redirect_stdout(io.StringIO()):
        from gaussian_processes_b200 import utils
    monkeypatch.setattr(utils, "native", fake_native)
    return utils


def fit_args(fx, prob):
    n, m = int(fx["n"]), int(fx["m"])
    theta = synth.theta_tensors(prob.theta, requires_grad=True)
    return dict(
        fit_parameters=dict(ntilde=m, maxiter=int(fx["maxiter"]), nEstep=int(fx["nEstep"]), nMstep=int(fx["nMstep"]),
                            nFparamstep=int(fx["nFparamstep"]), kernfun="acosker", cellid=0, n_px_side=n,
                            display_hyper=False),
        xtilde=prob.xtilde, hyperparams_tuple=(theta, dict(synth.THETA_LIMS_LO), dict(synth.THETA_LIMS_HI)),
        f_params={"logA": torch.tensor(prob.logA), "lambda0": torch.tensor(prob.lambda0)})


def run_fit(U, name="fit_small_n10", **over):
    fx = load(name)
    keys = [str(k) for k in fx["theta_keys"]]
    prob = synth.make_problem(int(fx["N"]), int(fx["n"]), int(fx["m"]), cfg=int(fx["cfg"]),
                              theta=dict(zip(keys, fx["theta0"].tolist())))
    args = fit_args(fx, prob)
    args.update(over)
    with contextlib.redirect_stdout(io.StringIO()):
        fit, err = U.varGP(prob.X, prob.r, **args)
    return fx, keys, prob, args, fit, err


@pytest.mark.parametrize("name", ["fit_small_n10", "fit_c1_n20"])
def test_vargp_tracks_match_reference(U, name):
    """Loop-level parity: loss / theta / f-param tracks of the reference's own varGP (the small case and
    BASELINE config 1: N = 2000, 20 x 20, m = 100)."""
    fx, keys, prob, args, fit, err = run_fit(U, name)
    assert not err["is_error"], err
    vt = fit["values_track"]
    tol = 1e-6   # LBFGS line searches amplify round-off differences of the adjoint vs forward-mode gradients
    assert np.allclose(vt["loss_track"]["logmarginal"].numpy(), fx["logmarginal"], rtol=tol)
    assert np.allclose(vt["loss_track"]["loglikelihood"].numpy(), fx["loglikelihood"], rtol=tol)
    assert np.allclose(vt["loss_track"]["KL"].numpy(), fx["KL"], rtol=tol, atol=1e-8)
    got = np.stack([vt["theta_track"][k].numpy() for k in keys])
    assert np.allclose(got, fx["theta_track"], rtol=tol, atol=1e-8)
    assert np.allclose(vt["f_par_track"]["logA"].numpy(), fx["logA_track"], rtol=tol)
    assert np.allclose(vt["f_par_track"]["lambda0"].numpy(), fx["lambda0_track"], rtol=tol)
    assert fit["B"].shape[1] == int(fx["q"])
    assert np.allclose(fit["V_b"].numpy(), fx["V_b"], rtol=1e-5, atol=1e-8)


def test_vargp_result_layout_and_no_mutation(U):
    fx, keys, prob, args, fit, err = run_fit(U)
    assert set(fit) == {"fit_parameters", "final_kernel", "err_dict", "xtilde", "hyperparams_tuple", "f_params", "m_b",
                        "V_b", "C", "mask", "K_tilde_b", "K_tilde_inv_b", "K_b", "Kvec", "B", "values_track"}
    assert set(fit["final_kernel"]) == {"C", "mask", "K_tilde", "K", "Kvec", "eigvecs"}
    vt = fit["values_track"]
    assert set(vt) == {"loss_track", "theta_track", "f_par_track", "variation_par_track"}
    assert len(vt["variation_par_track"]["V_b"]) == int(fx["maxiter"])
    assert fit["fit_parameters"]["min_tolerance"] == 1e-11 and fit["fit_parameters"]["eigval_tol"] == 1e-4
    # caller's dicts are not mutated (U:1632, U:1653-1660)
    th0 = args["hyperparams_tuple"][0]
    assert all(float(th0[k]) == fx["theta0"][i] for i, k in enumerate(keys))
    assert float(args["f_params"]["logA"]) == float(fx["logA0"])
    assert "min_tolerance" not in args["fit_parameters"]
    # one fused call per M-step closure, none elsewhere
    assert fake_native.calls["elbo_grad"] > 0


def test_vargp_warm_start_tolerates_extra_keys(U):
    """The active loop feeds the previous fit_model back as kwargs (cell 17): extra keys must be ignored."""
    fx, keys, prob, args, fit, err = run_fit(U)
    again = dict(fit)
    again["fit_parameters"] = dict(fit["fit_parameters"], maxiter=2)
    with contextlib.redirect_stdout(io.StringIO()):
        fit2, err2 = U.varGP(prob.X, prob.r, **again)
    assert not err2["is_error"]
    assert fit2["values_track"]["loss_track"]["logmarginal"].shape[0] == 2


def test_vargp_error_rollback(U, monkeypatch):
    """An exception inside the loop rolls back to the last tracked iteration and is reported via err_dict."""
    real = fake_native.estep_terms
    n = {"c": 0}

    def flaky(*a, **k):
        n["c"] += 1
        if n["c"] > 7:      # fails inside iteration 3 (3 E-steps per iteration)
            raise RuntimeError("injected")
        return real(*a, **k)
    monkeypatch.setattr(fake_native, "estep_terms", flaky)
    fx, keys, prob, args, fit, err = run_fit(U)
    assert err["is_error"] and isinstance(err["error"], RuntimeError)
    assert fit["fit_parameters"]["maxiter"] == 3
    vt = fit["values_track"]
    assert vt["loss_track"]["logmarginal"].shape[0] == 3
    for i, k in enumerate(keys):     # theta restored to the value tracked at iteration maxiter-1
        assert float(fit["hyperparams_tuple"][0][k]) == float(vt["theta_track"][k][2])


def test_closure_out_of_bounds_returns_inf(U):
    fx = load("closure_dense_n8")
    prob, st = case_inputs(fx)
    th = synth.theta_tensors(prob.theta)
    th["eps_0x"] = torch.tensor(1.5)
    with pytest.raises(ValueError):
        U.localker(th, synth.THETA_LIMS_HI, synth.THETA_LIMS_LO, prob.n_px_side)
    with pytest.raises(ValueError):
        U.elbo_and_grad(th, synth.THETA_LIMS_LO, synth.THETA_LIMS_HI, prob.n_px_side, prob.X, prob.xtilde, prob.r,
                        st["B"], st["m_b"], st["V_b"], {"logA": torch.tensor(st["logA"]),
                                                        "lambda0": torch.tensor(st["lambda0"])})


@pytest.mark.parametrize("name", ["closure_square_n8", "closure_local_n12"])
def test_forward_mode_closure_matches_reference(U, name):
    """ntilde == nt (a = B) and the fallback path: the reference-sequenced closure on the function-level surface."""
    fx = load(name)
    prob, st = case_inputs(fx)
    th = synth.theta_tensors(prob.theta)
    fp = {"logA": torch.tensor(st["logA"]), "lambda0": torch.tensor(st["lambda0"])}
    same = prob.X.shape[0] == prob.xtilde.shape[0]
    ll, kl, g = U._closure_forward_mode(th, synth.THETA_LIMS_LO, synth.THETA_LIMS_HI, prob.n_px_side, prob.X,
                                        prob.xtilde, prob.r, U.acosker, same, st["B"], st["m_b"], st["V_b"], fp)
    assert abs(ll - float(fx["loglik"])) <= 1e-10 * abs(float(fx["loglik"]))
    assert abs(kl - float(fx["KL"])) <= 1e-10 * abs(float(fx["KL"]))
    keys = [str(k) for k in fx["theta_keys"]]
    ref = fx["dloglik"] - fx["dKL"]
    scale = max(np.abs(fx["dloglik"]).max(), np.abs(fx["dKL"]).max())
    assert np.abs(np.array([g[k] for k in keys]) - ref).max() <= 1e-9 * scale


def test_function_level_surface_against_reference(U):
    fx = load("closure_local_n12")
    prob, st = case_inputs(fx)
    th = synth.theta_tensors(prob.theta)
    C, mask, dC = U.localker(th, synth.THETA_LIMS_HI, synth.THETA_LIMS_LO, prob.n_px_side, grad=True)
    assert list(dC.keys()) == ["Amp", "-2log2beta", "-log2rho2", "eps_0x", "eps_0y"]
    xm, um = prob.X[:, mask], prob.xtilde[:, mask]
    K, dK = U.acosker(th, xm, um, C=C, dC=dC)
    assert list(dK.keys())[0] == "sigma_0"
    for i, k in enumerate([str(s) for s in fx["dK_keys"]]):
        assert np.allclose(dK[k].numpy(), fx["dK"][i], rtol=1e-9, atol=1e-12), k
    Kvec, dKvec = U.acosker(th, xm, C=C, dC=dC, diag=True)
    assert np.allclose(Kvec.numpy(), fx["Kvec"], rtol=1e-12)
    fp = {"logA": torch.tensor(st["logA"]), "lambda0": torch.tensor(st["lambda0"])}
    lam_m, lam_var = torch.from_numpy(fx["lam_m"]), torch.from_numpy(fx["lam_var"])
    f = U.mean_f_given_lambda_moments(fp, lam_m, lam_var)
    assert np.allclose(f.numpy(), fx["f"], rtol=1e-12)
    ll, d = U.compute_loglikelihood(prob.r, f, lam_m, lam_var, fp, compute_grad_for_f_params=True)
    assert abs(float(ll) - float(fx["fpar_loglik"])) <= 1e-11 * abs(float(fx["fpar_loglik"]))
    assert abs(float(d["logA"]) - float(fx["fpar_dlogA"])) <= 1e-10 * abs(float(fx["fpar_dlogA"]))
    ll3 = U.compute_loglikelihood(prob.r, f, lam_m, lam_var, fp)
    assert len(ll3) == 3 and float(ll3[2]) == float(prob.r.sum())
    assert abs(float(U.lambda0_given_logA(fp["logA"], prob.r, lam_m, lam_var)) - float(fx["lambda0_closed"])) < 1e-12
    # E-step against the reference's Estep output
    Kt_b, Kinv = torch.from_numpy(fx["K_tilde_b"]), torch.from_numpy(fx["K_tilde_inv_b"])
    Cm, mk = C, mask
    a = (U.acosker(th, xm, um, C=Cm) @ st["B"]) @ Kinv
    m_new, V_new = U.Estep(r=prob.r, KKtilde_inv=a, m=st["m_b"], f_params=fp, f_mean=torch.from_numpy(fx["f"]),
                           K_tilde=Kt_b, K_tilde_inv=Kinv)
    assert np.allclose(m_new.numpy(), fx["estep_m"], rtol=1e-8, atol=1e-11)
    assert np.allclose(V_new.numpy(), fx["estep_V"], rtol=1e-8, atol=1e-11)
    KL = U.compute_KL_div(st["m_b"], st["V_b"], Kt_b, Kinv)
    assert abs(float(KL) - float(fx["KL"])) <= 1e-10 * abs(float(fx["KL"]))


def test_log_det_fallback_semantics(U):
    M = torch.diag(torch.tensor([2.0, 1.0, -1.0]))
    with pytest.warns(UserWarning):
        v = U.log_det(M, name="M")
    assert abs(float(v) - math.log(2.0)) < 1e-12          # only eigenvalues > max(2e-4, 1e-4): {2, 1}
    assert U.log_det(M, ignore_warning=True) is not None
    A = torch.tensor([[1.0, 2.0], [0.0, 1.0]])
    with pytest.warns(UserWarning):
        assert U.log_det(A) == 0                           # not symmetric -> 0 with a warning (U:1302-1304)


def test_utility_family_surface(U):
    """The nd_* and single-candidate utility functions (U:416-618): shapes, masking semantics, values."""
    from helpers import load
    fx = load("utility_n96")
    s2, mu, r = (torch.from_numpy(fx[k]) for k in ("sigma2", "mu", "r"))
    lam, keep = U.nd_lambda_r_mean(r, s2, mu)
    assert lam.shape == (100, 96) and keep.dtype == torch.bool and np.array_equal(keep.numpy(), fx["keep"])
    p, log_p, r2d, lrf = U.nd_p_r_given_xD(r, s2, mu)
    assert np.allclose(log_p.numpy(), fx["log_p"], rtol=1e-12, atol=1e-12)
    assert np.array_equal(r2d.numpy(), fx["r2d"]) and np.allclose(lrf.numpy(), fx["lrf"], rtol=1e-15)
    Uv = U.nd_utility(s2, mu, r)
    assert np.allclose(Uv.numpy(), fx["U"], rtol=1e-10, atol=1e-12)
    assert np.allclose(U.nd_mean_noise_entropy(p, lrf, s2, mu).numpy(),
                       (-torch.exp(mu + 0.5 * s2) * (mu + s2 - 1) + (p * lrf).sum(0)).numpy())
    j = int(fx["single_index"])
    assert U.nd_utility(s2[j], mu[j], r).shape == (1,)            # 0-dim inputs are promoted (U:508-510)
    p1, log_p1, r1 = U.p_r_given_xD(r, s2[j], mu[j])
    assert np.allclose(p1.numpy(), fx["p1"], rtol=1e-11) and np.array_equal(r1.numpy(), fx["r1"])
    assert abs(float(U.utility(s2[j], mu[j], r)) - float(fx["U1"])) <= 1e-10 * abs(float(fx["U1"]))
    # the single-candidate variants DROP overflowing counts instead of zeroing them (U:551-553)
    lam0, keep0 = U.lambda_r_mean(r, s2[0], mu[0])
    assert lam0.numel() == int(keep0.sum()) < 100


# ---- test() / lambda_moments_star / explained_variance and the checkpoint format ----------
def test_checkpoint_written_by_the_reference_loads_and_text_matches(U, tmp_path):
    """load_model reads a model pickled by the reference's own save_model; save_model writes the same description /
    metadata text for it, and refuses an existing directory (U:46-109, U:312-324)."""
    from helpers import reference_saved_model_dir
    d = reference_saved_model_dir()
    model = U.load_model(d)
    assert {"fit_parameters", "final_kernel", "values_track", "m_b", "V_b", "B", "description"} <= set(model)
    ref_text = open(f"{d}/metadata").read()
    assert model["description"] == ref_text
    model.pop("description")
    out = tmp_path / "resaved"
    U.save_model(model, str(out), additional_description="written by the reference's save_model (oracle/gen_golden.py)")
    assert (out / "metadata").read_text() == ref_text
    back = U.load_model(str(out))
    assert torch.equal(back["V_b"], model["V_b"]) and back["description"] == ref_text
    with pytest.raises(ValueError):
        U.save_model(model, str(out))


def test_test_function_matches_reference(U):
    """utils.test as the notebooks call it (`test(X_test, R_test, X_train=X, at_iteration=None, **fit_model)`) on a
    model fitted AND pickled by the reference: predicted rates at 1e-9, the deterministic r2 at 1e-9."""
    from helpers import reference_saved_model_dir
    fx = load("test_small_n10")
    model = U.load_model(reference_saved_model_dir())
    X_test, R_test = torch.from_numpy(fx["X_test"]), torch.from_numpy(fx["R_test"])
    with contextlib.redirect_stdout(io.StringIO()):
        R_cell, R_pred, r2, s2 = U.test(X_test, R_test, X_train=None, at_iteration=None, **model)
    assert np.allclose(R_pred.numpy(), fx["R_pred"], rtol=1e-9, atol=0)
    assert torch.equal(R_cell, R_test[:, :, 0]) and math.isfinite(float(r2)) and math.isfinite(float(s2))
    r2_ns, none = U.explained_variance(R_test[:, :, 0], R_pred, sigma=False)
    assert none is None and abs(float(r2_ns) - float(fx["r2_nosigma"])) <= 1e-9 * abs(float(fx["r2_nosigma"]))
    theta = model["hyperparams_tuple"][0]
    mask = model["mask"]
    xs = X_test.reshape(X_test.shape[0], -1)
    mu, s2s = U.lambda_moments_star(xs[:7, mask], model["xtilde"][:, mask], model["C"], theta, model["K_tilde_b"],
                                    model["K_tilde_inv_b"], model["m_b"], model["V_b"], model["B"], "acosker")
    assert np.allclose(mu.numpy(), fx["mu_star"][:7], rtol=1e-9, atol=1e-12)
    assert np.allclose(s2s.numpy(), fx["sigma_star2"][:7], rtol=1e-9, atol=1e-12)


def test_warm_start_from_a_reloaded_checkpoint(U, tmp_path):
    """save_model -> load_model -> varGP(**model) with (m, V) lifted back as the active loop does (notebook cell 17):
    the warm-started fit begins exactly where the saved one ended."""
    fx, keys, prob, args, fit, err = run_fit(U)
    assert not err["is_error"]
    U.save_model(fit, str(tmp_path / "ckpt"))
    model = U.load_model(str(tmp_path / "ckpt"))
    B = model["B"]
    model["m"] = B @ model["m_b"]
    model["V"] = B @ model["V_b"] @ B.T
    model["fit_parameters"] = dict(model["fit_parameters"], maxiter=2, nEstep=1, nMstep=0)
    with contextlib.redirect_stdout(io.StringIO()):
        fit2, err2 = U.varGP(prob.X, prob.r, **model)      # extra keys of the result dict are tolerated
    assert not err2["is_error"], err2
    last = float(fit["values_track"]["loss_track"]["logmarginal"][-1])
    first = float(fit2["values_track"]["loss_track"]["logmarginal"][0])
    assert abs(first - last) <= 1e-8 * abs(last)


def test_plot_entry_points_exist(U):
    """plot_loss_and_theta_notebook / plot_fit exist with the reference's signatures (U:111, U:1543) and return None
    with a warning when matplotlib is absent (it is, in this image) instead of failing at `import utils`."""
    import warnings
    from helpers import reference_saved_model_dir
    model = U.load_model(reference_saved_model_dir())
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        fig = U.plot_loss_and_theta_notebook(model, linestyle='-', marker='o', figsize=(10, 10), xlim=None,
                                             ylim_logmarg=None, ylim_lambda0=None, ylim_eigvals=None)
        fig2 = U.plot_fit(torch.ones(5), torch.ones(3, 5), 0.5, 0.1, 0)
    try:
        import matplotlib  # noqa: F401
        assert fig is not None and fig2 is not None
    except ImportError:
        assert fig is None and fig2 is None


@pytest.mark.parametrize("name", ["active_same_n10", "active_fixedN_n10"])
def test_active_learning_cycles_match_reference(U, name):
    """three add-one-image cycles (K_tilde grown by one column, eigh, warm start through `m`, `V`,
    `init_kernel`) against the tracks of the REAL reference run through the same driver (oracle/gen_golden.py)."""
    from active_flow import run_active_fixture
    worst = run_active_fixture(U, name)
    assert max(worst.values()) <= 1e-6, worst
