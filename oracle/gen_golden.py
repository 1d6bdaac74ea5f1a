"""Generate tests/golden/*.npz by running the REAL reference (``/root/reference``) on CPU.

Run only in the build container (the reference does not travel to the GPU box):

    python oracle/gen_golden.py

The reference module needs matplotlib at import time (U:21-22) which is not installed; three
stub modules are injected before ``import utils``.  The stored fixtures hold the
*outputs of the reference's own functions* on seeded synthetic inputs (inputs are regenerated
from the seeds by ``gaussian_processes_b200.synth``; only what cannot be regenerated is stored).

The body of ``closure_hyperparams`` (U:2030-2099) is a nested function and cannot be imported;
``ref_closure`` below re-states its *sequencing* around calls into the reference's own
localker / acosker / mean_f / compute_loglikelihood / compute_KL_div.
"""
from __future__ import annotations

import io
import contextlib
import math
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
REF = "/root/reference/Spatial_GP_repo"


def import_reference():
    for name in ("matplotlib", "matplotlib.pyplot", "matplotlib.ticker"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["matplotlib.ticker"].FuncFormatter = object
    sys.modules["matplotlib.ticker"].MaxNLocator = object
    sys.path.insert(0, REF)
    with contextlib.redirect_stdout(io.StringIO()):
        import utils as ref  # noqa
    torch.set_default_device("cpu")
    return ref


from gaussian_processes_b200 import synth  # noqa: E402


def ref_closure(ref, theta, X, r, xtilde, n_px, B, m_b, V_b, f_params):
    """Sequencing of U:2030-2099 around the reference's own functions."""
    lo, hi = synth.THETA_LIMS_LO, synth.THETA_LIMS_HI
    nt, ntilde = X.shape[0], xtilde.shape[0]
    C, mask, dC = ref.localker(theta=theta, theta_higher_lims=hi, theta_lower_lims=lo, n_px_side=n_px, grad=True)
    K_tilde, dK_tilde = ref.acosker(theta, xtilde[:, mask], xtilde[:, mask], C=C, dC=dC, diag=False)
    K, dK = ref.acosker(theta, X[:, mask], xtilde[:, mask], C=C, dC=dC, diag=False) if ntilde != nt else (K_tilde, dK_tilde)
    Kvec, dKvec = ref.acosker(theta, X[:, mask], x2=None, C=C, dC=dC, diag=True)
    K_tilde_b = B.T @ K_tilde @ B
    K_tilde_b = (K_tilde_b + K_tilde_b.T) * 0.5
    K_b = K @ B
    dK_tilde_b = {k: B.T @ dK_tilde[k] @ B for k in dK_tilde}
    dK_b = {k: dK[k] @ B for k in dK}
    K_tilde_inv_b = torch.linalg.solve(K_tilde_b, torch.eye(K_tilde_b.shape[0]))
    KKtilde_inv_b = K_b @ K_tilde_inv_b if ntilde != nt else B
    f_mean, lambda_m, lambda_var, dlambda_m, dlambda_var = ref.mean_f(
        f_params=f_params, calculate_moments=True, x=X[:, mask], K_tilde=K_tilde_b, KKtilde_inv=KKtilde_inv_b,
        Kvec=Kvec, K=K_b, C=C, m=m_b, V=V_b, theta=theta, kernfun=ref.acosker, lambda_m=None, lambda_var=None,
        dK=dK_b, dK_tilde=dK_tilde_b, dK_vec=dKvec, K_tilde_inv=K_tilde_inv_b)
    loglik, dloglik = ref.compute_loglikelihood(X.new_tensor(r) if not torch.is_tensor(r) else r, f_mean, lambda_m,
                                                lambda_var, f_params, dlambda_m=dlambda_m, dlambda_var=dlambda_var)
    KL, dKL = ref.compute_KL_div(m_b, V_b, K_tilde_b, K_tilde_inv=K_tilde_inv_b, dK_tilde=dK_tilde_b)
    keys = list(theta.keys())
    return dict(C=C, mask=mask, K_tilde=K_tilde, K=K, Kvec=Kvec, K_tilde_b=K_tilde_b, K_tilde_inv_b=K_tilde_inv_b,
                a=KKtilde_inv_b, K_b=K_b, lam_m=lambda_m, lam_var=lambda_var, f=f_mean, loglik=loglik, KL=KL,
                dloglik=torch.stack([dloglik[k] for k in keys]), dKL=torch.stack([dKL[k] for k in keys]),
                dC=dC, dK_tilde=dK_tilde, dK=dK, dKvec=dKvec)


def setup_case(ref, N, n, m, cfg, theta_dict):
    prob = synth.make_problem(N, n, m, cfg=cfg, theta=theta_dict)
    theta = synth.theta_tensors(prob.theta)
    lo, hi = synth.THETA_LIMS_LO, synth.THETA_LIMS_HI
    C, mask = ref.localker(theta=theta, theta_higher_lims=hi, theta_lower_lims=lo, n_px_side=n, grad=False)
    Kt = ref.acosker(theta, prob.xtilde[:, mask], prob.xtilde[:, mask], C=C, dC=None, diag=False)
    ev, evec = torch.linalg.eigh(Kt, UPLO="L")
    keep = ev > max(ev.max() * ref.EIGVAL_TOL, ref.EIGVAL_TOL)
    B = evec[:, keep]
    Kt_b = torch.diag(ev[keep])
    m_b, V_b = synth.make_variational_state(Kt_b, cfg)
    return prob, theta, B, m_b, V_b, ev


def gen_closure_case(ref, name, N, n, m, cfg, theta_dict, store_big=False, perturb=None):
    prob, theta, B, m_b, V_b, ev = setup_case(ref, N, n, m, cfg, theta_dict)
    if perturb:  # evaluate the closure at a theta different from the one B was built at (frozen B, U:2036-2049)
        theta = synth.theta_tensors({k: prob.theta[k] + perturb.get(k, 0.0) for k in prob.theta})
    f_params = {"logA": torch.tensor(prob.logA), "lambda0": torch.tensor(prob.lambda0)}
    out = ref_closure(ref, theta, prob.X, prob.r, prob.xtilde, n, B, m_b, V_b, f_params)
    # E-step pieces at the same state (a10, a8, a7 f-param mode), with the *closure's* a / K_b
    a, K_b = out["a"], out["K_b"]
    m_new, V_new = ref.Estep(r=prob.r, KKtilde_inv=a, m=m_b, f_params=f_params, f_mean=out["f"],
                             K_tilde=out["K_tilde_b"], K_tilde_inv=out["K_tilde_inv_b"], update_V_inv=False, alpha=1)
    A = torch.exp(f_params["logA"])
    g = A * a.T @ (prob.r - out["f"])
    G = A * A * a.T @ (a * out["f"][:, None])
    lam0 = ref.lambda0_given_logA(f_params["logA"], prob.r, out["lam_m"], out["lam_var"])
    ll_f, dll_f = ref.compute_loglikelihood(prob.r, out["f"], out["lam_m"], out["lam_var"], f_params,
                                            compute_grad_for_f_params=True)
    lam_m2, lam_var2 = ref.lambda_moments(None, out["K_tilde_b"], a, out["Kvec"], K_b, None, m_new, V_new, theta)
    keys = list(theta.keys())
    fx = dict(
        N=N, n=n, m=m, cfg=cfg, theta=np.array([float(theta[k]) for k in keys]), theta_keys=np.array(keys),
        theta_B=np.array([prob.theta[k] for k in keys]),
        q=B.shape[1], p=int(out["mask"].sum()), mask=out["mask"].numpy(),
        B=B.numpy(), m_b=m_b.numpy(), V_b=V_b.numpy(), logA=prob.logA, lambda0=prob.lambda0,
        loglik=float(out["loglik"]), KL=float(out["KL"]), dloglik=out["dloglik"].numpy(), dKL=out["dKL"].numpy(),
        lam_m=out["lam_m"].numpy(), lam_var=out["lam_var"].numpy(), f=out["f"].numpy(), Kvec=out["Kvec"].numpy(),
        K_tilde_b=out["K_tilde_b"].numpy(), K_tilde_inv_b=out["K_tilde_inv_b"].numpy(),
        estep_g=g.numpy(), estep_G=G.numpy(), estep_m=m_new.numpy(), estep_V=V_new.numpy(),
        lam_m_after=lam_m2.numpy(), lam_var_after=lam_var2.numpy(),
        lambda0_closed=float(lam0), fpar_loglik=float(ll_f), fpar_dlogA=float(dll_f["logA"]),
        fpar_dlambda0=float(dll_f["lambda0"]), eigvals=ev.numpy(),
    )
    if store_big:
        dkeys = list(out["dC"].keys())
        fx.update(C=out["C"].numpy(), K_tilde=out["K_tilde"].numpy(), K=out["K"].numpy(),
                  dC_keys=np.array(dkeys), dC=np.stack([out["dC"][k].numpy() for k in dkeys]),
                  dK_keys=np.array(list(out["dK"].keys())),
                  dK=np.stack([out["dK"][k].numpy() for k in out["dK"]]),
                  dK_tilde=np.stack([out["dK_tilde"][k].numpy() for k in out["dK_tilde"]]),
                  dKvec=np.stack([out["dKvec"][k].numpy() for k in out["dKvec"]]))
    path = os.path.join(ROOT, "tests", "golden", f"{name}.npz")
    np.savez_compressed(path, **fx)
    print(f"{name}: N={N} n={n} m={m} p={fx['p']} q={fx['q']} loglik={fx['loglik']:.6f} KL={fx['KL']:.6f} "
          f"-> {os.path.getsize(path)/1024:.0f} KiB")


def gen_big_closure_case(ref, name, N, n, m, cfg, theta_dict, perturb=None):
    """Benchmark-size closure fixture (n = 40: the sep_quad_full_kernel<5,.> / K = 1600 oz_gemm instantiations the
    small fixtures never reach).  Only what cannot be regenerated from the seeds is stored: B (eigh is not
    reproducible across machines), the eigenvalues, and the reference's outputs.  m_b / V_b are regenerated by
    ``synth.make_variational_state(diag(ev[keep]), cfg)`` (seeded CPU generator); a few of their entries are kept
    as a guard against generator drift."""
    prob, theta, B, m_b, V_b, ev = setup_case(ref, N, n, m, cfg, theta_dict)
    if perturb:
        theta = synth.theta_tensors({k: prob.theta[k] + perturb.get(k, 0.0) for k in prob.theta})
    f_params = {"logA": torch.tensor(prob.logA), "lambda0": torch.tensor(prob.lambda0)}
    out = ref_closure(ref, theta, prob.X, prob.r, prob.xtilde, n, B, m_b, V_b, f_params)
    keys = list(theta.keys())
    keep = ev > max(ev.max() * ref.EIGVAL_TOL, ref.EIGVAL_TOL)
    fx = dict(
        N=N, n=n, m=m, cfg=cfg, theta=np.array([float(theta[k]) for k in keys]), theta_keys=np.array(keys),
        theta_B=np.array([prob.theta[k] for k in keys]), q=B.shape[1], p=int(out["mask"].sum()),
        B=B.numpy(), eigvals=ev.numpy(), keep=keep.numpy(), m_b=m_b.numpy(),
        V_b_guard=np.array([float(V_b[0, 0]), float(V_b[-1, -1]), float(V_b[0, -1]), float(V_b.sum())]),
        logA=prob.logA, lambda0=prob.lambda0,
        loglik=float(out["loglik"]), KL=float(out["KL"]), dloglik=out["dloglik"].numpy(), dKL=out["dKL"].numpy(),
        lam_m=out["lam_m"].numpy(), lam_var=out["lam_var"].numpy(), f=out["f"].numpy(), Kvec=out["Kvec"].numpy(),
    )
    path = os.path.join(ROOT, "tests", "golden", f"{name}.npz")
    np.savez_compressed(path, **fx)
    print(f"{name}: N={N} n={n} m={m} p={fx['p']} q={fx['q']} loglik={fx['loglik']:.6f} KL={fx['KL']:.6f} "
          f"-> {os.path.getsize(path)/1024:.0f} KiB")


def gen_config5(ref, total=1000000, chunk=100000):
    """BASELINE config 5 (10^6 images of 40 x 40, m = 500) through the real reference, on row chunks (the
    likelihood part of the ELBO and its gradient are sums over images, KL does not depend on them): the reference's closure
    on ten blocks of 100 000 images with the same q x q state, partial loglik / dloglik added.  ~25 minutes, ~30 GB."""
    n, m, cfg = 40, 500, 5
    prob, theta, B, m_b, V_b, ev = setup_case(ref, total, n, m, cfg, synth.THETA_DENSE)
    theta = synth.theta_tensors({k: prob.theta[k] + {"eps_0x": 0.01, "-2log2beta": 0.02}.get(k, 0.0) for k in prob.theta})
    f_params = {"logA": torch.tensor(prob.logA), "lambda0": torch.tensor(prob.lambda0)}
    keys = list(theta.keys())
    loglik, dloglik, KL, dKL = 0.0, None, None, None
    sample_m, sample_v, sample_f = [], [], []
    for c0 in range(0, total, chunk):
        out = ref_closure(ref, theta, prob.X[c0:c0 + chunk], prob.r[c0:c0 + chunk], prob.xtilde, n, B, m_b, V_b, f_params)
        loglik += float(out["loglik"])
        dloglik = out["dloglik"].clone() if dloglik is None else dloglik + out["dloglik"]
        if KL is None:
            KL, dKL = float(out["KL"]), out["dKL"].clone()
        sample_m.append(out["lam_m"][::1000].clone()); sample_v.append(out["lam_var"][::1000].clone())
        sample_f.append(out["f"][::1000].clone())
        print(f"  chunk at {c0}: partial loglik {float(out['loglik']):.6f}", flush=True)
        del out
    keep = ev > max(ev.max() * ref.EIGVAL_TOL, ref.EIGVAL_TOL)
    fx = dict(N=total, n=n, m=m, cfg=cfg, chunk=chunk, theta=np.array([float(theta[k]) for k in keys]), theta_keys=np.array(keys),
              theta_B=np.array([prob.theta[k] for k in keys]), q=B.shape[1], p=n * n, B=B.numpy(), eigvals=ev.numpy(),
              keep=keep.numpy(), m_b=m_b.numpy(),
              V_b_guard=np.array([float(V_b[0, 0]), float(V_b[-1, -1]), float(V_b[0, -1]), float(V_b.sum())]),
              logA=prob.logA, lambda0=prob.lambda0, loglik=loglik, KL=KL, dloglik=dloglik.numpy(), dKL=dKL.numpy(),
              lam_m_every_1000=torch.cat(sample_m).numpy(), lam_var_every_1000=torch.cat(sample_v).numpy(),
              f_every_1000=torch.cat(sample_f).numpy())
    path = os.path.join(ROOT, "tests", "golden", "closure_c5_full_1M.npz")
    np.savez_compressed(path, **fx)
    print(f"closure_c5_full_1M: loglik={loglik:.6f} KL={KL:.6f} -> {os.path.getsize(path)/1024:.0f} KiB")


def gen_config4_cells(ref, N=200000, chunk=100000):
    """BASELINE config 4's shape through the real reference: three cells of a population that share N = 200 000 stimuli of
    40 x 40 (m = 500), each with its own receptive-field centre (all crop the grid: p < P), spikes, subspace and variational
    state.  The reference's closure per cell, on two blocks of 100 000 images (partial loglik / dloglik added)."""
    n, m = 40, 500
    base = dict(synth.THETA_DENSE)
    base["-2log2beta"] = 1.9
    eps = [(0.0, 0.0), (0.45, -0.4), (-0.5, 0.5)]
    prob = synth.make_problem(N, n, m, cfg=4, theta=base)
    g = torch.Generator().manual_seed(4000)
    lo, hi = synth.THETA_LIMS_LO, synth.THETA_LIMS_HI
    fx = dict(N=N, n=n, m=m, cfg=4, ncells=len(eps), theta_keys=np.array(list(synth.THETA_KEYS)), spike_seed=4000)
    for c, (ex, ey) in enumerate(eps):
        th = dict(base, eps_0x=ex, eps_0y=ey)
        theta_B = synth.theta_tensors(th)
        C, mask = ref.localker(theta=theta_B, theta_higher_lims=hi, theta_lower_lims=lo, n_px_side=n, grad=False)
        Kt = ref.acosker(theta_B, prob.xtilde[:, mask], prob.xtilde[:, mask], C=C, dC=None, diag=False)
        ev, evec = torch.linalg.eigh(Kt, UPLO="L")
        keep = ev > max(ev.max() * ref.EIGVAL_TOL, ref.EIGVAL_TOL)
        B = evec[:, keep]
        m_b, V_b = synth.make_variational_state(torch.diag(ev[keep]), 40 + c)
        r = torch.poisson(torch.full((N,), 0.6 + 0.2 * c, dtype=torch.float64), generator=g)
        th_eval = dict(th)
        th_eval["-log2rho2"] += 0.02 * (c + 1)                      # evaluate away from the theta B was built at
        theta = synth.theta_tensors(th_eval)
        f_params = {"logA": torch.tensor(math.log(0.05) + 0.01 * c), "lambda0": torch.tensor(-0.5 - 0.05 * c)}
        loglik, dloglik, KL, dKL, sm, sv = 0.0, None, None, None, [], []
        for c0 in range(0, N, chunk):
            out = ref_closure(ref, theta, prob.X[c0:c0 + chunk], r[c0:c0 + chunk], prob.xtilde, n, B, m_b, V_b, f_params)
            loglik += float(out["loglik"])
            dloglik = out["dloglik"].clone() if dloglik is None else dloglik + out["dloglik"]
            if KL is None:
                KL, dKL, p = float(out["KL"]), out["dKL"].clone(), int(out["mask"].sum())
            sm.append(out["lam_m"][::200].clone()); sv.append(out["lam_var"][::200].clone())
            del out
        keys = list(theta.keys())
        fx.update({f"theta_{c}": np.array([float(theta[k]) for k in keys]), f"B_{c}": B.numpy(), f"eigvals_{c}": ev.numpy(),
                   f"keep_{c}": keep.numpy(), f"m_b_{c}": m_b.numpy(),
                   f"V_b_guard_{c}": np.array([float(V_b[0, 0]), float(V_b[-1, -1]), float(V_b[0, -1]), float(V_b.sum())]),
                   f"logA_{c}": float(f_params["logA"]), f"lambda0_{c}": float(f_params["lambda0"]), f"p_{c}": p,
                   f"r_sum_{c}": float(r.sum()), f"loglik_{c}": loglik, f"KL_{c}": KL, f"dloglik_{c}": dloglik.numpy(),
                   f"dKL_{c}": dKL.numpy(), f"lam_m_every_200_{c}": torch.cat(sm).numpy(),
                   f"lam_var_every_200_{c}": torch.cat(sv).numpy()})
        print(f"  cell {c}: eps=({ex},{ey}) p={p} q={B.shape[1]} loglik={loglik:.6f} KL={KL:.6f}", flush=True)
    path = os.path.join(ROOT, "tests", "golden", "cells_c4_full.npz")
    np.savez_compressed(path, **fx)
    print(f"cells_c4_full -> {os.path.getsize(path)/1024:.0f} KiB")


def gen_big(ref):
    """Fixtures at the sizes bench.py times: config 2 dense / localised, config-3 shape."""
    gen_big_closure_case(ref, "closure_c2_dense_n40", N=8192, n=40, m=500, cfg=2, theta_dict=synth.THETA_DENSE,
                         perturb={"eps_0x": 0.01, "-log2rho2": -0.02, "Amp": 0.05})
    gen_big_closure_case(ref, "closure_c2_local_n40", N=8192, n=40, m=500, cfg=2,
                         theta_dict=synth.theta_localised(40), perturb={"eps_0x": 0.01, "eps_0y": -0.01})
    gen_big_closure_case(ref, "closure_c3_m1000_n40", N=4096, n=40, m=1000, cfg=3, theta_dict=synth.THETA_DENSE,
                         perturb={"eps_0y": 0.01, "sigma_0": 0.05})


def gen_square_cases(ref):
    """ntilde == nt (K = K_tilde, a = B) beyond the full-rank toy case: a rank-deficient one (q < m, so the reference's
    `da` formula does not vanish), a perturbed-theta one, and one at the active-learning notebook's scale (m = N = 300)."""
    rd = {**synth.THETA_NOTEBOOK, "-2log2beta": 3.0, "-log2rho2": 3.5}
    gen_closure_case(ref, "closure_square_rankdef_n10", N=90, n=10, m=90, cfg=15, theta_dict=rd, store_big=False,
                     perturb={"eps_0y": -0.02, "-log2rho2": 0.1, "sigma_0": 0.05})
    gen_closure_case(ref, "closure_square_local_n12", N=70, n=12, m=70, cfg=16, theta_dict=synth.theta_localised(12),
                     store_big=False, perturb={"eps_0x": 0.03, "Amp": 0.2, "-2log2beta": -0.05})
    gen_closure_case(ref, "closure_square_n20_m300", N=300, n=20, m=300, cfg=17, theta_dict=synth.theta_localised(20),
                     store_big=False, perturb={"eps_0x": 0.01, "eps_0y": -0.01})


def gen_active_case(ref, name, n, m0, n_add, same, N_fixed=None, cfg=31, fit=(3, 2, 2, 2)):
    """Active-learning cycles (one_cell_active_training.ipynb cell 17) run by the REAL reference: a start fit with m0
    inducing images, then `n_add` times: append one image, grow K_tilde (and K) by one column, eigh, and the
    warm-started `varGP(X_in_use, R_in_use, **active_model)` with `m`, `V`, `init_kernel`.  The cell is notebook code;
    `gaussian_processes_b200.active.grow_model` restates it and is driven here with U = the reference module.
    same: ntilde == nt (the notebook's case: every image in use is an inducing image); otherwise the training set is the
    fixed first N_fixed images and only the inducing set grows.  Stored: the tracks of every warm-started fit."""
    from gaussian_processes_b200 import active
    maxiter, nE, nM, nF = fit
    npool = m0 + n_add if same else N_fixed
    prob = synth.make_problem(max(npool, m0 + n_add), n, m0, cfg=cfg, theta=synth.theta_localised(n))
    X, R = prob.X, prob.r
    theta = synth.theta_tensors(prob.theta, requires_grad=True)
    X_in_use, R_in_use = (X[:m0], R[:m0]) if same else (X[:N_fixed], R[:N_fixed])
    args = dict(
        fit_parameters=dict(ntilde=m0, maxiter=maxiter, nEstep=nE, nMstep=nM, nFparamstep=nF, kernfun="acosker", cellid=0,
                            n_px_side=n, display_hyper=False),
        xtilde=X[:m0].clone(), hyperparams_tuple=(theta, dict(synth.THETA_LIMS_LO), dict(synth.THETA_LIMS_HI)),
        f_params={"logA": torch.tensor(prob.logA), "lambda0": torch.tensor(prob.lambda0)})
    with contextlib.redirect_stdout(io.StringIO()):
        model, err = ref.varGP(X_in_use, R_in_use, **args)
    assert not err["is_error"], err
    keys = list(model["hyperparams_tuple"][0].keys())
    fx = dict(n=n, m0=m0, n_add=n_add, same=int(same), N_fixed=N_fixed or 0, cfg=cfg, maxiter=maxiter, nEstep=nE, nMstep=nM,
              nFparamstep=nF, theta_keys=np.array(keys), logA0=prob.logA, lambda00=prob.lambda0,
              start_logmarginal=model["values_track"]["loss_track"]["logmarginal"].numpy())
    for j in range(n_add):
        xt_new = X[:m0 + j + 1].clone()
        if same:
            X_in_use, R_in_use = X[:m0 + j + 1], R[:m0 + j + 1]
        with contextlib.redirect_stdout(io.StringIO()):
            model, err = active.add_image_and_refit(ref, model, X_in_use, R_in_use, xt_new)
        assert not err["is_error"], err
        vt = model["values_track"]
        fx[f"logmarginal_{j}"] = vt["loss_track"]["logmarginal"].numpy()
        fx[f"loglikelihood_{j}"] = vt["loss_track"]["loglikelihood"].numpy()
        fx[f"KL_{j}"] = vt["loss_track"]["KL"].numpy()
        fx[f"theta_track_{j}"] = np.stack([vt["theta_track"][k].numpy() for k in keys])
        fx[f"logA_track_{j}"] = vt["f_par_track"]["logA"].numpy()
        fx[f"lambda0_track_{j}"] = vt["f_par_track"]["lambda0"].numpy()
        fx[f"q_{j}"] = model["B"].shape[1]
    path = os.path.join(ROOT, "tests", "golden", f"{name}.npz")
    np.savez_compressed(path, **fx)
    print(f"{name}: start {fx['start_logmarginal'][-1]:.4f}; " +
          ", ".join(f"step {j}: {fx['logmarginal_%d' % j][0]:.4f} -> {fx['logmarginal_%d' % j][-1]:.4f} (q={fx['q_%d' % j]})"
                    for j in range(n_add)))


def gen_active(ref):
    gen_active_case(ref, "active_same_n10", n=10, m0=40, n_add=3, same=True)
    gen_active_case(ref, "active_fixedN_n10", n=10, m0=40, n_add=3, same=False, N_fixed=400, cfg=32)
    # larger cycles: 20 x 20 pixels, a fixed training set above the int8 threshold (N = 5000) with m = 100 -> 103, and the
    # notebook's ntilde == nt configuration at m = 150 -> 153
    gen_active_case(ref, "active_fixedN_n20_N5000", n=20, m0=100, n_add=3, same=False, N_fixed=5000, cfg=33)
    gen_active_case(ref, "active_same_n20_m150", n=20, m0=150, n_add=3, same=True, cfg=34)


def gen_kat(ref):
    """The toy known-answer vectors of moments_gradients.ipynb cells 1-3, recomputed with the
    reference's own lambda_moments / compute_loglikelihood / log_det and compared with the
    numbers stored in the notebook output."""
    a_mat = torch.tensor([[1.0, 2.0, 3.0, 4], [5.0, 6.0, 7.0, 8], [9.0, 10.0, 11.0, 12]])
    Sigma = torch.tensor([[1.0, 2.0, 3.0], [5.0, 6.0, 7.0], [8.0, 7.0, 6.0]]); Sigma = Sigma + Sigma.T
    dSigma = torch.tensor([[4.0, 5.0, 6.0], [5.0, 6.0, 7.0], [6.0, 7.0, 8.0]]); dSigma = dSigma + dSigma.T
    m = torch.tensor([1.0, 2.0, 3.0])
    dki = torch.tensor([[6.0, 7.0, 8.0, 9.0], [2.0, 3.0, 4.0, 5.0], [9.0, 8.0, 7.0, 6]])
    dkstar = torch.tensor([3.0, 4, 5, 6])
    invV = Sigma * 3
    a, K, dK = a_mat.T, dki.T, dki.T
    K_tilde_inv = torch.linalg.pinv(Sigma)
    V = torch.linalg.inv(invV)
    lam_m, lam_var, dlm, dlv = ref.lambda_moments(None, Sigma, a, dkstar, K, None, m, V, None, dK={"k": dK},
                                                  dK_tilde={"k": dSigma}, dK_vec={"k": dkstar}, K_tilde_inv=K_tilde_inv)
    f = torch.tensor([55.0, 4.0, 22.0, 5.0]); r = torch.tensor([23.0, 47.0, 2.0, 1.0])
    lam_m_toy = torch.tensor([1.0, 2.0, 3.0, 4.0])
    fp = {"logA": torch.tensor(0.0), "lambda0": torch.tensor(0.0)}
    ll, dll = ref.compute_loglikelihood(r, f, lam_m_toy, None, fp, dlambda_m=dlm, dlambda_var=dlv)
    c = V @ K_tilde_inv; b = K_tilde_inv @ m; Bm = dSigma @ K_tilde_inv
    I10 = 10 * torch.eye(3)
    KLD = -0.5 * (ref.log_det(V + I10) - ref.log_det(Sigma + I10)) + 0.5 * (m @ b) + 0.5 * torch.trace(c)
    dKL = 0.5 * torch.trace(Bm) - 0.5 * torch.trace(c @ Bm) - 0.5 * b @ (Bm @ m)
    got = dict(dlam_m=dlm["k"].numpy(), dlam_var=dlv["k"].numpy(), logLK=float(ll), dlogLK=float(dll["k"]),
               KLD=float(KLD), dKL=float(dKL))
    # numbers printed in the notebook's stored output
    assert np.allclose(got["dlam_m"], [13.75, 19.9167, 26.0833, 32.25], atol=1e-4)
    assert np.allclose(got["dlam_var"], [-2199.1944, -3161.6944, -4291.75, -5589.3611], atol=1e-4)
    assert got["logLK"] == 41.0 and abs(got["dlogLK"] - 127749.63888888987) < 1e-6
    assert abs(got["KLD"] - 34.91913283450331) < 1e-10 and abs(got["dKL"] - 110.47222222222383) < 1e-9
    print("KAT (moments_gradients.ipynb) reproduced by the reference:", {k: v for k, v in got.items() if np.isscalar(v)})


def gen_fit_case(ref, name, N, n, m, cfg, theta_dict, maxiter, nE, nM, nF):
    """Loop-level fixture: the reference's own varGP on a small synthetic problem."""
    prob = synth.make_problem(N, n, m, cfg=cfg, theta=theta_dict)
    theta = synth.theta_tensors(prob.theta, requires_grad=True)
    args = dict(
        fit_parameters=dict(ntilde=m, maxiter=maxiter, nEstep=nE, nMstep=nM, nFparamstep=nF, kernfun="acosker",
                            cellid=0, n_px_side=n, display_hyper=False),
        xtilde=prob.xtilde, hyperparams_tuple=(theta, dict(synth.THETA_LIMS_LO), dict(synth.THETA_LIMS_HI)),
        f_params={"logA": torch.tensor(prob.logA), "lambda0": torch.tensor(prob.lambda0)},
    )
    with contextlib.redirect_stdout(io.StringIO()):
        fit, err = ref.varGP(prob.X, prob.r, **args)
    assert not err["is_error"], err
    vt = fit["values_track"]
    keys = list(fit["hyperparams_tuple"][0].keys())
    fx = dict(N=N, n=n, m=m, cfg=cfg, maxiter=maxiter, nEstep=nE, nMstep=nM, nFparamstep=nF,
              theta0=np.array([prob.theta[k] for k in keys]), theta_keys=np.array(keys),
              logA0=prob.logA, lambda00=prob.lambda0,
              logmarginal=vt["loss_track"]["logmarginal"].numpy(), loglikelihood=vt["loss_track"]["loglikelihood"].numpy(),
              KL=vt["loss_track"]["KL"].numpy(),
              theta_track=np.stack([vt["theta_track"][k].numpy() for k in keys]),
              logA_track=vt["f_par_track"]["logA"].numpy(), lambda0_track=vt["f_par_track"]["lambda0"].numpy(),
              q=fit["B"].shape[1], m_b=fit["m_b"].numpy(), V_b=fit["V_b"].numpy())
    path = os.path.join(ROOT, "tests", "golden", f"{name}.npz")
    np.savez_compressed(path, **fx)
    print(f"{name}: logmarginal track {fx['logmarginal']}")
    return prob, fit


def gen_fit_big(ref):
    """Loop-level fixtures at sizes that reach the production kernels: BASELINE config 1 itself (separable quadratic forms,
    cluster Cholesky), and an N = 8192 / 40 x 40 / m = 200 fit whose image-sized products run on the int8-slice engine."""
    gen_fit_case(ref, "fit_c1_n20", N=2000, n=20, m=100, cfg=1, theta_dict=synth.theta_localised(20),
                 maxiter=3, nE=3, nM=3, nF=3)
    gen_fit_case(ref, "fit_n40_N8192_m200", N=8192, n=40, m=200, cfg=22, theta_dict=synth.THETA_DENSE,
                 maxiter=3, nE=2, nM=3, nF=3)      # maxiter = 3: the reference's first M-step runs in iteration 1


def gen_test_and_checkpoint(ref, prob, fit, name):
    """the reference's own test() / lambda_moments_star / explained_variance(sigma=False) on
    seeded test images for the model `fit` (the reference's varGP result of gen_fit_case), and that model written by
    the reference's own save_model (pickle + metadata) -- the on-disk format the replacement must read and write."""
    import shutil
    n = prob.n_px_side
    g = torch.Generator().manual_seed(5021)
    ntest, ntrials, ncells = 30, 10, 3
    X_test = torch.randn(ntest, n, n, 1, generator=g, dtype=torch.float64)
    R_test = torch.poisson(torch.full((ntrials, ntest, ncells), 1.5, dtype=torch.float64), generator=g)
    # give the cell of interest a stimulus-driven part so that the correlations of explained_variance are not ~0
    drive = torch.exp(0.5 * X_test.reshape(ntest, -1)[:, : n].sum(1))
    R_test[:, :, 0] = torch.poisson(drive.expand(ntrials, ntest).contiguous(), generator=g)
    with contextlib.redirect_stdout(io.StringIO()):
        R_cell, R_pred, r2, sr2 = ref.test(X_test, R_test, X_train=prob.X, at_iteration=None, **fit)   # as the notebooks call it
    mask = fit["mask"]
    xs = X_test.reshape(ntest, -1)
    theta = fit["hyperparams_tuple"][0]
    mu, s2 = [], []
    for i in range(ntest):   # the reference's per-image call (n1 = 1: never symmetrised)
        a, b = ref.lambda_moments_star(xs[i:i + 1, mask], prob.xtilde[:, mask], fit["C"], theta, fit["K_tilde_b"],
                                       fit["K_tilde_inv_b"], fit["m_b"], fit["V_b"], fit["B"], "acosker")
        mu.append(a); s2.append(b)
    r2_ns, _ = ref.explained_variance(R_test[:, :, 0], R_pred, sigma=False)
    keys = list(theta.keys())
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", f"{name}.npz"),
                        X_test=X_test.numpy(), R_test=R_test.numpy(), R_pred=R_pred.numpy(), mu_star=torch.cat(mu).numpy(),
                        sigma_star2=torch.cat(s2).numpy(), r2_nosigma=float(r2_ns), r2_boot=float(r2), sr2_boot=float(sr2),
                        theta=np.array([float(theta[k]) for k in keys]), theta_keys=np.array(keys),
                        mask=mask.numpy(), C=fit["C"].numpy(), B=fit["B"].numpy(), m_b=fit["m_b"].numpy(),
                        V_b=fit["V_b"].numpy(), K_tilde_b=fit["K_tilde_b"].numpy(),
                        K_tilde_inv_b=fit["K_tilde_inv_b"].numpy(), logA=float(fit["f_params"]["logA"]),
                        lambda0=float(fit["f_params"]["lambda0"]), N=prob.N, n=n, m=prob.m, cfg=prob.cfg)
    d = os.path.join(ROOT, "tests", "golden", f"{name}_saved_by_reference")
    shutil.rmtree(d, ignore_errors=True)
    ref.save_model(fit, d, additional_description="written by the reference's save_model (oracle/gen_golden.py)")
    print(f"{name}: R_pred[:4] = {R_pred[:4].numpy()}, r2(sigma=False) = {float(r2_ns):.6f}; model pickled to {d} "
          f"({os.path.getsize(os.path.join(d, 'model'))/1024:.0f} KiB)")


def gen_utility(ref):
    """nd_utility family (U:416-525) on seeded (mu, sigma2) pairs, including candidates whose z overflows for the
    larger counts (the reference's masking quirk) and the 0-dim call of the single-candidate variants."""
    g = torch.Generator().manual_seed(77)
    nstar = 96
    mu = -1.5 + 1.2 * torch.randn(nstar, generator=g, dtype=torch.float64)
    sigma2 = torch.exp(-1.0 + 1.3 * torch.randn(nstar, generator=g, dtype=torch.float64))
    sigma2[:6] = torch.tensor([8.0, 12.0, 20.0, 1e-6, 1e-3, 5.0], dtype=torch.float64)   # overflow / tiny variances
    mu[:6] = torch.tensor([2.0, -3.0, 0.5, -2.0, 1.0, 4.0], dtype=torch.float64)
    r = torch.arange(0, 100, dtype=torch.float64)
    lam, keep = ref.nd_lambda_r_mean(r, sigma2, mu)
    p, log_p, r2d, lrf = ref.nd_p_r_given_xD(r, sigma2, mu)
    U = ref.nd_utility(sigma2, mu, r)
    # single-candidate variants (U:540-611) for a candidate without overflow
    j = 10
    p1, log_p1, r1 = ref.p_r_given_xD(r, sigma2[j], mu[j])
    U1 = ref.utility(sigma2[j], mu[j], r)
    assert bool((~keep).any()), "fixture must contain masked entries"
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "utility_n96.npz"),
                        mu=mu.numpy(), sigma2=sigma2.numpy(), r=r.numpy(), lam=lam.numpy(), keep=keep.numpy(),
                        p=p.numpy(), log_p=log_p.numpy(), r2d=r2d.numpy(), lrf=lrf.numpy(), U=U.numpy(),
                        single_index=j, p1=p1.numpy(), log_p1=log_p1.numpy(), r1=r1.numpy(), U1=float(U1))
    print("utility_n96: masked entries", int((~keep).sum()), "U range", float(U.min()), float(U.max()))


def main():
    ref = import_reference()
    assert torch.pi == 3.1415927410125732  # U:25
    gen_kat(ref)
    gen_utility(ref)
    if len(sys.argv) > 1 and sys.argv[1] == "utility":
        return
    if len(sys.argv) > 1 and sys.argv[1] == "active":
        gen_active(ref)
        return
    if len(sys.argv) > 1 and sys.argv[1] == "square":
        gen_square_cases(ref)
        return
    if len(sys.argv) > 1 and sys.argv[1] == "big":
        gen_big(ref)
        return
    if len(sys.argv) > 1 and sys.argv[1] == "config4":
        gen_config4_cells(ref)
        return
    if len(sys.argv) > 1 and sys.argv[1] == "config5":
        gen_config5(ref)
        return
    if len(sys.argv) > 1 and sys.argv[1] == "bigfull":
        # the closure at BASELINE's full sizes: config 2 (N = 100 000, m = 500), dense and localised theta, and config 3's
        # end point (N = 50 000, m = 1000).  Minutes each and up to ~25 GB on the host.
        gen_big_closure_case(ref, "closure_c2_full_dense", N=100000, n=40, m=500, cfg=2, theta_dict=synth.THETA_DENSE,
                             perturb={"eps_0x": 0.01, "-log2rho2": -0.02, "Amp": 0.05})
        gen_big_closure_case(ref, "closure_c2_full_local", N=100000, n=40, m=500, cfg=2,
                             theta_dict=synth.theta_localised(40), perturb={"eps_0x": 0.01, "eps_0y": -0.01})
        gen_big_closure_case(ref, "closure_c3_full_m1000", N=50000, n=40, m=1000, cfg=3, theta_dict=synth.THETA_DENSE,
                             perturb={"eps_0y": 0.01, "sigma_0": 0.05})
        return
    if len(sys.argv) > 1 and sys.argv[1] == "fitc2":
        # BASELINE config 2 itself (N = 100 000, 40 x 40, m = 500, dense theta): ~10 minutes and ~15 GB on 8 host cores
        gen_fit_case(ref, "fit_c2_n40_N100000_m500", N=100000, n=40, m=500, cfg=2, theta_dict=synth.THETA_DENSE,
                     maxiter=3, nE=2, nM=2, nF=2)
        return
    if len(sys.argv) > 1 and sys.argv[1] == "fitbig":
        gen_fit_big(ref)
        return
    if len(sys.argv) > 1 and sys.argv[1] == "fit":
        prob_f, fit_f = gen_fit_case(ref, "fit_small_n10", N=400, n=10, m=40, cfg=21,
                                     theta_dict=synth.theta_localised(10), maxiter=4, nE=3, nM=3, nF=3)
        gen_test_and_checkpoint(ref, prob_f, fit_f, "test_small_n10")
        return
    # small closure cases, every intermediate stored
    gen_closure_case(ref, "closure_dense_n8", N=150, n=8, m=24, cfg=11, theta_dict=synth.THETA_DENSE, store_big=True)
    gen_closure_case(ref, "closure_local_n12", N=260, n=12, m=40, cfg=12, theta_dict=synth.theta_localised(12),
                     store_big=True, perturb={"eps_0x": 0.03, "-2log2beta": -0.05, "sigma_0": 0.1, "Amp": 0.2})
    # m > p: K_tilde is rank deficient, q < m (eigen-truncation active)
    gen_closure_case(ref, "closure_rankdef_n10", N=300, n=10, m=90, cfg=13,
                     theta_dict={**synth.THETA_NOTEBOOK, "-2log2beta": 3.0, "-log2rho2": 3.5}, store_big=False,
                     perturb={"eps_0y": -0.02, "-log2rho2": 0.1})
    # ntilde == nt (active-learning configuration: K = K_tilde, a = B)
    gen_closure_case(ref, "closure_square_n8", N=48, n=8, m=48, cfg=14, theta_dict=synth.THETA_DENSE, store_big=False)
    gen_square_cases(ref)
    # BASELINE config 1 (20x20, N=2000, m=100), dense and localised theta
    gen_closure_case(ref, "closure_c1_dense", N=2000, n=20, m=100, cfg=1, theta_dict=synth.THETA_DENSE)
    gen_closure_case(ref, "closure_c1_local", N=2000, n=20, m=100, cfg=1, theta_dict=synth.theta_localised(20),
                     perturb={"eps_0x": 0.01, "eps_0y": -0.01})
    prob_f, fit_f = gen_fit_case(ref, "fit_small_n10", N=400, n=10, m=40, cfg=21, theta_dict=synth.theta_localised(10),
                                 maxiter=4, nE=3, nM=3, nF=3)
    gen_test_and_checkpoint(ref, prob_f, fit_f, "test_small_n10")
    gen_active(ref)
    gen_big(ref)
    gen_fit_big(ref)


if __name__ == "__main__":
    main()
