"""Shared by the CPU (stand-in) and GPU tests: run the active-learning cycles of an `active_*` fixture through a
`utils`-compatible module and return the tracks of every warm-started fit."""
import contextlib
import io

import numpy as np
import torch

from gaussian_processes_b200 import active, synth
from helpers import load


def run_active_fixture(U, name, dev="cpu"):
    fx = load(name)
    n, m0, n_add, same = int(fx["n"]), int(fx["m0"]), int(fx["n_add"]), bool(int(fx["same"]))
    N_fixed = int(fx["N_fixed"])
    npool = m0 + n_add if same else N_fixed
    prob = synth.make_problem(max(npool, m0 + n_add), n, m0, cfg=int(fx["cfg"]), theta=synth.theta_localised(n))
    X, R = prob.X.to(dev), prob.r.to(dev)
    theta = synth.theta_tensors(prob.theta, device=dev, requires_grad=True)
    X_in_use, R_in_use = (X[:m0], R[:m0]) if same else (X[:N_fixed], R[:N_fixed])
    args = dict(
        fit_parameters=dict(ntilde=m0, maxiter=int(fx["maxiter"]), nEstep=int(fx["nEstep"]), nMstep=int(fx["nMstep"]),
                            nFparamstep=int(fx["nFparamstep"]), kernfun="acosker", cellid=0, n_px_side=n,
                            display_hyper=False),
        xtilde=X[:m0].clone(), hyperparams_tuple=(theta, dict(synth.THETA_LIMS_LO), dict(synth.THETA_LIMS_HI)),
        f_params={"logA": torch.tensor(prob.logA, device=dev), "lambda0": torch.tensor(prob.lambda0, device=dev)})
    with contextlib.redirect_stdout(io.StringIO()):
        model, err = U.varGP(X_in_use, R_in_use, **args)
    assert not err["is_error"], err
    keys = [str(k) for k in fx["theta_keys"]]
    worst = {"start": float(np.abs(model["values_track"]["loss_track"]["logmarginal"].cpu().numpy()
                                   - fx["start_logmarginal"]).max() / np.abs(fx["start_logmarginal"]).max())}
    for j in range(n_add):
        xt_new = X[:m0 + j + 1].clone()
        if same:
            X_in_use, R_in_use = X[:m0 + j + 1], R[:m0 + j + 1]
        with contextlib.redirect_stdout(io.StringIO()):
            model, err = active.add_image_and_refit(U, model, X_in_use, R_in_use, xt_new)
        assert not err["is_error"], err
        vt = model["values_track"]
        assert model["B"].shape[1] == int(fx[f"q_{j}"])

        def rel(a, b):
            return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-300))
        worst[f"logmarginal_{j}"] = rel(vt["loss_track"]["logmarginal"].cpu().numpy(), fx[f"logmarginal_{j}"])
        worst[f"KL_{j}"] = rel(vt["loss_track"]["KL"].cpu().numpy(), fx[f"KL_{j}"])
        worst[f"theta_{j}"] = rel(np.stack([vt["theta_track"][k].cpu().numpy() for k in keys]), fx[f"theta_track_{j}"])
        worst[f"logA_{j}"] = rel(vt["f_par_track"]["logA"].cpu().numpy(), fx[f"logA_track_{j}"])
    return worst
