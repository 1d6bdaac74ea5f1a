"""GPU probe: stage times of a short EM fit through utils.varGP at config 2 (the function's own timers)."""
import math
import os
import sys
import time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from gaussian_processes_b200 import synth, utils as U, native  # noqa: E402

dev = torch.device("cuda:0")
N, n, m = int(os.environ.get("N", 100000)), 40, int(os.environ.get("M", 500))
prob = synth.make_problem(N, n, m, cfg=2, theta=synth.THETA_DENSE)
X, r, xt = prob.X.to(dev), prob.r.to(dev), prob.xtilde.to(dev)
kw = dict(fit_parameters=dict(ntilde=m, maxiter=int(os.environ.get("MAXITER", 4)), nEstep=10, nMstep=10, nFparamstep=10,
                              kernfun="acosker", cellid=0, n_px_side=n, display_hyper=False),
          xtilde=xt, hyperparams_tuple=(synth.theta_tensors(prob.theta, device=dev, requires_grad=True),
                                        dict(synth.THETA_LIMS_LO), dict(synth.THETA_LIMS_HI)),
          f_params={"logA": torch.tensor(prob.logA, device=dev), "lambda0": torch.tensor(prob.lambda0, device=dev)})
for rep in range(2):
    torch.cuda.synchronize()
    native.launch_count(0, reset=True)
    t0 = time.perf_counter()
    fit, err = U.varGP(X, r, **kw)
    torch.cuda.synchronize()
    print("TOTAL wall %.3f s, launches %d, error %s" % (time.perf_counter() - t0, native.launch_count(0), err["is_error"]))
