"""GPU: the shape of the reference's REAL data (one_cell_fit.ipynb cell 4: 108 x 108 images, ~3160 of them, ntilde = 200)
at the notebook's theta (p = 500 of 11 664 pixels inside the mask, q = 124 < m) and at generate_theta's default (p = 848):
the fused closure against the CPU oracle (itself pinned to the real reference, tests/test_oracle_golden.py).  n = 108 is
beyond the separable kernel (n <= 48), so this is the dense-C route on a cropped support with ld = 11 664."""
import numpy as np
import pytest
import torch

import oracle.spatial_gp_oracle as orc
from gaussian_processes_b200 import synth
from helpers import record

pytestmark = pytest.mark.gpu
RTOL = 1e-9
_prob = {}


def problem():
    if not _prob:
        _prob["p"] = synth.make_problem(3160, 108, 200, cfg=108, theta=synth.THETA_NOTEBOOK)
    return _prob["p"]


@pytest.mark.parametrize("kind", ["notebook", "generate_theta", "ntilde_eq_nt"])
def test_real_data_shape_against_oracle(kind):
    from gaussian_processes_b200 import native
    prob = problem()
    n = 108
    theta0 = dict(synth.THETA_NOTEBOOK) if kind != "generate_theta" else synth.theta_localised(n)
    X, r = (prob.X, prob.r) if kind != "ntilde_eq_nt" else (prob.X[:200], prob.r[:200])   # active-learning configuration
    C, mask = orc.build_C(theta0, n)
    um = prob.xtilde[:, mask]
    B, Kt_b, _ = orc.eig_projection(orc.arccos_kernel(theta0, um, um, C))
    m_b, V_b = synth.make_variational_state(Kt_b, 108)
    theta = dict(theta0)
    theta["eps_0x"] += 0.004                      # evaluate away from the theta B was built at; the mask flips pixels
    theta["-2log2beta"] -= 0.03
    ref = orc.elbo_and_grad(theta, X, r, prob.xtilde, n, B, m_b, V_b, prob.logA, prob.lambda0, return_moments=True)
    d = torch.device("cuda:0")
    native.invalidate_cache(0)
    out = native.elbo_grad(X.to(d), r.to(d), prob.xtilde.to(d), n, theta, B.to(d), m_b.to(d), V_b.to(d), prob.logA,
                           prob.lambda0, want_moments=True)
    assert out["flags"] == 0 and out["p"] == ref["p"]
    if kind == "notebook":
        assert B.shape[1] < 200                   # the eigen-truncation is active at the notebook's theta (q = 124)
    e_ll = abs(out["loglik"] - ref["loglik"]) / abs(ref["loglik"])
    e_kl = abs(out["KL"] - ref["KL"]) / abs(ref["KL"])
    e_g = max(abs(out["grad"][k] - ref["grad"][k]) / max(abs(ref["dloglik"][k]), abs(ref["dKL"][k])) for k in orc.KEYS)
    e_lv = float((out["lam_var"].cpu() - ref["lam_var"]).abs().max() / ref["lam_var"].abs().max())
    record("realshape/" + kind, loglik=e_ll, KL=e_kl, grad_comp=e_g, lam_var=e_lv, p=out["p"], q=B.shape[1])
    assert e_ll <= RTOL and e_kl <= RTOL and e_g <= RTOL and e_lv <= RTOL
    assert np.allclose(out["lam_m"].cpu().numpy(), ref["lam_m"].numpy(), rtol=RTOL, atol=1e-12)
    native.invalidate_cache(0)
