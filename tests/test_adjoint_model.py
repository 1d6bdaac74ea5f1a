"""The reverse-mode formulation implemented by the CUDA path reproduces the reference's forward-mode
analytic gradient (fixtures from the real reference).  CPU only."""
import numpy as np
import pytest

import adjoint_model as adj
from helpers import case_inputs, load

CASES = ["closure_dense_n8", "closure_local_n12", "closure_rankdef_n10", "closure_c1_dense", "closure_c1_local"]


@pytest.mark.parametrize("name", CASES)
def test_adjoint_equals_reference_gradient(name):
    fx = load(name)
    prob, st = case_inputs(fx)
    out = adj.elbo_and_grad(prob.theta, prob.X, prob.r, prob.xtilde, prob.n_px_side, st["B"], st["m_b"], st["V_b"],
                            st["logA"], st["lambda0"])
    assert abs(out["loglik"] - float(fx["loglik"])) <= 1e-10 * abs(float(fx["loglik"]))
    assert abs(out["KL"] - float(fx["KL"])) <= 1e-10 * abs(float(fx["KL"]))
    keys = [str(k) for k in fx["theta_keys"]]
    ref = fx["dloglik"] - fx["dKL"]
    got = np.array([out["grad"][k] for k in keys])
    scale = max(np.abs(fx["dloglik"]).max(), np.abs(fx["dKL"]).max())
    assert np.abs(got - ref).max() < 1e-9 * scale, (got, ref)
    assert np.allclose(out["lam_m"].numpy(), fx["lam_m"], rtol=1e-9, atol=1e-12)
    assert np.allclose(out["lam_var"].numpy(), fx["lam_var"], rtol=1e-9, atol=1e-11)
    assert np.allclose(out["f"].numpy(), fx["f"], rtol=1e-9, atol=0)
