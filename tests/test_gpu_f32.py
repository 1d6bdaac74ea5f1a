"""GPU: fp32 mode (sgp_elbo_grad_f32: float32 images, fp32-class int8-slice engine) against the CPU oracle evaluated in
fp64 on the SAME fp32-valued images.  north_star's bar for fp32: 1e-4 relative for ELBO, moments and gradients.
(The reference itself is fp64 only, utils.py:31-33: there is no fp32 reference output to compare with.)"""
import numpy as np
import pytest
import torch

import oracle.spatial_gp_oracle as orc
from gaussian_processes_b200 import synth
from helpers import case_inputs, load, record

pytestmark = pytest.mark.gpu
TOL = 1e-4


def reference_on_fp32_images(prob, st):
    X32, U32 = prob.X.float(), prob.xtilde.float()
    ref = orc.elbo_and_grad_chunked(prob.theta, X32.double(), prob.r, U32.double(), prob.n_px_side, st["B"], st["m_b"],
                                    st["V_b"], st["logA"], st["lambda0"], chunk=4096)
    return X32, U32, ref


@pytest.mark.parametrize("slices", [-1, 8])
@pytest.mark.parametrize("name,slab", [("closure_c1_dense", 0), ("closure_c1_local", 512), ("closure_rankdef_n10", 128),
                                       ("closure_c2_dense_n40", 4096), ("closure_c2_local_n40", 4096),
                                       ("closure_square_local_n12", 0)])
def test_fp32_images_against_fp64_oracle(name, slab, slices):
    from gaussian_processes_b200 import native
    fx = load(name)
    prob, st = case_inputs(fx)
    X32, U32, ref = reference_on_fp32_images(prob, st)
    d = torch.device("cuda:0")
    native.set_slab_rows(0, slab)
    native.set_option(0, "i8_slices", slices)       # -1: the mode's own choice (5 slices where the int8 engine applies)
    try:
        out = native.elbo_grad_f32(X32.to(d), prob.r.to(d), U32.to(d), prob.n_px_side, prob.theta, st["B"].to(d),
                                   st["m_b"].to(d), st["V_b"].to(d), st["logA"], st["lambda0"], want_moments=True)
    finally:
        native.set_slab_rows(0, 0)
        native.set_option(0, "i8_slices", -1)
    assert out["flags"] == 0 and out["p"] == ref["p"]
    e_ll = abs(out["loglik"] - ref["loglik"]) / abs(ref["loglik"])
    e_kl = abs(out["KL"] - ref["KL"]) / abs(ref["KL"])
    e_g = max(abs(out["grad"][k] - ref["grad"][k]) / max(abs(ref["dloglik"][k]), abs(ref["dKL"][k])) for k in orc.KEYS)
    record(f"f32/{name}/S{slices}", loglik=e_ll, KL=e_kl, grad_comp=e_g)
    bar = TOL if slices < 0 else 1e-9                # with 8 slices the fp32 entry is as exact as the fp64 one
    assert e_ll <= bar and e_kl <= bar and e_g <= bar, (e_ll, e_kl, e_g)
    if name != "closure_square_local_n12":
        # moments against the oracle on a prefix (the chunked oracle does not return them)
        nn = min(prob.N, 1024)
        full = orc.elbo_and_grad(prob.theta, X32[:nn].double(), prob.r[:nn], U32.double(), prob.n_px_side, st["B"],
                                 st["m_b"], st["V_b"], st["logA"], st["lambda0"], return_moments=True)
        assert np.allclose(out["lam_m"][:nn].cpu().numpy(), full["lam_m"].numpy(), rtol=bar, atol=bar * 1e-2)
        assert np.allclose(out["lam_var"][:nn].cpu().numpy(), full["lam_var"].numpy(), rtol=bar, atol=bar * 1e-2)
        assert np.allclose(out["f"][:nn].cpu().numpy(), full["f"].numpy(), rtol=bar, atol=0)


def test_fp32_entry_rejects_fp64_images():
    from gaussian_processes_b200 import native
    fx = load("closure_dense_n8")
    prob, st = case_inputs(fx)
    d = torch.device("cuda:0")
    with pytest.raises(TypeError):
        native.elbo_grad_f32(prob.X.to(d), prob.r.to(d), prob.xtilde.to(d), prob.n_px_side, prob.theta, st["B"].to(d),
                             st["m_b"].to(d), st["V_b"].to(d), st["logA"], st["lambda0"])


@pytest.mark.parametrize("name", ["closure_c2_full_dense", "closure_c3_full_m1000"])
def test_fp32_mode_at_full_size_against_the_real_reference(name):
    """fp32 mode at BASELINE's full sizes (config 2: N = 100 000, m = 500; config 3's end point: N = 50 000, m = 1000) against
    the REAL reference's fp64 closure on the fp64 images (oracle/gen_golden.py bigfull): rounding the images to float32 and
    running the products on five int8 slices must stay inside north_star's fp32 bar, 1e-4, on every output."""
    from gaussian_processes_b200 import native
    fx = load(name)
    prob, st = case_inputs(fx)
    d = torch.device("cuda:0")
    X32, U32 = prob.X.float().to(d), prob.xtilde.float().to(d)
    del prob.X
    native.invalidate_cache(0)
    try:
        out = native.elbo_grad_f32(X32, prob.r.to(d), U32, prob.n_px_side, prob.theta, st["B"].to(d), st["m_b"].to(d),
                                   st["V_b"].to(d), st["logA"], st["lambda0"], want_moments=True)
    finally:
        del X32
        native.release_cache(0)
        torch.cuda.empty_cache()
    keys = [str(k) for k in fx["theta_keys"]]
    e_ll = abs(out["loglik"] - float(fx["loglik"])) / abs(float(fx["loglik"]))
    e_kl = abs(out["KL"] - float(fx["KL"])) / abs(float(fx["KL"]))
    e_g = max(abs(out["grad"][k] - (fx["dloglik"][i] - fx["dKL"][i])) / max(abs(fx["dloglik"][i]), abs(fx["dKL"][i]))
              for i, k in enumerate(keys))
    e_m = float(np.abs(out["lam_m"].cpu().numpy() - fx["lam_m"]).max() / np.abs(fx["lam_m"]).max())
    e_v = float(np.abs(out["lam_var"].cpu().numpy() - fx["lam_var"]).max() / np.abs(fx["lam_var"]).max())
    record(f"f32_full/{name}", loglik=e_ll, KL=e_kl, grad_comp=e_g, lam_m=e_m, lam_var=e_v)
    assert out["flags"] == 0 and out["p"] == int(fx["p"])
    assert max(e_ll, e_kl, e_g, e_m, e_v) <= TOL, (e_ll, e_kl, e_g, e_m, e_v)
