"""GPU: edge shapes of the fused closure against the CPU oracle (pinned to the real reference): one training image, a
receptive-field mask of a handful of pixels, more inducing images than pixels in the mask, tiny ntilde == nt, an image
count that is not a multiple of any tile, and theta exactly on its limits."""
import pytest
import torch

import oracle.spatial_gp_oracle as orc
from gaussian_processes_b200 import synth

pytestmark = pytest.mark.gpu
RTOL = 1e-9
D = torch.device("cuda:0")


def run(N, n, m, theta0, dtheta=None, cfg=60, i8=-1, slab=0, xt_rows=None):
    from gaussian_processes_b200 import native
    prob = synth.make_problem(max(N, m), n, m, cfg=cfg, theta=theta0)
    X, r = prob.X[:N], prob.r[:N]
    xt = prob.xtilde if xt_rows is None else prob.X[xt_rows]
    C, mask = orc.build_C(theta0, n)
    um = xt[:, mask]
    B, Kt_b, _ = orc.eig_projection(orc.arccos_kernel(theta0, um, um, C))
    m_b, V_b = synth.make_variational_state(Kt_b, cfg)
    theta = dict(theta0)
    theta.update(dtheta or {})
    ref = orc.elbo_and_grad(theta, X, r, xt, n, B, m_b, V_b, prob.logA, prob.lambda0, return_moments=True)
    native.set_option(0, "i8_slices", i8)
    native.set_slab_rows(0, slab)
    native.invalidate_cache(0)
    try:
        out = native.elbo_grad(X.to(D), r.to(D), xt.to(D), n, theta, B.to(D), m_b.to(D), V_b.to(D), prob.logA,
                               prob.lambda0, want_moments=True, lims=(synth.THETA_LIMS_LO, synth.THETA_LIMS_HI))
    finally:
        native.set_option(0, "i8_slices", -1)
        native.set_slab_rows(0, 0)
    assert out["flags"] == 0 and out["p"] == ref["p"], (out["flags"], out["p"], ref["p"])
    assert abs(out["loglik"] - ref["loglik"]) <= RTOL * max(abs(ref["loglik"]), 1e-300)
    assert abs(out["KL"] - ref["KL"]) <= RTOL * abs(ref["KL"])
    for k in orc.KEYS:
        sc = max(abs(ref["dloglik"][k]), abs(ref["dKL"][k]), 1e-300)
        assert abs(out["grad"][k] - ref["grad"][k]) <= RTOL * sc, (k, out["grad"][k], ref["grad"][k], sc)
    assert torch.allclose(out["lam_m"].cpu(), ref["lam_m"], rtol=RTOL, atol=1e-12)
    assert torch.allclose(out["lam_var"].cpu(), ref["lam_var"], rtol=RTOL, atol=1e-10)
    return out, ref, B.shape[1]


def test_single_training_image():
    run(1, 8, 6, synth.THETA_DENSE, {"eps_0x": 0.02})


def test_two_inducing_images():
    run(37, 8, 2, synth.THETA_DENSE, {"Amp": 0.3})


@pytest.mark.parametrize("i8", [0, 8])
def test_mask_of_a_few_pixels(i8):
    """A very narrow receptive field: p = 5 pixels of 144 (and m = 30 > p: K_tilde is far from full rank in the pixel
    space, the arc-cosine nonlinearity keeps q well above p)."""
    th = {**synth.THETA_NOTEBOOK, "-2log2beta": 6.9, "-log2rho2": 5.0, "eps_0x": 0.09, "eps_0y": -0.09}
    out, ref, q = run(333, 12, 30, th, {"eps_0x": 0.094}, i8=i8, slab=128)
    assert out["p"] <= 9 and q <= 30


def test_empty_mask_is_refused():
    """No pixel reaches alpha >= 1e-3: the reference would index with an empty mask and fail inside the kernel function;
    the library refuses the evaluation with SGP_E_ARG and says why."""
    from gaussian_processes_b200 import _lib, native
    th = {**synth.THETA_NOTEBOOK, "-2log2beta": 6.9, "-log2rho2": 5.0, "eps_0x": 0.004, "eps_0y": -0.09}
    assert int(native.mask_host(th, 12).sum()) == 0
    prob = synth.make_problem(64, 12, 8, cfg=62, theta=synth.theta_localised(12))
    e8 = torch.eye(8, dtype=torch.float64, device=D)
    with pytest.raises(_lib.SgpError) as e:
        native.elbo_grad(prob.X.to(D), prob.r.to(D), prob.xtilde.to(D), 12, th, e8, e8[0].contiguous(), e8, 0.0, 0.0)
    assert "empty" in str(e.value)


def test_tiny_ntilde_equals_nt():
    run(5, 8, 5, synth.theta_localised(8), {"sigma_0": 0.1})


def test_ragged_image_count_multi_slab():
    """N = 1001 with 128-row slabs: 7 full slabs and a ragged one of 105 rows, int8 engine forced."""
    run(1001, 12, 40, synth.theta_localised(12), {"eps_0y": 0.01}, i8=8, slab=128)


def test_theta_on_its_limits_is_inside():
    """The reference's bounds check is inclusive (`lo <= value <= hi`, U:866): eps_0 = 1 and Amp at a tiny positive value
    are evaluated, not rejected."""
    th = dict(synth.THETA_DENSE)
    th.update({"eps_0x": 1.0, "eps_0y": -1.0})
    run(200, 8, 10, th)


def test_more_inducing_images_than_the_backward_kernel_supports_is_refused():
    from gaussian_processes_b200 import _lib, native
    prob = synth.make_problem(2100, 8, 2100, cfg=61, theta=synth.THETA_DENSE)
    X = prob.X.to(D)
    B = torch.eye(2100, dtype=torch.float64, device=D)[:, :8].contiguous()
    with pytest.raises(_lib.SgpError) as e:
        native.elbo_grad(X[:50], prob.r[:50].to(D), X, 8, prob.theta, B, torch.zeros(8, dtype=torch.float64, device=D),
                         torch.eye(8, dtype=torch.float64, device=D), 0.0, 0.0)
    assert e.value.code == _lib.SGP_E_UNSUPPORTED


def _state(prob, theta0, n):
    C, mask = orc.build_C(theta0, n)
    um = prob.xtilde[:, mask]
    B, Kt_b, _ = orc.eig_projection(orc.arccos_kernel(theta0, um, um, C))
    m_b, V_b = synth.make_variational_state(Kt_b, 63)
    return B.to(D), m_b.to(D), V_b.to(D)


@pytest.mark.parametrize("N,slab,i8", [(700, 0, 0), (9000, 4096, 8)])
def test_nan_in_the_images_is_flagged_and_does_not_outlive_the_call(N, slab, i8):
    """One NaN pixel: the evaluation comes back (no hang, no exception from the library) with the NaN status bit -- the
    reference raises on NaN in f_mean (U:1923), utils maps the bit to that -- and the NEXT call, on clean images of the same
    shape at the same address, is exactly what a fresh process computes: nothing of the poisoned call survives in the
    cached copies, the int8 slices or the work areas."""
    from gaussian_processes_b200 import _lib, native
    n, m = 12, 40
    th = synth.theta_localised(n)
    prob = synth.make_problem(N, n, m, cfg=63, theta=th)
    B, m_b, V_b = _state(prob, th, n)
    X, r, xt = prob.X.to(D), prob.r.to(D), prob.xtilde.to(D)
    native.set_option(0, "i8_slices", i8)
    native.set_slab_rows(0, slab)
    try:
        native.invalidate_cache(0)
        clean = native.elbo_grad(X, r, xt, n, th, B, m_b, V_b, prob.logA, prob.lambda0)
        assert clean["flags"] == 0
        mask = native.mask_host(th, n)
        px = int(torch.nonzero(mask)[3])
        X[N // 2, px] = float("nan")
        bad = native.elbo_grad(X, r, xt, n, th, B, m_b, V_b, prob.logA, prob.lambda0)
        assert bad["flags"] & _lib.FLAG_NAN
        X[N // 2, px] = prob.X[N // 2, px]
        again = native.elbo_grad(X, r, xt, n, th, B, m_b, V_b, prob.logA, prob.lambda0)
        assert again["flags"] == 0 and again["loglik"] == clean["loglik"] and again["KL"] == clean["KL"]
        assert all(again["grad"][k] == clean["grad"][k] for k in clean["grad"])
    finally:
        native.set_option(0, "i8_slices", -1)
        native.set_slab_rows(0, 0)
        native.invalidate_cache(0)


def test_the_handle_survives_refused_calls():
    """Out-of-bounds theta (the reference's ValueError, U:865-867), an empty mask and a bad argument are refused with their
    codes; the same handle then evaluates correctly."""
    from gaussian_processes_b200 import _lib, native
    n, m, N = 12, 30, 500
    th = synth.theta_localised(n)
    prob = synth.make_problem(N, n, m, cfg=64, theta=th)
    B, m_b, V_b = _state(prob, th, n)
    X, r, xt = prob.X.to(D), prob.r.to(D), prob.xtilde.to(D)
    lims = (synth.THETA_LIMS_LO, synth.THETA_LIMS_HI)
    ref = native.elbo_grad(X, r, xt, n, th, B, m_b, V_b, prob.logA, prob.lambda0, lims=lims)
    out_of_bounds = dict(th, eps_0x=1.5)
    with pytest.raises(_lib.SgpError) as e:
        native.elbo_grad(X, r, xt, n, out_of_bounds, B, m_b, V_b, prob.logA, prob.lambda0, lims=lims)
    assert e.value.code == _lib.SGP_E_BOUNDS
    empty = {**synth.THETA_NOTEBOOK, "-2log2beta": 6.9, "-log2rho2": 5.0, "eps_0x": 0.004, "eps_0y": -0.09}
    with pytest.raises(_lib.SgpError):
        native.elbo_grad(X, r, xt, n, empty, B, m_b, V_b, prob.logA, prob.lambda0)
    with pytest.raises(ValueError):
        native.elbo_grad(X, r, xt, n + 1, th, B, m_b, V_b, prob.logA, prob.lambda0)       # P != n_px^2: Python-side check
    with pytest.raises(ValueError):
        native.elbo_grad(X, r[:-1], xt, n, th, B, m_b, V_b, prob.logA, prob.lambda0)      # one spike count short
    with pytest.raises(ValueError):
        native.elbo_grad(X, r, xt, n, th, B, m_b[:-1], V_b, prob.logA, prob.lambda0)
    with pytest.raises(_lib.SgpError) as e:                                                # and the C ABI's own check
        import ctypes as C
        h = _lib.handle_for(0)
        out = (C.c_double * 16)()
        P_ = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
        h.check(_lib.lib.sgp_elbo_grad(h.ptr, P_(X), X.stride(0), P_(r), N, P_(xt), xt.stride(0), m, n + 1,
                                       native.theta_array(th), None, None, P_(B), B.stride(0), B.shape[1], P_(m_b), P_(V_b),
                                       V_b.stride(0), float(prob.logA), float(prob.lambda0), None, None, None, None, out,
                                       C.c_void_p(torch.cuda.current_stream(D).cuda_stream)))
    assert e.value.code == _lib.SGP_E_ARG
    out = native.elbo_grad(X, r, xt, n, th, B, m_b, V_b, prob.logA, prob.lambda0, lims=lims)
    assert out["flags"] == 0 and out["loglik"] == ref["loglik"]
    assert all(out["grad"][k] == ref["grad"][k] for k in ref["grad"])


def test_function_level_entry_points_refuse_bad_shapes():
    """Sizes <= 0 and leading dimensions smaller than the row length come back as SGP_E_ARG with a message, through the C
    ABI, before any kernel is launched (the buffers below would be overrun otherwise)."""
    import ctypes as C
    from gaussian_processes_b200 import _lib
    lib, h = _lib.lib, _lib.handle_for(0)
    a = torch.ones(64, 8, dtype=torch.float64, device=D)
    v = torch.ones(64, dtype=torch.float64, device=D)
    P = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
    st = C.c_void_p(torch.cuda.current_stream(D).cuda_stream)
    out = (C.c_double * 16)()
    calls = [
        lambda: lib.sgp_gemm_f64(h.ptr, 0, 0, P(a), 4, P(a), 8, P(a), 8, 8, 8, 8, 1.0, 0, 0, st),          # lda < K
        lambda: lib.sgp_gemm_f64(h.ptr, 0, 0, P(a), 8, P(a), 8, P(a), 8, 0, 8, 8, 1.0, 0, 0, st),          # M == 0
        lambda: lib.sgp_gemm_i8emu(h.ptr, 0, P(a), 8, 8, P(a), 8, 8, 8, 8, P(a), 4, st),                    # ldc < N
        lambda: lib.sgp_cholesky(h.ptr, P(a), 4, 8, None, st),                                              # ld < n
        lambda: lib.sgp_spd_inverse(h.ptr, P(a), 8, 8, P(a), 4, None, None, st),                            # ldi < n
        lambda: lib.sgp_acosker(h.ptr, P(a), 8, 4, P(a), 8, 8, 8, P(a), 8, 1.0, P(a), 8, 0, st),            # ld1 < p
        lambda: lib.sgp_acosker(h.ptr, P(a), 8, 8, P(a), 8, 8, 8, P(a), 8, 1.0, P(a), 4, 0, st),            # ldk < n2
        lambda: lib.sgp_acosker_diag(h.ptr, P(a), -1, 8, 8, P(a), 8, 1.0, P(v), st),                        # n1 < 0
        lambda: lib.sgp_acosker_known(h.ptr, P(a), 8, 8, P(v), P(a), 1, 4, 8, P(a), 8, 1.0, P(a), 8, st),   # ld2 < p
        lambda: lib.sgp_lambda_moments(h.ptr, P(a), 4, P(a), 8, P(v), 8, 8, P(v), P(a), 8, P(v), P(v), st),  # lda < q
        lambda: lib.sgp_estep_terms(h.ptr, P(a), 8, P(v), P(v), 8, 8, 0.0, P(v), P(a), 4, st),              # ldg < q
        lambda: lib.sgp_fparam_eval(h.ptr, P(v), P(v), P(v), 0, 0.0, 0.0, None, out, st),                   # N == 0
        lambda: lib.sgp_localker(h.ptr, (C.c_double * 6)(1, 0, 0, 1, 1, 1), 0, P(a), 8, None, 0, st),       # n_px == 0
    ]
    before = a.clone()
    for i, call in enumerate(calls):
        rc = call()
        assert rc == _lib.SGP_E_ARG, (i, rc)
        assert lib.sgp_last_error(h.ptr)
    torch.cuda.synchronize()
    assert torch.equal(a, before)
