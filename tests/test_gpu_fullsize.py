"""GPU: the fused ELBO+grad closure at BASELINE.json's full sizes, through properties that do not need the oracle
(which cannot finish these sizes in seconds):

* additivity over image blocks -- loglik and the six gradients of the whole image set equal the sums over disjoint
  shards evaluated with the same replicated state, the KL part counted once (this is also the multi-GPU contract,
  SURVEY 8e);
* the two GEMM engines (tcgen05 int8 slices / fp64 DMMA) and the two entry points (device-resident / host-streamed)
  agree at the parity tolerance;
* the gradient is the derivative of the value: a central difference of loglik - KL along a random direction in
  theta matches grad . d (grad = d(loglik - KL)/d theta).
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

RTOL = 1e-9
KEYS = ("sigma_0", "eps_0x", "eps_0y", "-2log2beta", "-log2rho2", "Amp")


def _problem(N, n, m, seed, theta=None):
    from gaussian_processes_b200 import native, synth
    D = torch.device("cuda:0")
    g = torch.Generator(device=D).manual_seed(seed)
    X = torch.randn(N, n * n, generator=g, dtype=torch.float64, device=D)
    lin = torch.linspace(-1, 1, n, dtype=torch.float64, device=D)
    w = torch.exp(-((lin[None, :] - 0.2) ** 2 + (lin[:, None] + 0.1) ** 2) / (2 * 0.15 ** 2)).reshape(-1)
    z = X @ w
    r = torch.poisson(torch.exp(0.8 * z / z.std() - 0.5), generator=g)
    xt = X[:m].clone()
    theta = dict(theta if theta is not None else synth.THETA_DENSE)
    Cm, mask = native.localker(theta, n, D)
    um = xt if bool(mask.all()) else xt[:, mask].contiguous()
    Kt = native.acosker(theta["sigma_0"], um, um, Cm, True)
    ev, evec = torch.linalg.eigh(Kt)
    keep = ev > max(float(ev.max()) * 1e-4, 1e-4)
    B = evec[:, keep].contiguous()
    q = B.shape[1]
    m_b = 0.1 * torch.randn(q, generator=g, dtype=torch.float64, device=D)
    W = torch.randn(q, q, generator=g, dtype=torch.float64, device=D)
    V_b = 0.5 * torch.diag(ev[keep]) + 0.01 * (W @ W.T)
    V_b = 0.5 * (V_b + V_b.T)
    return dict(X=X, r=r, xt=xt, n=n, theta=theta, B=B, m_b=m_b, V_b=V_b, logA=float(np.log(0.05)), lambda0=-0.5)


def _eval(P, rows=slice(None), theta=None, kl_weight=1.0):
    from gaussian_processes_b200 import native
    native.invalidate_cache(0)
    return native.elbo_grad(P["X"][rows], P["r"][rows], P["xt"], P["n"], theta or P["theta"], P["B"], P["m_b"],
                            P["V_b"], P["logA"], P["lambda0"], kl_weight=kl_weight)


def _gvec(out):
    return np.array([out["grad"][k] for k in KEYS])


def _assert_additive(whole, parts):
    ll = sum(p["loglik"] for p in parts)
    assert abs(ll - whole["loglik"]) <= 1e-11 * abs(whole["loglik"])
    g = sum(_gvec(p) for p in parts)
    assert np.abs(g - _gvec(whole)).max() <= 1e-10 * np.abs(_gvec(whole)).max()
    assert all(p["flags"] == 0 for p in parts) and whole["flags"] == 0


@pytest.fixture(scope="module")
def config2():
    return _problem(100000, 40, 500, seed=1002)


def test_config2_additive_over_image_blocks(config2):
    P = config2
    whole = _eval(P)
    cut = 37123                                   # not a multiple of any tile or slab size
    parts = [_eval(P, slice(0, cut)), _eval(P, slice(cut, None), kl_weight=0.0)]
    _assert_additive(whole, parts)
    assert parts[0]["KL"] == whole["KL"]


def test_config2_engines_and_entry_points_agree(config2):
    from gaussian_processes_b200 import native
    P = config2
    a = _eval(P)
    native.set_option(0, "i8_slices", 0)
    try:
        b = _eval(P)
    finally:
        native.set_option(0, "i8_slices", -1)
    assert abs(a["loglik"] - b["loglik"]) <= RTOL * abs(b["loglik"])
    assert np.abs(_gvec(a) - _gvec(b)).max() <= RTOL * np.abs(_gvec(b)).max()
    h = native.elbo_grad_host(P["X"].cpu().pin_memory(), P["r"].cpu(), P["xt"].cpu(), P["n"], P["theta"],
                              P["B"].cpu(), P["m_b"].cpu(), P["V_b"].cpu(), P["logA"], P["lambda0"])
    assert abs(a["loglik"] - h["loglik"]) <= RTOL * abs(a["loglik"])
    assert np.abs(_gvec(a) - _gvec(h)).max() <= RTOL * np.abs(_gvec(a)).max()


def test_config2_gradient_is_derivative_of_value(config2):
    P = config2
    out = _eval(P)
    rng = np.random.default_rng(7)
    d = rng.standard_normal(len(KEYS))
    d /= np.linalg.norm(d)
    h = 1e-5

    def elbo(sign):
        th = {k: P["theta"][k] + sign * h * d[i] for i, k in enumerate(KEYS)}
        o = _eval(P, theta=th)
        return o["loglik"] - o["KL"]
    fd = (elbo(+1) - elbo(-1)) / (2 * h)
    an = float(_gvec(out) @ d)
    assert abs(fd - an) <= 2e-6 * max(abs(an), np.abs(_gvec(out)).max())


def _engines_agree(P):
    from gaussian_processes_b200 import native
    a = _eval(P)
    native.set_option(0, "i8_slices", 0)
    try:
        b = _eval(P)
    finally:
        native.set_option(0, "i8_slices", -1)
    assert abs(a["loglik"] - b["loglik"]) <= RTOL * abs(b["loglik"])
    assert np.abs(_gvec(a) - _gvec(b)).max() <= RTOL * np.abs(_gvec(b)).max()
    return a


def test_config3_thousand_inducing_images():
    """BASELINE config 3 at its largest inducing set: N = 50 000, m = 1 000."""
    P = _problem(50000, 40, 1000, seed=1003)
    whole = _engines_agree(P)
    _assert_additive(whole, [_eval(P, slice(0, 20001)), _eval(P, slice(20001, None), kl_weight=0.0)])
    del P
    torch.cuda.empty_cache()


def test_config4_one_cell_localised_receptive_field():
    """One cell of BASELINE config 4: N = 200 000 shared images, receptive field off-centre so the mask crops the
    pixel grid (p < 1600) and the images go through the masked gather."""
    from gaussian_processes_b200 import synth
    th = synth.theta_localised(40)
    th.update({"eps_0x": 0.31, "eps_0y": -0.42})
    P = _problem(200000, 40, 500, seed=1004, theta=th)
    whole = _engines_agree(P)
    assert 0 < whole["p"] < 1600
    _assert_additive(whole, [_eval(P, slice(0, 131072)), _eval(P, slice(131072, None), kl_weight=0.0)])
    del P
    torch.cuda.empty_cache()


def test_config5_one_million_images_additive():
    """BASELINE config 5 (N = 1 000 000, 40x40, m = 500) on one GPU: 12.8 GB of images, ~42 GB with the cached slices."""
    P = _problem(1000000, 40, 500, seed=1005)
    whole = _eval(P)
    parts = [_eval(P, slice(0, 400000)), _eval(P, slice(400000, 700001), kl_weight=0.0),
             _eval(P, slice(700001, None), kl_weight=0.0)]
    _assert_additive(whole, parts)
    del P
    torch.cuda.empty_cache()
