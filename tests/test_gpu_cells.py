"""GPU: population fit (BASELINE configs[3]) -- sgp_elbo_grad_cells evaluates several independent cells that share the
stimuli in one call, all on the same resident images / int8 slices; it must give what per-cell calls of the fused
closure give (which tests/test_gpu_closure.py / test_gpu_bigsize.py pin to the reference)."""
import numpy as np
import pytest
import torch

from gaussian_processes_b200 import synth
from helpers import record

pytestmark = pytest.mark.gpu
D = torch.device("cuda:0")


def make_cells(N, n, m, ncells, base_theta, eps, seed=4000):
    """`ncells` cells on shared images: own receptive-field centre (eps[c]), own spikes, own subspace / variational state."""
    from gaussian_processes_b200 import native
    prob = synth.make_problem(N, n, m, cfg=40, theta=base_theta)
    X, xt = prob.X.to(D), prob.xtilde.to(D)
    g = torch.Generator().manual_seed(seed)
    cells = []
    for c in range(ncells):
        th = dict(base_theta)
        th["eps_0x"], th["eps_0y"] = eps[c]
        Cm, mask = native.localker(th, n, D)
        um = xt[:, mask].contiguous()
        Kt = native.acosker(th["sigma_0"], um, um, Cm, True)
        ev, evec = torch.linalg.eigh(Kt, UPLO="L")
        keep = ev > max(float(ev.max()) * 1e-4, 1e-4)
        B = evec[:, keep].contiguous()
        m_b, V_b = synth.make_variational_state(torch.diag(ev[keep]).cpu(), 40 + c)
        r = torch.poisson(torch.full((N,), 0.6 + 0.2 * c, dtype=torch.float64), generator=g)
        # evaluate away from the theta the subspace was built at (frozen B)
        th_eval = dict(th)
        th_eval["-log2rho2"] += 0.02 * (c + 1)
        cells.append(dict(theta=th_eval, B=B, m_b=m_b.to(D), V_b=V_b.to(D), r=r.to(D), logA=np.log(0.05) + 0.01 * c,
                          lambda0=-0.5 - 0.05 * c, p=int(native.mask_host(th_eval, n).sum())))
    return prob, X, xt, cells


@pytest.mark.parametrize("cfg", [dict(N=2000, n=20, m=100, i8=8, slab=512), dict(N=8192, n=40, m=200, i8=-1, slab=4096),
                                 dict(N=1500, n=20, m=64, i8=0, slab=0)])
def test_cells_equal_single_cell_calls(cfg):
    from gaussian_processes_b200 import native
    # a narrower RF than THETA_DENSE so that off-centre cells crop the grid (p < P) while the central one keeps it
    base = dict(synth.THETA_DENSE)
    base["-2log2beta"] = 1.9
    eps = [(0.0, 0.0), (0.45, -0.4), (-0.5, 0.5), (0.2, 0.3)]
    prob, X, xt, cells = make_cells(cfg["N"], cfg["n"], cfg["m"], 4, base, eps)
    P = cfg["n"] ** 2
    assert any(c["p"] < P for c in cells) and len({c["p"] for c in cells}) > 1, [c["p"] for c in cells]
    native.set_option(0, "i8_slices", cfg["i8"])
    native.set_option(0, "path", 2)
    native.set_slab_rows(0, cfg["slab"])
    try:
        native.invalidate_cache(0)
        singles = [native.elbo_grad(X, c["r"], xt, cfg["n"], c["theta"], c["B"], c["m_b"], c["V_b"], c["logA"],
                                    c["lambda0"]) for c in cells]
        native.invalidate_cache(0)
        native.launch_count(0, reset=True)
        batch = native.elbo_grad_cells(X, [c["r"] for c in cells], xt, cfg["n"], [c["theta"] for c in cells],
                                       [c["B"] for c in cells], [c["m_b"] for c in cells], [c["V_b"] for c in cells],
                                       [c["logA"] for c in cells], [c["lambda0"] for c in cells])
        again = native.elbo_grad_cells(X, [c["r"] for c in cells], xt, cfg["n"], [c["theta"] for c in cells],
                                       [c["B"] for c in cells], [c["m_b"] for c in cells], [c["V_b"] for c in cells],
                                       [c["logA"] for c in cells], [c["lambda0"] for c in cells])
    finally:
        native.set_option(0, "i8_slices", -1)
        native.set_option(0, "path", 0)
        native.set_slab_rows(0, 0)
        native.invalidate_cache(0)
    worst_ll = worst_g = 0.0
    for c, s, b, a in zip(cells, singles, batch, again):
        assert b["flags"] == 0 and s["flags"] == 0
        assert b["p"] == c["p"] == s["p"]
        assert b["loglik"] == a["loglik"] and b["grad"] == a["grad"]        # back-to-back cells are deterministic
        worst_ll = max(worst_ll, abs(b["loglik"] - s["loglik"]) / abs(s["loglik"]))
        assert abs(b["KL"] - s["KL"]) <= 1e-12 * abs(s["KL"])
        gmax = max(abs(v) for v in s["grad"].values())
        worst_g = max(worst_g, max(abs(b["grad"][k] - s["grad"][k]) for k in s["grad"]) / gmax)
    record(f"cells_vs_single/n{cfg['n']}_i8_{cfg['i8']}", loglik=worst_ll, grad=worst_g)
    # two valid evaluations of the same numbers (whole-grid support vs mask support: different zero padding and summation
    # order in the products): measured 7e-16 / 1e-12, a thousand times inside the parity bar
    assert worst_ll <= 1e-12 and worst_g <= 1e-11


def test_cells_argument_checks():
    from gaussian_processes_b200 import native
    base = dict(synth.THETA_DENSE)
    prob, X, xt, cells = make_cells(600, 12, 32, 2, base, [(0.0, 0.0), (0.1, 0.1)])
    with pytest.raises(ValueError):
        native.elbo_grad_cells(X, [cells[0]["r"]], xt, 12, [c["theta"] for c in cells], [c["B"] for c in cells],
                               [c["m_b"] for c in cells], [c["V_b"] for c in cells], [0.0, 0.0], [0.0, 0.0])
    with pytest.raises(RuntimeError):
        native.elbo_grad_cells(X.cpu(), [c["r"] for c in cells], xt, 12, [c["theta"] for c in cells],
                               [c["B"] for c in cells], [c["m_b"] for c in cells], [c["V_b"] for c in cells],
                               [0.0, 0.0], [0.0, 0.0])


def test_cells_at_config4_size_against_the_reference():
    """BASELINE config 4's shape against the REAL reference (oracle/gen_golden.py config4): three cells of a population on
    N = 200 000 shared stimuli of 40 x 40, m = 500 -- own receptive-field centre (all three crop the grid: p = 1240 /
    883 / 802), spikes, subspace and state each; the reference's closure per cell on two 100 000-image blocks.  One sgp_elbo_grad_cells call."""
    from gaussian_processes_b200 import native
    from helpers import load
    fx = load("cells_c4_full")
    N, n, m, nc = int(fx["N"]), int(fx["n"]), int(fx["m"]), int(fx["ncells"])
    keys = [str(k) for k in fx["theta_keys"]]
    prob = synth.make_problem(N, n, m, cfg=int(fx["cfg"]), theta=synth.THETA_DENSE)
    X, xt = prob.X.to(D), prob.xtilde.to(D)
    g = torch.Generator().manual_seed(int(fx["spike_seed"]))
    cells = []
    for c in range(nc):
        ev = torch.from_numpy(fx[f"eigvals_{c}"])[torch.from_numpy(fx[f"keep_{c}"])]
        m_b, V_b = synth.make_variational_state(torch.diag(ev), 40 + c)
        got = np.array([float(V_b[0, 0]), float(V_b[-1, -1]), float(V_b[0, -1]), float(V_b.sum())])
        assert np.allclose(got, fx[f"V_b_guard_{c}"], rtol=1e-13, atol=0) and np.array_equal(m_b.numpy(), fx[f"m_b_{c}"])
        r = torch.poisson(torch.full((N,), 0.6 + 0.2 * c, dtype=torch.float64), generator=g)
        assert float(r.sum()) == float(fx[f"r_sum_{c}"])          # the seeded spike trains are the reference's
        cells.append(dict(theta=dict(zip(keys, fx[f"theta_{c}"].tolist())), B=torch.from_numpy(fx[f"B_{c}"]).to(D),
                          m_b=m_b.to(D), V_b=V_b.to(D), r=r.to(D), logA=float(fx[f"logA_{c}"]),
                          lambda0=float(fx[f"lambda0_{c}"])))
    native.invalidate_cache(0)
    try:
        outs = native.elbo_grad_cells(X, [c["r"] for c in cells], xt, n, [c["theta"] for c in cells],
                                      [c["B"] for c in cells], [c["m_b"] for c in cells], [c["V_b"] for c in cells],
                                      [c["logA"] for c in cells], [c["lambda0"] for c in cells])
        ps = []
        for c, o in enumerate(outs):
            dl, dk = fx[f"dloglik_{c}"], fx[f"dKL_{c}"]
            gerr = max(abs(o["grad"][k] - (dl[i] - dk[i])) / max(abs(dl[i]), abs(dk[i])) for i, k in enumerate(keys))
            e_ll = abs(o["loglik"] - float(fx[f"loglik_{c}"])) / abs(float(fx[f"loglik_{c}"]))
            e_kl = abs(o["KL"] - float(fx[f"KL_{c}"])) / abs(float(fx[f"KL_{c}"]))
            record(f"cells_c4_full/cell{c}", loglik=e_ll, KL=e_kl, grad_comp=gerr, p=o["p"])
            assert o["flags"] == 0 and o["p"] == int(fx[f"p_{c}"])
            # measured 1e-10 .. 7e-10 per gradient component: the largest is eps_0x of the cell centred at (0, 0), a sum over
            # 200 000 images that cancels to ~10 by symmetry (spike trains independent of the stimuli); fixed-order
            # reductions make the figure reproducible
            assert e_ll <= 1e-9 and e_kl <= 1e-9 and gerr <= 1e-9, (c, e_ll, e_kl, gerr)
            ps.append(o["p"])
        assert min(ps) < n * n                                     # at least one cell's mask crops the grid
        # the same cells one by one, with the per-image moments (every 200th image is in the fixture)
        for c, cell in enumerate(cells):
            o = native.elbo_grad(X, cell["r"], xt, n, cell["theta"], cell["B"], cell["m_b"], cell["V_b"], cell["logA"],
                                 cell["lambda0"], want_moments=True)
            lm, lv = o["lam_m"][::200].cpu().numpy(), o["lam_var"][::200].cpu().numpy()
            rm, rv = fx[f"lam_m_every_200_{c}"], fx[f"lam_var_every_200_{c}"]
            assert np.allclose(lm, rm, rtol=1e-9, atol=1e-11 * np.abs(rm).max())
            assert np.allclose(lv, rv, rtol=1e-9, atol=1e-10)
            assert abs(o["loglik"] - outs[c]["loglik"]) <= 1e-12 * abs(o["loglik"])
    finally:
        del X
        native.release_cache(0)
        torch.cuda.empty_cache()
