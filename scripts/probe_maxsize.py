"""GPU probe: BASELINE config 5 (10^6 images of 40 x 40, m = 500) RESIDENT on one GPU -- 64-bit indexing everywhere -- against
the sum of eight 125 000-image evaluations (the ELBO's likelihood part and its gradient are sums over images; KL counted
once through kl_weight).  Images are generated on the device."""
import os
import sys
import time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from gaussian_processes_b200 import native, synth  # noqa: E402

dev = torch.device("cuda:0")
N, n, m, G = int(os.environ.get("N", 1000000)), 40, 500, 8
g = torch.Generator(device=dev).manual_seed(5)
X = torch.randn(N, n * n, generator=g, dtype=torch.float64, device=dev)
prob = synth.make_problem(4096, n, m, cfg=5, theta=synth.THETA_DENSE)
xt = X[:m].clone()
w = prob.X[:, :].new_zeros(1)
r = torch.poisson(torch.full((N,), 0.8, dtype=torch.float64, device=dev), generator=g)
theta = dict(prob.theta)
Cm, mask = native.localker(theta, n, dev)
Kt = native.acosker(theta["sigma_0"], xt, xt, Cm, True)
ev, evec = torch.linalg.eigh(Kt, UPLO="L")
keep = ev > max(float(ev.max()) * 1e-4, 1e-4)
B = evec[:, keep].contiguous()
m_b, V_b = synth.make_variational_state(torch.diag(ev[keep]).cpu(), 5)
m_b, V_b = m_b.to(dev), V_b.to(dev)
theta["eps_0x"] += 0.01
torch.cuda.synchronize()
t0 = time.perf_counter()
whole = native.elbo_grad(X, r, xt, n, theta, B, m_b, V_b, prob.logA, prob.lambda0, want_moments=True)
torch.cuda.synchronize()
t1 = time.perf_counter()
whole2 = native.elbo_grad(X, r, xt, n, theta, B, m_b, V_b, prob.logA, prob.lambda0)
torch.cuda.synchronize()
t2 = time.perf_counter()
print(f"N = {N}: first call {t1 - t0:.3f} s, second {1e3 * (t2 - t1):.1f} ms, flags {whole['flags']}, "
      f"memory in use {torch.cuda.mem_get_info(0)[1] / 2**30 - torch.cuda.mem_get_info(0)[0] / 2**30:.1f} GiB")
native.release_cache(0)
ll, gsum = 0.0, {k: 0.0 for k in native.THETA_KEYS}
lam_m = []
for s in range(G):
    lo, hi = s * N // G, (s + 1) * N // G
    o = native.elbo_grad(X[lo:hi], r[lo:hi], xt, n, theta, B, m_b, V_b, prob.logA, prob.lambda0, want_moments=True,
                         kl_weight=1.0 if s == 0 else 0.0, same=False)
    ll += o["loglik"]
    lam_m.append(o["lam_m"])
    for k in gsum:
        gsum[k] += o["grad"][k]
gmax = max(abs(v) for v in whole["grad"].values())
e_ll = abs(ll - whole["loglik"]) / abs(whole["loglik"])
e_g = max(abs(gsum[k] - whole["grad"][k]) for k in gsum) / gmax
e_m = float((torch.cat(lam_m) - whole["lam_m"]).abs().max() / whole["lam_m"].abs().max())
same2 = whole2["loglik"] == whole["loglik"]
print(f"whole vs sum of {G} shards: loglik {e_ll:.2e}, grad {e_g:.2e}, lam_m {e_m:.2e}; repeat bit-identical: {same2}")
ok = e_ll < 1e-12 and e_g < 1e-10 and e_m < 1e-12 and whole["flags"] == 0 and same2
print("OK" if ok else "MISMATCH")
sys.exit(0 if ok else 1)
