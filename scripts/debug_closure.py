"""Compare every intermediate of sgp_elbo_grad with the CPU adjoint model for one fixture (GPU debugging aid)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import adjoint_model as adj
from helpers import case_inputs, load
from gaussian_processes_b200 import native

name = sys.argv[1] if len(sys.argv) > 1 else "closure_rankdef_n10"
fx = load(name); prob, st = case_inputs(fx)
d = torch.device("cuda:0")
out = native.elbo_grad(prob.X.to(d), prob.r.to(d), prob.xtilde.to(d), prob.n_px_side, prob.theta, st["B"].to(d),
                       st["m_b"].to(d), st["V_b"].to(d), st["logA"], st["lambda0"], want_moments=True)
ref = adj.elbo_and_grad(prob.theta, prob.X, prob.r, prob.xtilde, prob.n_px_side, st["B"], st["m_b"], st["V_b"],
                        st["logA"], st["lambda0"])
I = ref["inter"]
print("loglik", out["loglik"], ref["loglik"], "KL", out["KL"], ref["KL"], "flags", out["flags"])
pairs = [("C", "C"), ("X2", "X2"), ("KtRaw", "KtRaw"), ("Kt", "Kt"), ("Ktb", "Kt_b"), ("Kinv", "Kinv"), ("bb", "bb"),
         ("Mfull", "Mfull"), ("K_last", "K"), ("y", "y"), ("beta", "beta"), ("Sig2", "Sig2"), ("GKtb", "GKt_b"),
         ("Y_last", "Y")]
for gname, rname in pairs:
    g = native.debug_get(d, gname).cpu()
    r = I[rname]
    if r.dim() == 1: r = r[None, :]
    if g.shape != r.shape:
        print(f"{gname:8s} shape mismatch {tuple(g.shape)} vs {tuple(r.shape)}"); continue
    err = (g - r).abs().max().item() / max(r.abs().max().item(), 1e-300)
    print(f"{gname:8s} relerr {err:.3e}  (|ref|max {r.abs().max().item():.3e})")
