"""Shared test helpers: load golden fixtures, rebuild their seeded inputs."""
import os

import numpy as np
import torch

from gaussian_processes_b200 import synth

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CLOSURE_CASES = ["closure_dense_n8", "closure_local_n12", "closure_rankdef_n10", "closure_square_n8",
                 "closure_c1_dense", "closure_c1_local"]
# fixtures at the sizes bench.py times (n = 40; m = 500 / 1000): outputs of the real reference, B stored, V_b regenerated
BIG_CASES = ["closure_c2_dense_n40", "closure_c2_local_n40", "closure_c3_m1000_n40"]


def load(name):
    return dict(np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False))


def case_inputs(fx):
    """Rebuild the seeded inputs of a closure fixture (X, r, xtilde regenerated from the seeds)."""
    N, n, m, cfg = int(fx["N"]), int(fx["n"]), int(fx["m"]), int(fx["cfg"])
    keys = [str(k) for k in fx["theta_keys"]]
    prob = synth.make_problem(N, n, m, cfg=cfg, theta=dict(zip(keys, fx["theta"].tolist())))
    if "V_b" in fx:
        V_b = torch.from_numpy(fx["V_b"])
    else:   # big fixtures: V_b regenerated from the stored eigenvalues and the seed (oracle/gen_golden.py gen_big_closure_case)
        ev = torch.from_numpy(fx["eigvals"])[torch.from_numpy(fx["keep"])]
        m_chk, V_b = synth.make_variational_state(torch.diag(ev), cfg)
        g = fx["V_b_guard"]
        got = np.array([float(V_b[0, 0]), float(V_b[-1, -1]), float(V_b[0, -1]), float(V_b.sum())])
        assert np.allclose(got, g, rtol=1e-13, atol=0), "seeded generator drifted: V_b cannot be regenerated"
        assert np.array_equal(m_chk.numpy(), fx["m_b"])
    st = dict(B=torch.from_numpy(fx["B"]), m_b=torch.from_numpy(fx["m_b"]), V_b=V_b,
              logA=float(fx["logA"]), lambda0=float(fx["lambda0"]))
    return prob, st


def rel_err(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))
