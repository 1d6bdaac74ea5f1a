"""Shared test helpers: load golden fixtures, rebuild their seeded inputs."""
import os

import numpy as np
import torch

from gaussian_processes_b200 import synth

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CLOSURE_CASES = ["closure_dense_n8", "closure_local_n12", "closure_rankdef_n10", "closure_square_n8",
                 "closure_c1_dense", "closure_c1_local", "closure_square_rankdef_n10", "closure_square_local_n12",
                 "closure_square_n20_m300"]
# ntilde == nt (K = K_tilde, a = B: the configuration the reference's active-learning notebook always runs)
SQUARE_CASES = ["closure_square_n8", "closure_square_rankdef_n10", "closure_square_local_n12", "closure_square_n20_m300"]
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


def record(name, **values):
    """Append measured errors to gpurun_out/parity_report.jsonl (scratch; read back after a GPU run to see how far
    below the asserted tolerance each comparison actually is)."""
    import json
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    try:
        os.makedirs(os.path.join(root, "gpurun_out"), exist_ok=True)
        with open(os.path.join(root, "gpurun_out", "parity_report.jsonl"), "a") as f:
            f.write(json.dumps({"test": name, **{k: float(v) for k, v in values.items()}}) + "\n")
    except Exception:
        pass


def to_device(obj, dev):
    """Recursively move the tensors of a (nested) fit_model dict / tuple / list to `dev` (requires_grad preserved)."""
    if torch.is_tensor(obj):
        t = obj.detach().to(dev)
        return t.requires_grad_(True) if obj.requires_grad else t
    if isinstance(obj, dict):
        return {k: to_device(v, dev) for k, v in obj.items()}
    if isinstance(obj, tuple):
        return tuple(to_device(v, dev) for v in obj)
    if isinstance(obj, list):
        return [to_device(v, dev) for v in obj]
    return obj


def reference_saved_model_dir():
    return os.path.join(GOLDEN, "test_small_n10_saved_by_reference")
