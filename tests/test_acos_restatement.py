"""CPU check of the rational arc-cosine used by the device helpers (csrc/sgp_common.cuh: acos_nr).

The device code cannot run here; this restates its arithmetic in numpy fp64 (same coefficients, same branches, same
association) and compares with numpy.arccos, so a mistyped coefficient or a wrong branch formula fails on the CPU
before it reaches the GPU parity tests."""
import re
from pathlib import Path

import numpy as np

SRC = Path(__file__).resolve().parents[1] / "gaussian_processes_b200" / "csrc" / "sgp_common.cuh"


def _coefficients():
    txt = SRC.read_text()
    def arr(name):
        body = re.search(name + r"\[\d+\] = \{([^}]*)\}", txt).group(1)
        return [float(v) for v in body.replace("\n", " ").split(",")]
    hi = float(re.search(r"kPio2Hi = ([0-9.e+-]+)", txt).group(1))
    lo = float(re.search(r"kPio2Lo = ([0-9.e+-]+)", txt).group(1))
    pi = float(re.search(r"kPiTrue = ([0-9.e+-]+)", txt).group(1))
    return arr("kAcosP"), arr("kAcosQ"), hi, lo, pi


def acos_nr(c):
    P, Q, hi, lo, pi = _coefficients()
    a = np.abs(c)
    big = a >= 0.5
    z = np.where(big, 0.5 * (1.0 - a), c * c)
    pz = z * P[5] + P[4]
    for k in (3, 2, 1, 0):
        pz = z * pz + P[k]
    pz = pz * z
    qz = z * Q[3] + Q[2]
    qz = z * qz + Q[1]
    qz = z * qz + Q[0]
    qz = z * qz + 1.0
    r = pz / qz
    s = np.sqrt(z)
    small = hi - (c - (lo - c * r))
    bigp = 2.0 * (s * r + s)
    bign = pi - 2.0 * (s + (r * s - lo))
    return np.where(big, np.where(c > 0.0, bigp, bign), small)


def test_coefficients_parse():
    P, Q, hi, lo, pi = _coefficients()
    assert len(P) == 6 and len(Q) == 4
    assert hi + lo == np.pi / 2 and pi == np.pi


def test_acos_restatement_within_two_ulp():
    rng = np.random.default_rng(0)
    c = np.concatenate([rng.uniform(-1, 1, 2_000_000),
                        1 - 10 ** rng.uniform(-16, 0, 200_000), -1 + 10 ** rng.uniform(-16, 0, 200_000),
                        np.array([1.0, -1.0, 0.0, 0.5, -0.5, np.nextafter(0.5, 0), np.nextafter(-0.5, 0)])])
    c = np.clip(c, -1.0, 1.0)
    ref = np.arccos(c)
    got = acos_nr(c)
    ulp = np.spacing(np.maximum(ref, np.finfo(float).tiny))
    assert np.all(np.isfinite(got))
    assert (np.abs(got - ref) / ulp).max() <= 2.0
    assert got[c == 1.0].max() == 0.0
