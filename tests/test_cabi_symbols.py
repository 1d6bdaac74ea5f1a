"""CPU: the C-ABI library loads and exports every symbol include/sgp_b200.h declares; host-only entry
points (mask) agree with the oracle; compute entry points refuse CPU tensors (no CPU fallback)."""
import os
import re

import pytest
import torch

import oracle.spatial_gp_oracle as orc
from gaussian_processes_b200 import _lib, native, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "sgp_b200.h")).read()
    return sorted(set(re.findall(r"SGP_API\s+[\w\s\*]+?\b(sgp_\w+)\s*\(", src)))


def test_every_declared_symbol_is_exported_and_bound():
    names = declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(_lib.lib, n), f"{n} declared in sgp_b200.h but not exported by libsgp_b200.so"
        assert n in _lib.SIGNATURES, f"{n} has no ctypes signature"
    assert _lib.lib.sgp_version() >= 100


@pytest.mark.parametrize("n,theta", [(8, synth.THETA_DENSE), (12, synth.theta_localised(12)),
                                     (20, synth.theta_localised(20)), (40, synth.theta_localised(40)),
                                     (40, synth.THETA_NOTEBOOK), (40, synth.THETA_DENSE), (27, synth.THETA_NOTEBOOK),
                                     (108, synth.THETA_NOTEBOOK)])
def test_host_mask_matches_oracle(n, theta):
    _, mask = orc.build_C(theta, n)
    got = native.mask_host(theta, n)
    assert torch.equal(got, mask)


def test_cpu_tensors_are_refused():
    x = torch.zeros(4, 4, dtype=torch.float64)
    with pytest.raises(RuntimeError):
        native.acosker(1.0, x, x, x, False)
    with pytest.raises(RuntimeError):
        native.elbo_grad(x, x[0], x, 2, synth.THETA_DENSE, x, x[0], x, 0.0, 0.0)


def test_handle_create_destroy_without_gpu():
    h = _lib.Handle(0)
    assert h.launch_count() == 0
    h.close()
