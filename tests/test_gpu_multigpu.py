"""GPU (needs >= 2 devices; skipped on a 1-GPU box): the three exchange variants of the image-sharded evaluation agree
with one evaluation of all the images on one GPU (tests/multigpu_worker.py under torchrun)."""
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("local", [0, 1])
def test_image_shards_exchange_variants(local):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    env = dict(os.environ, MG_N="12000", MG_M="128")
    if local:
        env["MG_LOCAL"] = "1"
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", str(29650 + local), os.path.join(ROOT, "tests", "multigpu_worker.py")]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    rep = json.loads([ln for ln in p.stdout.splitlines() if ln.startswith("{")][-1])
    assert rep["ok"] and rep["peer_memory_identical_on_all_ranks"]
