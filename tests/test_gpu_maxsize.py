"""GPU: BASELINE config 5's image count (10^6 images of 40 x 40, m = 500; 12.8 GB of images, ~39 GiB with the cached int8
slices) RESIDENT on one GPU.  Size-independent property: the likelihood part of the ELBO and its gradient are sums over
images, so one evaluation of all the images equals the sum of eight 125 000-image evaluations (KL counted once) up to the
summation order -- which exercises the 64-bit indexing of the kernels (element offsets row * ld pass 2^31 at image 1 342 178 for
ld = 1600; byte offsets into the images and the slice buffers pass 2^32 at image 335 545)."""
import subprocess
import sys
import os

import pytest
import torch

pytestmark = pytest.mark.gpu


def test_one_million_images_resident_equals_the_sum_of_its_shards():
    free, _ = torch.cuda.mem_get_info(0)
    if free < (70 << 30):
        pytest.skip("needs 70 GiB of free device memory")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    # own process: 39 GiB of library caches should not stay behind in the test session's handle
    res = subprocess.run([sys.executable, os.path.join(root, "scripts", "probe_maxsize.py")], cwd=root,
                         capture_output=True, text=True, timeout=900)
    assert res.returncode == 0 and "OK" in res.stdout, res.stdout[-2000:] + res.stderr[-2000:]
