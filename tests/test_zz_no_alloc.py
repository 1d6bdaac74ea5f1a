"""GPU: sgp_cholesky_noalloc (scratch in a file-scope `static __device__` array, no cudaMalloc) against the existing entry
point sgp_cholesky, through the C ABI, on two calls in a row.  No counterpart in the reference: the entry point was added
for callers that cannot allocate (CUDA graph capture, serving loops).

Collected last on purpose (the file name sorts after every other test file): a
failure here must not hide other tests.  Passes alone in a fresh process and after the rest of the suite: it uses no
handle state, and the static array's status word is cleared by every call."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
RTOL = 1e-9          # the parity bar of the GPU tests (north_star: 1e-9 relative in fp64)
D = "cuda:0"


def spd(n, seed, ld=None):
    """Well-conditioned SPD matrix in a row-major buffer with an even leading dimension."""
    g = torch.Generator().manual_seed(seed)
    W = torch.randn(n, n, generator=g, dtype=torch.float64)
    A = W @ W.T / n + torch.eye(n, dtype=torch.float64) * 0.5
    ld = ld or (n + 1) // 2 * 2
    buf = torch.zeros(n, ld, dtype=torch.float64)
    buf[:, :n] = A
    return A, buf


def noalloc(buf_dev, n, want_flags=True):
    from gaussian_processes_b200 import _lib
    flags = torch.full((1,), -1, dtype=torch.int32, device=buf_dev.device) if want_flags else None
    err = C.create_string_buffer(256)
    torch.cuda.set_device(buf_dev.device)
    rc = _lib.lib.sgp_cholesky_noalloc(C.c_void_p(buf_dev.data_ptr()), buf_dev.stride(0), int(n),
                                       C.c_void_p(flags.data_ptr()) if want_flags else None,
                                       C.c_void_p(torch.cuda.current_stream(buf_dev.device).cuda_stream), err, 256)
    return rc, (int(flags.item()) if want_flags else None), err.value.decode()


def compare(n, seed):
    from gaussian_processes_b200 import native
    A, buf = spd(n, seed)
    L_ref, fl_ref = native.cholesky(A.to(D))                       # existing entry point: sgp_cholesky
    work = buf.to(D)
    rc, fl, msg = noalloc(work, n)
    assert rc == 0, msg
    torch.cuda.synchronize()
    L = torch.tril(work[:, :n])
    assert fl == 0 and fl_ref == 0
    err = float((L - L_ref).abs().max() / L_ref.abs().max())
    assert err <= RTOL, err
    # and it is a Cholesky factor of A
    rec = float((L @ L.T - A.to(D)).abs().max() / A.abs().max())
    assert rec <= 1e-12, rec
    return err


def test_two_calls_in_a_row_match_the_allocating_entry_point():
    e1 = compare(300, 1)          # 5 diagonal blocks, ragged last block
    e2 = compare(1000, 2)         # 16 blocks: the whole static array is in use
    e3 = compare(64, 3)           # a single block after a large one: stale block inverses must not matter
    assert max(e1, e2, e3) <= RTOL


def test_status_word_is_cleared_between_calls():
    """The static array keeps the last call's contents: a failed factorisation (status bit 1) must not leak into the
    next call's status."""
    A, buf = spd(200, 4)
    bad = buf.clone()
    bad[150, 150] = -5.0                                   # not positive definite
    rc, fl, msg = noalloc(bad.to(D), 200)
    assert rc == 0 and (fl & 1) == 1, (rc, fl, msg)
    rc, fl, msg = noalloc(buf.to(D), 200)
    assert rc == 0 and fl == 0, (rc, fl, msg)
    rc, fl, msg = noalloc(buf.to(D), 200, want_flags=False)   # flags pointer is optional
    assert rc == 0, msg


def test_sizes_beyond_the_maximum_are_rejected_with_a_clear_error():
    from gaussian_processes_b200 import _lib
    A, buf = spd(8, 5)
    work = buf.to(D)
    before = work.clone()
    err = C.create_string_buffer(256)
    rc = _lib.lib.sgp_cholesky_noalloc(C.c_void_p(work.data_ptr()), 1026, 1025, None, None, err, 256)
    assert rc == _lib.SGP_E_UNSUPPORTED
    assert "1024" in err.value.decode() and "maximum" in err.value.decode()
    rc = _lib.lib.sgp_cholesky_noalloc(C.c_void_p(work.data_ptr()), 7, 8, None, None, err, 256)     # odd ld < n
    assert rc == _lib.SGP_E_ARG and err.value
    rc = _lib.lib.sgp_cholesky_noalloc(None, 8, 8, None, None, None, 0)                               # no error buffer
    assert rc == _lib.SGP_E_ARG
    torch.cuda.synchronize()
    assert torch.equal(work, before)                       # rejected calls touch nothing


def test_the_call_allocates_no_device_memory():
    """Driver-level check: free device memory is the same before and after a (warmed-up) call."""
    A, buf = spd(512, 6)
    work = buf.to(D)
    noalloc(work, 512)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info(0)
    for _ in range(3):
        work.copy_(buf)
        rc, fl, msg = noalloc(work, 512)
        assert rc == 0 and fl == 0, msg
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info(0)
    assert free1 == free0
    assert np.isfinite(work.cpu().numpy()).all()
