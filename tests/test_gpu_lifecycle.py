"""GPU: handle lifecycle and re-entrancy (SURVEY 8b "Threading / device" and "Ownership" rows).

* one notebook variant (one_cell_active_training.ipynb cell 19) calls acosker / lambda_moments / compute_loglikelihood from
  a second Python thread while varGP runs: entry points are re-entrant per handle, handles are per (thread, device);
* the .so owns only scratch: creating, using and destroying handles returns every byte to the driver."""
import threading

import pytest
import torch

from gaussian_processes_b200 import synth

pytestmark = pytest.mark.gpu
D = torch.device("cuda:0")


def problem(N=6000, n=20, m=120, cfg=70):
    from gaussian_processes_b200 import native
    prob = synth.make_problem(N, n, m, cfg=cfg, theta=synth.theta_localised(n))
    X, r, xt = prob.X.to(D), prob.r.to(D), prob.xtilde.to(D)
    Cm, mask = native.localker(prob.theta, n, D)
    um = xt[:, mask].contiguous()
    Kt = native.acosker(prob.theta["sigma_0"], um, um, Cm, True)
    ev, evec = torch.linalg.eigh(Kt, UPLO="L")
    keep = ev > max(float(ev.max()) * 1e-4, 1e-4)
    B = evec[:, keep].contiguous()
    m_b, V_b = synth.make_variational_state(torch.diag(ev[keep]).cpu(), cfg)
    return prob, X, r, xt, Cm, mask, B, m_b.to(D), V_b.to(D)


def test_second_thread_calls_function_level_entry_points_while_the_closure_runs():
    from gaussian_processes_b200 import _lib, native
    prob, X, r, xt, Cm, mask, B, m_b, V_b = problem()
    xm, um = X[:, mask].contiguous(), xt[:, mask].contiguous()
    s0 = prob.theta["sigma_0"]

    def closure(k):
        th = dict(prob.theta)
        th["eps_0x"] += 0.01 * (k % 3)
        return native.elbo_grad(X, r, xt, prob.n_px_side, th, B, m_b, V_b, prob.logA, prob.lambda0)

    def side(k):
        K, Kvec = native.acosker_kvec(s0, xm[: 2000 + 100 * (k % 3)], um, Cm, False)
        return float(K.sum()), float(Kvec.sum())
    ref_main = [closure(k) for k in range(3)]
    ref_side = [side(k) for k in range(3)]
    torch.cuda.synchronize()
    errors, got_side, handles = [], [], []

    def worker():
        try:
            torch.cuda.set_device(0)
            handles.append(_lib.handle_for(0))
            for k in range(30):
                got_side.append((k % 3, side(k)))
        except Exception as e:  # pragma: no cover
            errors.append(e)
    t = threading.Thread(target=worker)
    t.start()
    got_main = [(k % 3, closure(k)) for k in range(30)]
    t.join()
    assert not errors, errors
    assert handles[0] is not _lib.handle_for(0)            # its own handle: own scratch, own caches
    for k, o in got_main:                                  # deterministic kernels: the interleaving changes nothing
        assert o["loglik"] == ref_main[k]["loglik"] and o["KL"] == ref_main[k]["KL"] and o["flags"] == 0
        assert all(o["grad"][key] == ref_main[k]["grad"][key] for key in o["grad"])
    for k, v in got_side:
        assert v == ref_side[k]


def test_handles_return_their_memory():
    """create -> evaluate (work areas, cached image copies, int8 slices) -> destroy, 12 times: free device memory comes back
    to where it started, and a destroyed handle's caches are not inherited by the next one."""
    from gaussian_processes_b200 import _lib
    from gaussian_processes_b200.native import THETA_KEYS  # noqa: F401  (package import builds the default handle first)
    prob, X, r, xt, Cm, mask, B, m_b, V_b = problem(N=9000, n=20, m=100, cfg=71)
    import ctypes as C
    lib = _lib.lib
    theta = (C.c_double * 6)(*[float(prob.theta[k]) for k in THETA_KEYS])
    q = B.shape[1]

    def one_cycle():
        h = _lib.Handle(0)
        out = (C.c_double * 16)()
        rc = lib.sgp_elbo_grad(h.ptr, C.c_void_p(X.data_ptr()), X.stride(0), C.c_void_p(r.data_ptr()), X.shape[0],
                               C.c_void_p(xt.data_ptr()), xt.stride(0), xt.shape[0], prob.n_px_side, theta, None, None,
                               C.c_void_p(B.data_ptr()), B.stride(0), q, C.c_void_p(m_b.data_ptr()),
                               C.c_void_p(V_b.data_ptr()), V_b.stride(0), float(prob.logA), float(prob.lambda0), None,
                               None, None, None, out, C.c_void_p(torch.cuda.current_stream(D).cuda_stream))
        h.check(rc)
        h.close()
        return out[0]
    first = one_cycle()
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info(0)
    vals = [one_cycle() for _ in range(12)]
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info(0)
    assert all(v == first for v in vals)
    assert free0 - free1 < (8 << 20), (free0, free1)        # nothing accumulates (driver bookkeeping aside)


def test_no_kernel_depends_on_stale_scratch():
    """SGP_POISON=1 fills every work area (and every fresh allocation) with 0xFF bytes -- NaN as a double, -1 as an int8
    digit or an exponent -- before each entry point uses it.  The closure / edge / cells parity tests must pass unchanged in
    a process started that way: no kernel reads scratch it did not write in the same call (compute-sanitizer's initcheck
    is not available on the GPU pool; this is the library's own version of it)."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, SGP_POISON="1")
    files = ["tests/test_gpu_edges.py", "tests/test_gpu_closure.py", "tests/test_gpu_cells.py", "tests/test_gpu_f32.py"]
    res = subprocess.run([sys.executable, "-m", "pytest", *files, "-m", "gpu", "-x", "-q", "-p", "no:cacheprovider"],
                         cwd=root, env=env, capture_output=True, text=True, timeout=900)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-2000:]


def test_evaluations_are_bitwise_reproducible():
    """Every reduction in the library runs in a fixed order (split-K partials summed by slice_reduce, per-block partials by
    tickets, column scales by atomicMax -- order-independent): the same call gives the same bits, call after call, with cold
    and warm caches, from a fresh handle in another thread, and on the int8 and the DMMA engine each."""
    from gaussian_processes_b200 import native
    n, m, N = 40, 200, 20000
    prob = synth.make_problem(N, n, m, cfg=72, theta=synth.THETA_DENSE)
    X, r, xt = prob.X.to(D), prob.r.to(D), prob.xtilde.to(D)
    Cm, mask = native.localker(prob.theta, n, D)
    Kt = native.acosker(prob.theta["sigma_0"], xt, xt, Cm, True)
    ev, evec = torch.linalg.eigh(Kt, UPLO="L")
    keep = ev > max(float(ev.max()) * 1e-4, 1e-4)
    B = evec[:, keep].contiguous()
    m_b, V_b = synth.make_variational_state(torch.diag(ev[keep]).cpu(), 72)
    m_b, V_b = m_b.to(D), V_b.to(D)
    theta = dict(prob.theta, eps_0x=0.07)

    def ev_():
        o = native.elbo_grad(X, r, xt, n, theta, B, m_b, V_b, prob.logA, prob.lambda0, want_moments=True)
        return (o["loglik"], o["KL"], tuple(o["grad"][k] for k in native.THETA_KEYS), o["flags"]), o["lam_var"].clone()
    for i8 in (-1, 0):
        native.set_option(0, "i8_slices", i8)
        native.set_slab_rows(0, 8192)                          # three slabs, a ragged last one
        try:
            native.invalidate_cache(0)
            first, lv0 = ev_()                                 # cold caches
            assert first[3] == 0
            for rep in range(25):
                if rep == 10:
                    native.invalidate_cache(0)
                got, lv = ev_()
                assert got == first and torch.equal(lv, lv0), (i8, rep)
            res = []

            def other_thread():
                torch.cuda.set_device(0)
                native.set_option(0, "i8_slices", i8)          # this thread's own handle
                native.set_slab_rows(0, 8192)
                res.append(ev_())
            t = threading.Thread(target=other_thread)
            t.start(); t.join()
            assert res and res[0][0] == first and torch.equal(res[0][1], lv0)
        finally:
            native.set_option(0, "i8_slices", -1)
            native.set_slab_rows(0, 0)
            native.invalidate_cache(0)
