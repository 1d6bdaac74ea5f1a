"""Launched by torchrun on >= 2 GPUs (tests/test_gpu_multigpu.py, or by hand under `gpurun --gpus N`):
image-sharded ELBO+gradient evaluation with the three exchange variants -- peer-memory exchange fused into the
result-assembly kernel (sgp_comm_*), device-side NCCL all-reduce (sgp_elbo_grad_dev), host-side reduction -- against ONE
evaluation of all the images on one GPU.  Prints one JSON line on rank 0; exit code 0 = all agree."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from gaussian_processes_b200 import native, sharded, synth  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device(f"cuda:{local}")
    dist.init_process_group("nccl", device_id=dev)
    N, n, m = int(os.environ.get("MG_N", 24000)), int(os.environ.get("MG_NPX", 40)), int(os.environ.get("MG_M", 200))
    th0 = synth.theta_localised(n) if os.environ.get("MG_LOCAL") else synth.THETA_DENSE
    prob = synth.make_problem(N, n, m, cfg=77, theta=th0)
    Xall, rall, xt = prob.X.to(dev), prob.r.to(dev), prob.xtilde.to(dev)
    Cm, mask = native.localker(prob.theta, n, dev)
    um = xt if bool(mask.all()) else xt[:, mask].contiguous()
    Kt = native.acosker(prob.theta["sigma_0"], um, um, Cm, True)
    ev, evec = torch.linalg.eigh(Kt, UPLO="L")
    keep = ev > max(float(ev.max()) * 1e-4, 1e-4)
    B = evec[:, keep].contiguous()
    m_b, V_b = synth.make_variational_state(torch.diag(ev[keep]).cpu(), 77)
    m_b, V_b = m_b.to(dev), V_b.to(dev)
    for t in (B, m_b, V_b):
        dist.broadcast(t, 0)
    theta = dict(prob.theta)
    theta["eps_0x"] += 0.01
    sl = sharded.shard_rows(N, rank, world)
    X, r = Xall[sl].contiguous(), rall[sl].contiguous()
    sg = sharded.ShardGroup(device=dev)
    whole = native.elbo_grad(Xall, rall, xt, n, theta, B, m_b, V_b, prob.logA, prob.lambda0)   # every rank, no exchange yet
    native.invalidate_cache(local)
    res, ms = {}, {}

    def timed(fn, name, reps=20):
        for _ in range(3):
            out = fn()
        dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(reps):
            out = fn()
        dist.barrier()
        torch.cuda.synchronize()
        ms[name] = (time.perf_counter() - t0) / reps * 1e3
        res[name] = out

    def ev_(**kw):
        return sharded.elbo_grad(X, r, xt, n, theta, B, m_b, V_b, prob.logA, prob.lambda0, sg=sg, **kw)
    timed(ev_, "nccl_device")

    class HostOnly:                        # round-1 protocol: read back, re-pack, all-reduce, read back again
        THETA_KEYS = native.THETA_KEYS
        elbo_grad = staticmethod(native.elbo_grad)
    timed(lambda: ev_(native_mod=HostOnly), "nccl_host")
    assert sg.enable_peer_exchange()
    timed(ev_, "peer_memory")
    sg.disable_peer_exchange()
    after = native.elbo_grad(X, r, xt, n, theta, B, m_b, V_b, prob.logA, prob.lambda0,
                             kl_weight=1.0 if rank == 0 else 0.0)      # local again once the mode is closed
    ok = True
    rep = {"world": world, "N": N, "ms_per_eval": ms, "rel_err_vs_single_gpu": {}}
    gmax = max(abs(v) for v in whole["grad"].values())
    for name, o in res.items():
        e_ll = abs(o["loglik"] - whole["loglik"]) / abs(whole["loglik"])
        e_kl = abs(o["KL"] - whole["KL"]) / abs(whole["KL"])
        e_g = max(abs(o["grad"][k] - whole["grad"][k]) for k in whole["grad"]) / gmax
        rep["rel_err_vs_single_gpu"][name] = {"loglik": e_ll, "KL": e_kl, "grad": e_g, "flags": o["flags"]}
        ok = ok and e_ll < 1e-12 and e_kl < 1e-12 and e_g < 1e-10 and o["flags"] == 0
    # every rank must hold identical reduced numbers
    v = torch.tensor([res["peer_memory"]["loglik"]] + [res["peer_memory"]["grad"][k] for k in native.THETA_KEYS],
                     dtype=torch.float64, device=dev)
    lo, hi = v.clone(), v.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    rep["peer_memory_identical_on_all_ranks"] = bool(torch.equal(lo, hi))
    rep["local_after_close_differs"] = after["loglik"] != res["peer_memory"]["loglik"]
    ok = ok and rep["peer_memory_identical_on_all_ranks"] and rep["local_after_close_differs"]
    # ---- the whole EM fit, SPMD over the image shards (utils.varGP with sharded.enable): tracks must equal the fit of
    #      ALL the images on one GPU.  M-step closures reduce inside the evaluation's last kernel (peer memory), the
    #      E-step's g / G and the logA-closure sums go through NCCL.
    if os.environ.get("MG_FIT", "1") != "0":
        import contextlib
        import io
        with contextlib.redirect_stdout(io.StringIO()):
            from gaussian_processes_b200 import utils as U

            def fit_kw():
                return dict(fit_parameters=dict(ntilde=m, maxiter=3, nEstep=2, nMstep=3, nFparamstep=3, kernfun="acosker",
                                                cellid=0, n_px_side=n, display_hyper=False),
                            xtilde=xt, hyperparams_tuple=(synth.theta_tensors(prob.theta, device=dev, requires_grad=True),
                                                          dict(synth.THETA_LIMS_LO), dict(synth.THETA_LIMS_HI)),
                            f_params={"logA": torch.tensor(prob.logA, device=dev),
                                      "lambda0": torch.tensor(prob.lambda0, device=dev)})
            single, e1 = U.varGP(Xall, rall, **fit_kw())                  # every rank: all images, no sharding
            sgf = sharded.enable(device=dev)
            sgf.enable_peer_exchange()
            shard, e2 = U.varGP(X, r, **fit_kw())
            sgf.disable_peer_exchange()
            sharded.disable()
        fit_ok = not e1["is_error"] and not e2["is_error"]
        errs = {}
        for key in ("logmarginal", "loglikelihood", "KL"):
            a_, b_ = shard["values_track"]["loss_track"][key], single["values_track"]["loss_track"][key]
            errs[key] = float((a_ - b_).abs().max() / b_.abs().max())
        ta = torch.stack([shard["values_track"]["theta_track"][k] for k in native.THETA_KEYS])
        tb = torch.stack([single["values_track"]["theta_track"][k] for k in native.THETA_KEYS])
        errs["theta"] = float((ta - tb).abs().max() / tb.abs().max())
        rep["sharded_fit_vs_single_gpu"] = errs
        ok = ok and fit_ok and max(errs.values()) < 1e-9
    okt = torch.tensor([1.0 if ok else 0.0], device=dev)
    dist.all_reduce(okt, op=dist.ReduceOp.MIN)
    if rank == 0:
        rep["ok"] = bool(okt.item() > 0)
        print(json.dumps(rep))
    dist.destroy_process_group()
    sys.exit(0 if okt.item() > 0 else 1)


if __name__ == "__main__":
    main()
