"""Worker of tests/test_sharded_gloo.py: one rank of an image-sharded run on CPU (gloo), with the CUDA library
replaced by the oracle-backed stand-in (tests/fake_native.py).  Usage: shard_worker.py <mode> <out.json>"""
import contextlib
import io
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))

import torch
import torch.distributed as dist

import fake_native
from gaussian_processes_b200 import sharded, synth
from helpers import case_inputs, load


def main():
    mode, out_path = sys.argv[1], sys.argv[2]
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    with contextlib.redirect_stdout(io.StringIO()):
        from gaussian_processes_b200 import utils
    utils.native = fake_native
    sg = sharded.enable()
    res = {"rank": rank, "world": world}
    if mode == "closure":
        fx = load("closure_c1_local")
        prob, st = case_inputs(fx)
        rows = sharded.shard_rows(prob.N, rank, world)
        o = sharded.elbo_grad(prob.X[rows], prob.r[rows], prob.xtilde, prob.n_px_side, prob.theta, st["B"], st["m_b"],
                              st["V_b"], st["logA"], st["lambda0"], sg=sg, native_mod=fake_native)
        res.update(loglik=o["loglik"], KL=o["KL"], grad=[o["grad"][k] for k in fake_native.THETA_KEYS],
                   reduces=sg.calls)
    elif mode == "fit":
        fx = load("fit_small_n10")
        keys = [str(k) for k in fx["theta_keys"]]
        prob = synth.make_problem(int(fx["N"]), int(fx["n"]), int(fx["m"]), cfg=int(fx["cfg"]),
                                  theta=dict(zip(keys, fx["theta0"].tolist())))
        rows = sharded.shard_rows(prob.N, rank, world)
        args = dict(
            fit_parameters=dict(ntilde=int(fx["m"]), maxiter=int(fx["maxiter"]), nEstep=int(fx["nEstep"]),
                                nMstep=int(fx["nMstep"]), nFparamstep=int(fx["nFparamstep"]), kernfun="acosker",
                                cellid=0, n_px_side=int(fx["n"]), display_hyper=False),
            xtilde=prob.xtilde,
            hyperparams_tuple=(synth.theta_tensors(prob.theta, requires_grad=True), dict(synth.THETA_LIMS_LO),
                               dict(synth.THETA_LIMS_HI)),
            f_params={"logA": torch.tensor(prob.logA), "lambda0": torch.tensor(prob.lambda0)})
        with contextlib.redirect_stdout(io.StringIO()):
            fit, err = utils.varGP(prob.X[rows], prob.r[rows], **args)
        vt = fit["values_track"]
        res.update(is_error=bool(err["is_error"]), error=repr(err.get("error")),
                   logmarginal=vt["loss_track"]["logmarginal"].tolist(),
                   theta_track=[vt["theta_track"][k].tolist() for k in keys],
                   logA_track=vt["f_par_track"]["logA"].tolist(), lambda0_track=vt["f_par_track"]["lambda0"].tolist(),
                   reduces=sg.calls)
    json.dump(res, open(f"{out_path}.{rank}", "w"))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
