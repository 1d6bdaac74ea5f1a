"""Image sharding of the Spatial_GP hot path across GPUs (north_star item 4).

One process per GPU (``torchrun``); rank k holds a contiguous block of the training images ``x`` / spike counts
``r``; the small state (theta, xtilde, B, m_b, V_b, f_params) is replicated and recomputed bit-identically on
every rank.  Every place the path contracts over images is a plain sum, so the exchange step is ONE all-reduce
of a few doubles per evaluation:

* M-step closure (``elbo_grad``): ``[loglik, 6 gradients, K_tilde_b-not-PD flag]``; the N-independent KL term and
  its gradient are computed on every rank and counted once (``kl_weight`` = 1 on rank 0, 0 elsewhere);
* E-step (``utils.Estep``): ``g`` (q) and ``G`` (q x q);
* logA closure / lambda0 / loss tracking: 4-6 scalar sums.

``enable()`` installs the reducer into ``gaussian_processes_b200.utils`` so that ``utils.varGP`` runs SPMD: all
ranks execute the same host loop on their shard and see identical reduced numbers, hence identical LBFGS
trajectories.  Backend: NCCL over NVLink on GPUs; ``gloo`` is used by the CPU tests of this host logic.
"""
from __future__ import annotations

from typing import Optional, Sequence

import torch
import torch.distributed as dist


class ShardGroup:
    """Sum-reducer over a torch.distributed process group."""

    def __init__(self, group=None, device: Optional[torch.device] = None):
        if not dist.is_initialized():
            raise RuntimeError("sharded.enable(): torch.distributed is not initialised")
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        backend = dist.get_backend(group)
        if device is None:
            device = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
        self.device = torch.device(device)
        self.calls = 0
        self.peer_exchange = False

    def enable_peer_exchange(self, native_mod=None) -> bool:
        """Fuse the exchange step of the M-step closure into its result-assembly kernel: every rank's exchange buffer is
        mapped into every other rank (CUDA IPC over NVLink / NVSwitch peer access; the 64-byte handles travel through
        this process group once), after which `native.elbo_grad` returns numbers already summed over the image shards --
        no NCCL launch and a single host read-back per evaluation.  NCCL stays for the E-step's q x q all-reduce."""
        if self.world == 1 or self.device.type != "cuda":
            return False
        if native_mod is None:
            from . import native as native_mod
        dev = self.device.index if self.device.index is not None else torch.cuda.current_device()
        mine = native_mod.comm_export(dev, self.world)
        allh = [None] * self.world
        dist.all_gather_object(allh, mine, group=self.group)
        native_mod.comm_import(dev, self.rank, self.world, b"".join(allh))
        dist.barrier(group=self.group)
        self.peer_exchange = True
        return True

    def disable_peer_exchange(self, native_mod=None) -> None:
        if not self.peer_exchange:
            return
        if native_mod is None:
            from . import native as native_mod
        dist.barrier(group=self.group)
        native_mod.comm_close(self.device.index if self.device.index is not None else torch.cuda.current_device())
        self.peer_exchange = False

    def sum_floats(self, vals: Sequence[float]):
        t = torch.tensor([float(v) for v in vals], dtype=torch.float64, device=self.device)
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        self.calls += 1
        return t.tolist()

    def sum_tensor_(self, t: torch.Tensor) -> torch.Tensor:
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        self.calls += 1
        return t

    def reduce_result_(self, out_dev: torch.Tensor):
        """The one exchange step of an image-sharded evaluation, on the DEVICE and in stream order: the 16 result doubles
        of sgp_elbo_grad_dev stay where the kernel wrote them, entries [0..8] (loglik, KL, 6 gradients, flags) are
        all-reduced in place over NCCL on the launching stream, and the host reads them back ONCE.  KL is identical on
        every rank and must count once: ranks > 0 zero their copy first (the KL part of their gradient is already
        weighted by kl_weight = 0)."""
        if self.rank != 0:
            out_dev[1:2].zero_()
        # status bits do not survive a sum; what the callers need is "K_tilde_b was not positive definite on some rank"
        # (bits 1 | 2): reduce that as a count
        fl = out_dev[8:9]
        fl.copy_(torch.remainder(fl, 4.0).gt(0.0))
        dist.all_reduce(out_dev[:9], op=dist.ReduceOp.SUM, group=self.group)
        self.calls += 1
        return out_dev.cpu().tolist()          # the only host synchronisation of the evaluation


def enable(group=None, device=None) -> ShardGroup:
    """Make ``utils`` treat its ``x`` / ``r`` arguments as this rank's block of the images."""
    from . import utils
    sg = ShardGroup(group, device)
    utils._SHARD = sg
    return sg


def disable() -> None:
    from . import utils
    utils._SHARD = None


def shard_rows(N: int, rank: int, world: int) -> slice:
    """Contiguous block of rank `rank` out of N images (the first N % world ranks hold one image more)."""
    base, extra = divmod(N, world)
    lo = rank * base + min(rank, extra)
    return slice(lo, lo + base + (1 if rank < extra else 0))


def elbo_grad(X_local, r_local, xtilde, n_px, theta, B, m_b, V_b, logA, lambda0, sg: Optional[ShardGroup] = None,
              native_mod=None, logdetV=None):
    """One image-sharded ELBO+gradient evaluation: the fused call on the local block, then ONE all-reduce.
    Returns dict(loglik, KL, grad{key}, flags) identical on every rank."""
    if native_mod is None:
        from . import native as native_mod
    keys = native_mod.THETA_KEYS
    rank = sg.rank if sg is not None else 0
    if sg is not None and sg.world > 1 and getattr(sg, "peer_exchange", False):
        # the exchange happened inside the evaluation's last kernel: the 16 doubles read back are already reduced
        out = native_mod.elbo_grad(X_local, r_local, xtilde, n_px, theta, B, m_b, V_b, logA, lambda0, logdetV=logdetV,
                                   kl_weight=1.0 if rank == 0 else 0.0, same=False)
        return {"loglik": out["loglik"], "KL": out["KL"], "grad": out["grad"], "flags": out["flags"],
                "any_flag": out["flags"] != 0, "p": out["p"]}
    if sg is not None and sg.world > 1 and X_local.is_cuda and hasattr(native_mod, "elbo_grad_dev"):
        dev = native_mod.elbo_grad_dev(X_local, r_local, xtilde, n_px, theta, B, m_b, V_b, logA, lambda0,
                                       logdetV=logdetV, kl_weight=1.0 if rank == 0 else 0.0)
        v = sg.reduce_result_(dev)
        return {"loglik": v[0], "KL": v[1], "grad": dict(zip(keys, v[2:8])), "flags": int(v[8] > 0),
                "any_flag": v[8] > 0, "p": int(v[9])}
    out = native_mod.elbo_grad(X_local, r_local, xtilde, n_px, theta, B, m_b, V_b, logA, lambda0, logdetV=logdetV,
                               kl_weight=1.0 if rank == 0 else 0.0)
    vec = [out["loglik"]] + [out["grad"][k] for k in keys] + [float(out["flags"] != 0)]
    if sg is not None and sg.world > 1:
        vec = sg.sum_floats(vec)
    return {"loglik": vec[0], "KL": out["KL"], "grad": dict(zip(keys, vec[1:7])), "flags": out["flags"],
            "any_flag": vec[7] > 0, "p": out.get("p")}
