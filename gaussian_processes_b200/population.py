"""Population fit: independent cells that share the stimuli, sharded across GPUs (BASELINE.json configs[3]; SURVEY.md
section 8e last row; north_star item 4 "sharding of ... independent cells across the 8 GPUs").

The reference fits one cell per ``varGP`` call (utils.py:1568; ``cellid`` in ``fit_parameters``, utils.py:1636-1643).
Cells are independent given the images: each has its own theta (hence its own receptive-field mask, C, K_tilde, B) and
its own (m_b, V_b, f_params); the only shared objects are the images X and the inducing images.  So the cells are
split over the ranks in contiguous blocks -- no data-path collective -- every rank keeps ONE resident copy of X (and one
set of int8 slices of it), and evaluates its cells back to back through ``sgp_elbo_grad_cells``; one gather of the
16-double results at the end.
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch

from . import native


def cells_for_rank(ncells: int, rank: int, world: int) -> range:
    """Contiguous block of cell indices owned by `rank` (the first ncells % world ranks own one more)."""
    base, extra = divmod(ncells, world)
    lo = rank * base + min(rank, extra)
    return range(lo, lo + base + (1 if rank < extra else 0))


def evaluate_cells(X: torch.Tensor, xtilde: torch.Tensor, n_px: int, cells: Sequence[dict]) -> List[dict]:
    """ELBO + gradient of every cell in `cells` (dicts with theta, r, B, m_b, V_b, logA, lambda0[, logdetV]) on this
    rank's GPU: one C call, the cells evaluated back to back on the current stream."""
    if not cells:
        return []
    ldv = [c.get("logdetV") for c in cells]
    return native.elbo_grad_cells(X, [c["r"] for c in cells], xtilde, n_px, [c["theta"] for c in cells],
                                  [c["B"] for c in cells], [c["m_b"] for c in cells], [c["V_b"] for c in cells],
                                  [c["logA"] for c in cells], [c["lambda0"] for c in cells],
                                  logdetV=ldv if all(v is not None for v in ldv) else None)


def gather_results(local: List[dict], group=None) -> Optional[List[dict]]:
    """All cells' results on every rank, in cell order (ranks hold contiguous blocks).  No-op without a process group."""
    import torch.distributed as dist
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return list(local)
    parts = [None] * dist.get_world_size(group)
    dist.all_gather_object(parts, local, group=group)
    return [r for part in parts for r in part]
