"""Scenario-batch sharding across ranks: every scenario is an independent NLP, so rank r owns a
contiguous slice of the batch and there is no collective inside the solve; one all_gather of
(action, exitflag, iters) per call returns the whole batch's result to every rank."""
from __future__ import annotations

from typing import Tuple

import torch
import torch.distributed as dist


def shard_bounds(B: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous split of B scenarios into `world` shards of ceil(B/world) (the last may be short or empty)."""
    per = (B + world - 1) // world
    lo = min(B, rank * per)
    return lo, min(B, lo + per)


def gather_results(action: torch.Tensor, exitflag: torch.Tensor, iters: torch.Tensor, B: int):
    """all_gather of the per-rank results (padded to the common shard size); returns full-batch tensors."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return action, exitflag, iters
    world = dist.get_world_size()
    per = (B + world - 1) // world
    nu = action.shape[1]
    # one packed float64 buffer per rank: [per, nu + 2]
    buf = torch.zeros(per, nu + 2, dtype=torch.float64, device=action.device)
    n = action.shape[0]
    buf[:n, :nu] = action
    buf[:n, nu] = exitflag.to(torch.float64)
    buf[:n, nu + 1] = iters.to(torch.float64)
    out = torch.empty(world * per, nu + 2, dtype=torch.float64, device=action.device)
    dist.all_gather_into_tensor(out, buf)
    out = out[:B]
    return out[:, :nu].contiguous(), out[:, nu].to(torch.int32), out[:, nu + 1].to(torch.int32)
