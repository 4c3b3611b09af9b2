"""Scenario-batch sharding across ranks: every scenario is an independent NLP, so rank r owns a
contiguous slice of the batch and there is no collective inside the solve; one all_gather of
(action, exitflag, iters) per call returns the whole batch's result to every rank."""
from __future__ import annotations

from typing import Tuple

import os

import torch
import torch.distributed as dist


def init_process_group(backend: str = "nccl", **kw):
    """torch.distributed.init_process_group with NCCL held to ONE channel unless the caller's environment says otherwise.
    The only collective on this path is the per-step all_gather of 24-40 bytes per scenario; its kernel is resident from
    the moment the rank's own solve is done until every peer's is, and each NCCL channel is a 512-thread CTA whose
    registers displace the solve kernel's one-warp blocks (7 per SM) from the SM it sits on.  Measured on 4 B200s,
    boxer 4096 per GPU (profiles/r1n_nccl_ab.txt): default channels 6.26 ms per step, one channel 5.62 (one GPU without
    any gather: 5.54)."""
    if backend == "nccl":
        os.environ.setdefault("NCCL_MAX_NCHANNELS", "1")
    return dist.init_process_group(backend, **kw)


def shard_bounds(B: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous split of B scenarios into `world` shards of ceil(B/world) (the last may be short or empty)."""
    per = (B + world - 1) // world
    lo = min(B, rank * per)
    return lo, min(B, lo + per)


class PendingGather:
    """An all_gather in flight on NCCL's own stream; wait() makes the current stream wait for it and unpacks."""

    def __init__(self, work, out, B, nu):
        self._work, self._out, self._B, self._nu = work, out, B, nu

    def wait(self):
        if self._work is not None:
            self._work.wait()
        out, nu = self._out[: self._B], self._nu
        return out[:, :nu].contiguous(), out[:, nu].to(torch.int32), out[:, nu + 1].to(torch.int32)


def gather_results(action: torch.Tensor, exitflag: torch.Tensor, iters: torch.Tensor, B: int, async_op: bool = False):
    """all_gather of the per-rank results (padded to the common shard size); returns full-batch tensors, or with
    async_op a PendingGather: the stream that ran the solve is then not held up until every rank has arrived (a
    throughput loop waits for its gathers once, at the end; a control loop that needs the actions waits at once)."""
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
    work = dist.all_gather_into_tensor(out, buf, async_op=async_op)
    pending = PendingGather(work if async_op else None, out, B, nu)
    return pending if async_op else pending.wait()
