"""Scenario-batch sharding across ranks: every scenario is an independent NLP, so rank r owns a
contiguous slice of the batch and there is no collective inside the solve.  What a control loop wants back on every rank
is the whole batch's (action, exitflag, iters): `PeerExchange` delivers it with the library's own kernels over peer memory
(NVLink / NVSwitch, csrc/rmpc_exchange.cuh); `gather_results` is the same through one NCCL / gloo all_gather per step and
is what runs where CUDA IPC is not available (CPU tests, ranks on different nodes)."""
from __future__ import annotations

from typing import List, Optional, Tuple

import ctypes

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


class PendingCollect:
    """A collect in flight on the exchange's own stream; wait() makes the current stream wait for it and returns the
    whole batch's (action[B, nu], exitflag[B], iters[B]).  Keep the object until wait(): it owns the output tensors."""

    def __init__(self, xchg, out, B):
        self._x, self._out, self._B = xchg, out, B

    def wait(self):
        self._x.join()
        a, e, i = self._out
        return a[: self._B], e[: self._B], i[: self._B]


class PeerExchange:
    """Every rank's results of a step on every rank, through the solver library's exchange kernels (include/rmpc.h
    `rmpc_exchange_*`): each rank allocates one region, the ranks swap its CUDA IPC handle once (all_gather_object), and from
    then on a step costs one small store kernel behind the solve and one waiting warp + unpack on a side stream; no NCCL
    kernel sits on the SMs.  `regions` lets one process play several ranks on one GPU (tests): a list of this many device
    pointers, shared by the exchanges of all its ranks."""

    def __init__(self, solver, rows_per_rank: int, depth: int = 8, rank: Optional[int] = None, world: Optional[int] = None,
                 regions: Optional[List[int]] = None, group=None):
        self.lib, self.device, self.nu = solver.lib, solver.device, solver.nu
        self._err = solver._err
        self._group = group
        if regions is None:
            world = dist.get_world_size(group)
            rank = dist.get_rank(group)
        self.rank, self.world, self.rows, self.depth = rank, world, rows_per_rank, depth
        self._opened: List[int] = []
        self._own = None
        self._h = None
        if regions is None:
            nbytes = self.lib.rmpc_exchange_region_bytes(world, rows_per_rank, depth)
            if nbytes == 0:
                raise ValueError("PeerExchange: bad world / rows_per_rank / depth")
            ptr, handle = ctypes.c_void_p(), ctypes.create_string_buffer(64)
            if self.lib.rmpc_ipc_alloc(self.device, nbytes, ctypes.byref(ptr), handle) != 0:
                raise RuntimeError(self._err())
            self._own = ptr.value
            handles = [None] * world
            dist.all_gather_object(handles, (rank, handle.raw), group=group)
            regions = [0] * world
            failure = ""
            for r, raw in handles:
                if r == rank:
                    regions[r] = self._own
                    continue
                p = ctypes.c_void_p()
                if self.lib.rmpc_ipc_open(self.device, raw, ctypes.byref(p)) != 0:
                    failure = "cannot map the region of rank %d: %s" % (r, self._err())
                    break
                regions[r] = p.value
                self._opened.append(p.value)
            # all ranks or none: a rank that cannot map a peer (no peer access, another node, IPC forbidden) takes everyone back
            ok = torch.tensor([0 if failure else 1], dtype=torch.int32, device=torch.device("cuda", self.device))
            dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
            if int(ok) == 0:
                for p in self._opened:
                    self.lib.rmpc_ipc_close(self.device, p)
                self._opened = []
                dist.barrier(group=group)
                self.lib.rmpc_ipc_free(self.device, self._own)
                self._own = None
                raise RuntimeError("PeerExchange: " + (failure or "a peer could not map this rank's region"))
        arr = (ctypes.c_void_p * world)(*regions)
        self._h = self.lib.rmpc_exchange_create(self.device, world, rank, rows_per_rank, depth, arr)
        if not self._h:
            raise RuntimeError(self._err())
        if self._own is not None:
            dist.barrier(group=group)   # every region is mapped everywhere before the first store

    @staticmethod
    def alloc_regions(solver, world: int, rows_per_rank: int, depth: int = 8) -> List[int]:
        """`world` regions on the solver's device for ranks that live in one process (tests)."""
        nbytes = solver.lib.rmpc_exchange_region_bytes(world, rows_per_rank, depth)
        out = []
        for _ in range(world):
            p = ctypes.c_void_p()
            if solver.lib.rmpc_ipc_alloc(solver.device, nbytes, ctypes.byref(p), None) != 0:
                raise RuntimeError(solver._err())
            out.append(p.value)
        return out

    def publish(self, action: torch.Tensor, exitflag: torch.Tensor, iters: torch.Tensor) -> int:
        """Enqueued behind what the current stream holds now, i.e. behind the solve that writes the three tensors (keep them
        until the step is collected).  Returns the step number."""
        assert action.is_cuda and action.dtype == torch.float64 and action.is_contiguous() and action.shape[1] == self.nu
        assert exitflag.dtype == torch.int32 and iters.dtype == torch.int32
        stream = torch.cuda.current_stream(action.device).cuda_stream
        seq = self.lib.rmpc_exchange_publish(self._h, action.shape[0], action.data_ptr(), exitflag.data_ptr(), iters.data_ptr(), stream)
        if seq < 0:
            raise RuntimeError(self._err())
        return seq

    def collect(self, seq: int, B: int) -> PendingCollect:
        """The world's results of step `seq` (steps are collected in order); the wait for the slowest rank and the unpack run
        on the exchange's own stream (the unpack behind what the current stream holds now), the returned object's wait() joins
        the current stream to them."""
        dev = torch.device("cuda", self.device)
        n = self.world * self.rows
        out = (torch.empty(n, self.nu, dtype=torch.float64, device=dev), torch.empty(n, dtype=torch.int32, device=dev),
               torch.empty(n, dtype=torch.int32, device=dev))
        stream = torch.cuda.current_stream(dev).cuda_stream
        if self.lib.rmpc_exchange_collect(self._h, seq, out[0].data_ptr(), out[1].data_ptr(), out[2].data_ptr(), stream) != 0:
            raise RuntimeError(self._err())
        return PendingCollect(self, out, B)

    def join(self):
        """The current stream waits for every collect enqueued so far."""
        if self.lib.rmpc_exchange_join(self._h, torch.cuda.current_stream(torch.device("cuda", self.device)).cuda_stream) != 0:
            raise RuntimeError(self._err())

    def status(self) -> int:
        return self.lib.rmpc_exchange_status(self._h)

    def close(self):
        if self._h:
            self.lib.rmpc_exchange_destroy(self._h)
            self._h = None
        for p in self._opened:
            self.lib.rmpc_ipc_close(self.device, p)
        self._opened = []
        if self._own is not None:
            if dist.is_initialized():
                dist.barrier(group=self._group)   # nobody still stores into a region that is about to be freed
            self.lib.rmpc_ipc_free(self.device, self._own)
            self._own = None
