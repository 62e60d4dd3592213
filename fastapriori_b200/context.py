"""MiningContext — what the reference's SparkContext is to its operators (Main.scala:17-22):
the handle every operator call receives.  Here it owns one Engine (one GPU, one process) and,
when torch.distributed is initialised with more than one rank, the row-sharding plumbing:
every rank holds a contiguous shard of the transactions, support counts are summed across
ranks on the device (NCCL all-reduce), and everything derived from global counts is
recomputed identically on every rank.
"""
from __future__ import annotations

import os

import numpy as np


class _DevInt32:
    """CUDA-array-interface view of `count` int32 at a raw device pointer (no copy)."""

    def __init__(self, ptr: int, count: int):
        self.__cuda_array_interface__ = {
            "shape": (count,), "typestr": "<i4", "data": (ptr, False), "version": 3, "strides": None,
        }


class MiningContext:
    def __init__(self, device: int | None = None, engine=None, group=None, verbose: bool = False, distributed: bool = True,
                 exchange: str | None = None):
        import torch

        self.verbose = verbose
        self.group = group
        self.stream = None
        self._dist = None
        try:
            import torch.distributed as dist
            if distributed and dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
                self._dist = dist
        except Exception:
            self._dist = None
        self.rank = self._dist.get_rank(group) if self._dist else 0
        self.world_size = self._dist.get_world_size(group) if self._dist else 1
        if engine is not None:
            self.engine = engine  # tests inject a stand-in to exercise the sharding logic without a GPU
            self.device = getattr(engine, "device", 0)
        else:
            if not torch.cuda.is_available():
                raise RuntimeError("fastapriori_b200 needs a CUDA device: the mining path has no CPU fallback")
            self.device = torch.cuda.current_device() if device is None else int(device)
            torch.cuda.set_device(self.device)
            from .engine import Engine
            # One explicit torch stream carries every kernel of this context.  (torch's default stream has handle 0,
            # which the C-ABI reads as "create your own": collectives issued by torch would then be unordered with
            # respect to the counting kernels.)  Events and collectives are issued with this stream current.
            self.stream = torch.cuda.Stream(device=self.device)
            self.engine = Engine(self.device, self.stream.cuda_stream)
            if self._dist:
                # "native": the library all-reduces over its own NCCL communicator (no Python per level); the id is
                # shipped through the torch group.  "callback": torch.distributed does the all-reduce via the hook.
                mode = exchange or os.environ.get("FA_EXCHANGE", "native")
                if mode == "native" and self._dist.get_backend(group) == "nccl":
                    box = [Engine.comm_unique_id() if self.rank == 0 else None]
                    src = self._dist.get_global_rank(group, 0) if group is not None else 0
                    self._dist.broadcast_object_list(box, src=src, group=group)
                    self.engine.comm_init(box[0], self.rank, self.world_size)
                elif mode in ("native", "callback"):
                    self.engine.set_exchange(self._allreduce_device)
                else:
                    raise ValueError(f"exchange must be 'native' or 'callback', not {mode!r}")
                self.exchange = mode

    # -- collectives -----------------------------------------------------------------------
    def _allreduce_device(self, ptr: int, count: int, stream: int):
        import torch
        # ProcessGroupNCCL orders the collective after the work already queued on the CURRENT stream and makes that
        # stream wait for the result, so the ctx stream has to be current here
        with torch.cuda.stream(self.stream):
            t = torch.as_tensor(_DevInt32(ptr, count), device=f"cuda:{self.device}")
            self._dist.all_reduce(t, op=self._dist.ReduceOp.SUM, group=self.group)

    def allreduce_host(self, arr: np.ndarray) -> np.ndarray:
        """Sum of a small host array over ranks (row counts, sizes)."""
        if not self._dist:
            return arr
        import torch
        t = torch.from_numpy(np.ascontiguousarray(arr).copy())
        if self._dist.get_backend(self.group) == "nccl":
            t = t.cuda(self.device)
        self._dist.all_reduce(t, op=self._dist.ReduceOp.SUM, group=self.group)
        return t.cpu().numpy()

    def allgather_object(self, obj):
        if not self._dist:
            return [obj]
        out = [None] * self.world_size
        self._dist.all_gather_object(out, obj, group=self.group)
        return out

    def shard_bounds(self, n: int):
        """Contiguous row range of this rank."""
        per = (n + self.world_size - 1) // self.world_size
        lo = min(n, self.rank * per)
        return lo, min(n, lo + per)

    def log(self, msg: str):
        if self.verbose and self.rank == 0:
            print(msg, flush=True)

    def close(self):
        if hasattr(self.engine, "close"):
            self.engine.close()
