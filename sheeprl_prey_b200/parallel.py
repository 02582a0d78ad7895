"""Data-parallel exchange steps of the training step over torch.distributed (NCCL on GPUs, gloo in CPU tests).
Reference: the DDP gradient all-reduce implicit in fabric.backward (sheeprl/algos/dreamer_v3/agent.py:1054-1063,
dreamer_v3.py:179,:318,:339) and fabric.all_gather inside Moments (algos/dreamer_v3/utils.py:35).  The engine reduces one
flat gradient arena per optimizer group instead of the reference's nine DDP reducers."""
from __future__ import annotations

import torch
import torch.distributed as dist


class TorchDist:
    def __init__(self, world_size: int):
        if not dist.is_initialized():
            raise RuntimeError("torch.distributed is not initialised")
        self.world_size = world_size
        # measurement switches (tools only, results are wrong with them): DV3_DP_SKIP=reduce,gather drops the named collectives so
        # that their cost on the critical path can be read off the step time
        import os
        self._skip = set(filter(None, os.environ.get("DV3_DP_SKIP", "").split(",")))

    def all_reduce_mean(self, t: torch.Tensor) -> None:
        if "reduce" in self._skip:
            return
        if dist.get_backend() == "nccl":
            dist.all_reduce(t, op=dist.ReduceOp.AVG)
        else:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
            t.div_(self.world_size)

    def start_reduce_mean(self, t: torch.Tensor):
        """Asynchronous all_reduce_mean: the collective runs on the process group's stream while the caller keeps enqueueing
        independent work; finish_reduce(handle) orders the caller's stream after it (inside a CUDA graph capture this becomes a
        fork / join of the graph)."""
        if "reduce" in self._skip:
            return (None, None)
        if dist.get_backend() == "nccl":
            return (dist.all_reduce(t, op=dist.ReduceOp.AVG, async_op=True), None)
        return (dist.all_reduce(t, op=dist.ReduceOp.SUM, async_op=True), t)

    def finish_reduce(self, handle) -> None:
        work, t = handle
        if work is None:
            return
        work.wait()
        if t is not None:
            t.div_(self.world_size)

    def all_gather(self, src: torch.Tensor, dst: torch.Tensor) -> None:
        if "gather" in self._skip:
            return
        dist.all_gather_into_tensor(dst, src)

    def start_all_gather(self, src: torch.Tensor, dst: torch.Tensor):
        """Asynchronous all_gather (handle for finish_reduce): the lambda-values of Moments travel while the critic's own loss and
        backward, which do not need the other ranks' values, are computed."""
        if "gather" in self._skip:
            return (None, None)
        return (dist.all_gather_into_tensor(dst, src, async_op=True), None)
