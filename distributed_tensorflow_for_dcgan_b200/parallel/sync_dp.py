"""Synchronous data-parallel optimizer step (BASELINE.json config 2; SURVEY.md 2.3 "DP -- synchronous").

Two interchangeable backends behind ``engine.comm``:

* ``FusedComm``  -- the product path: one hand-written kernel does reduce-scatter of the gradients over
  NVLink peer memory + gradient average + partitioned TF-form Adam on the rank's parameter shard (the
  "ps shard") + all-gather of the updated bf16 parameters into every peer.  No NCCL call.
* ``NcclComm``   -- the baseline / fallback: ``ncclAllReduce`` on the flat gradient buffer followed by the
  single-GPU fused Adam kernel.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from ..ops import kernels as K


class NcclComm:
    def __init__(self, device, sync_bn: bool = False, group=None):
        self.device, self.sync_bn = device, sync_bn
        self.group = group if group is not None else dist.group.WORLD   # the worker ranks (ps tasks excluded)
        self.world = dist.get_world_size(self.group)
        self.rank = dist.get_rank(self.group)
        self.chief = dist.get_global_rank(self.group, 0)

    def attach(self, engine) -> None:
        # identical initial weights on every replica (the chief's init in the reference, image_train.py:117-125)
        dist.broadcast(engine.arena.master, src=self.chief, group=self.group)
        engine.refresh_compute_weights()

    def allreduce_bn(self, buf: torch.Tensor) -> None:
        # mean over replicas: the batch-norm kernels divide by the LOCAL row count (FusedComm averages as well)
        dist.all_reduce(buf, group=self.group)
        buf.div_(self.world)

    def optimizer_step(self, engine, which: str = "both") -> None:
        A = engine.arena
        lo, hi = {"both": (0, A.padded), "g": (0, A.g_end), "d": (A.g_end, A.padded)}[which]
        dist.all_reduce(engine.grads[lo:hi], group=self.group)
        K.adam(A.master[lo:hi], engine.grads[lo:hi], A.m[lo:hi], A.v[lo:hi], A.bf16[lo:hi], engine.counters,
               engine.lr, engine.beta1, engine.beta2, engine.eps, 1.0 / self.world)


def make_comm(kind: str, device, sync_bn: bool = False, group=None):
    if kind == "nccl":
        return NcclComm(device, sync_bn, group=group)
    from .fused_comm import FusedComm
    return FusedComm(device, sync_bn, group=group)
