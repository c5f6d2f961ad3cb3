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
    def __init__(self, device, sync_bn: bool = False):
        self.device, self.sync_bn = device, sync_bn
        self.world = dist.get_world_size()
        self.rank = dist.get_rank()

    def attach(self, engine) -> None:
        # identical initial weights on every replica (the chief's init in the reference, image_train.py:117-125)
        dist.broadcast(engine.arena.master, src=0)
        engine.refresh_compute_weights()

    def allreduce_bn(self, buf: torch.Tensor) -> None:
        dist.all_reduce(buf)

    def optimizer_step(self, engine, which: str = "both") -> None:
        A = engine.arena
        lo, hi = {"both": (0, A.padded), "g": (0, A.g_end), "d": (A.g_end, A.padded)}[which]
        dist.all_reduce(engine.grads[lo:hi])
        K.adam(A.master[lo:hi], engine.grads[lo:hi], A.m[lo:hi], A.v[lo:hi], A.bf16[lo:hi], engine.counters,
               engine.lr, engine.beta1, engine.beta2, engine.eps, 1.0 / self.world)


def make_comm(kind: str, device, sync_bn: bool = False):
    if kind == "nccl":
        return NcclComm(device, sync_bn)
    from .fused_comm import FusedComm
    return FusedComm(device, sync_bn)
