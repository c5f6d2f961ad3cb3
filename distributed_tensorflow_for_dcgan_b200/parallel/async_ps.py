"""Asynchronous parameter-server mode on GPUs (BASELINE.json config 4).

Rank r *owns* shard r of the flat parameter space -- f32 master weights and both Adam slots -- in
peer-mapped HBM: that is the "variables live on the ps job" placement of the reference
(distriubted_model.py:66-72) with real sharding.  After its backward pass a worker runs one kernel
(csrc/comm_sm100.cu: ``async_apply_kernel``) that, for every element, loads (p, m, v) from the owner
over NVLink, applies TF-form Adam with its *own* gradient -- stale with respect to whatever the other
workers did meanwhile -- stores the result back and keeps the value as its fresh local bf16 copy.
There is no barrier, no lock and no aggregation, like the reference's ``use_locking=False`` applies
(SURVEY.md 2.3); ``global_step`` (rank 0) and the per-shard Adam step count advance once per push.
"""
from __future__ import annotations

import ctypes
import os
from typing import Dict

import torch
import torch.distributed as dist

from ..ops import _lib
from .fused_comm import _ptr_array


class AsyncPSComm:
    sync_bn = False

    def __init__(self, device, blocks: int = 148, group=None, overlap: bool = True, overlap_blocks: int = 64):
        import torch.distributed._symmetric_memory as symm_mem
        self.symm_mem = symm_mem
        self.device, self.blocks = device, blocks
        # like the synchronous path, the push of D's variables runs on a stream of its own under G's backward pass
        self.overlap = overlap and os.environ.get("DCGAN_COMM_OVERLAP", "1") != "0"
        self.overlap_blocks = int(os.environ.get("DCGAN_COMM_BLOCKS", overlap_blocks))
        self.stream = torch.cuda.Stream(device)
        self._fork_event, self._join_event, self._forked = torch.cuda.Event(), torch.cuda.Event(), False
        # Deferred generator push: the generator's gradients of step k are pushed at the START of step k + 1, on the
        # push stream, under that step's forward passes and D's backward pass (they are the last thing a step
        # computes, so pushed at once they would have nothing to hide behind).  The worker's own generator update
        # therefore lands one step late -- one more step of the staleness this mode is made of; its reads of the
        # generator's parameters meanwhile are as unlocked as the reference's (use_locking=False).
        self.deferred_g = self.overlap and os.environ.get("DCGAN_ASYNC_DEFER_G", "1") != "0"
        self.gprev = None
        self.group = group if group is not None else dist.group.WORLD   # the worker ranks (ps tasks excluded)
        self.world, self.rank = dist.get_world_size(self.group), dist.get_rank(self.group)
        self.chief = dist.get_global_rank(self.group, 0)
        if self.group is not dist.group.WORLD and hasattr(symm_mem, "enable_symm_mem_for_group"):
            symm_mem.enable_symm_mem_for_group(self.group.group_name)
        self.handles: Dict[str, object] = {}

    def _symmetric(self, name: str, numel: int, dtype) -> torch.Tensor:
        t = self.symm_mem.empty(numel, dtype=dtype, device=self.device)
        t.zero_()
        torch.cuda.synchronize(self.device)
        self.handles[name] = self.symm_mem.rendezvous(t, self.group.group_name)
        return t

    def alloc(self, name: str, numel: int, dtype) -> torch.Tensor:
        if name in ("master", "adam_m", "adam_v", "bf16"):
            return self._symmetric(name, numel, dtype)
        return torch.zeros(numel, dtype=dtype, device=self.device)

    def attach(self, engine) -> None:
        A = engine.arena
        dist.broadcast(A.master, src=self.chief, group=self.group)
        engine.refresh_compute_weights()
        self._p = _ptr_array(self.handles["master"].buffer_ptrs)
        self._b = _ptr_array(self.handles["bf16"].buffer_ptrs)
        self._m = _ptr_array(self.handles["adam_m"].buffer_ptrs)
        self._v = _ptr_array(self.handles["adam_v"].buffer_ptrs)
        # per-shard push counters ([0] = Adam t of this rank's shard, rank 0's [1] = global_step), symmetric
        self.shard_counters = self._symmetric("shard_counters", 2, torch.int64)
        self._counter_ptrs = torch.tensor([int(p) for p in self.handles["shard_counters"].buffer_ptrs],
                                          dtype=torch.int64, device=self.device)
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.group)

    def seed_counters(self, engine) -> None:
        """After a restore / the chief's broadcast: global_step and the per-shard Adam step live in the shard
        counters in this mode, so they continue from the restored values instead of restarting at zero."""
        adam_t, gstep = (int(v) for v in engine.counters.tolist())
        self.shard_counters[0] = adam_t
        self.shard_counters[1] = gstep if self.rank == 0 else 0
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.group)

    def _apply(self, engine, lo: int, hi: int, advance: bool, blocks: int, grads=None) -> None:
        A = engine.arena
        _lib.call("dcgan_comm_async_apply", self._p, self._b, self._m, self._v,
                  (engine.grads if grads is None else grads).data_ptr(),
                  A.bf16.data_ptr(), A.master.data_ptr(), self._counter_ptrs.data_ptr(), self.rank, self.world,
                  A.padded, lo, hi, int(advance), engine.lr, engine.beta1, engine.beta2, engine.eps, blocks)

    def optimizer_step(self, engine, which: str = "both") -> None:
        A = engine.arena
        lo, hi = {"both": (0, A.padded), "g": (0, A.g_end), "d": (A.g_end, A.padded)}[which]
        self._apply(engine, lo, hi, which != "d", self.blocks)   # one advance per step: with G's (or the only) push

    def launch_bucket(self, engine, which: str, side_stream: bool, also_wait=None) -> None:
        """Same protocol as FusedComm.launch_bucket: 'd' under G's backward pass on the push stream, 'g' last."""
        A = engine.arena
        lo, hi = {"g": (0, A.g_end), "d": (A.g_end, A.padded)}[which]
        if side_stream:
            cur = torch.cuda.current_stream(self.device)
            self._fork_event.record(cur)
            self.stream.wait_event(self._fork_event)
            if also_wait is not None:
                self.stream.wait_event(also_wait)
            with torch.cuda.stream(self.stream):
                self._apply(engine, lo, hi, which != "d", self.overlap_blocks)
                self._join_event.record(self.stream)
            self._forked = True
        else:
            self._apply(engine, lo, hi, which != "d", self.blocks)

    def join(self) -> None:
        if self._forked:
            torch.cuda.current_stream(self.device).wait_event(self._join_event)
            self._forked = False

    # ---- deferred generator push ----------------------------------------------------------------------------------
    def launch_deferred_g(self, engine) -> None:
        """Start of a step: push the generator gradients stashed by the previous step (zeros the first time: an
        Adam step on zero moments and a zero gradient moves nothing)."""
        A = engine.arena
        if self.gprev is None:
            self.gprev = torch.zeros(A.g_end, dtype=torch.float32, device=self.device)
        cur = torch.cuda.current_stream(self.device)
        self._fork_event.record(cur)
        self.stream.wait_event(self._fork_event)
        with torch.cuda.stream(self.stream):
            self._apply(engine, 0, A.g_end, True, self.overlap_blocks, grads=self.gprev)
            self._join_event.record(self.stream)
        self._forked = True

    def stash_g(self, engine) -> None:
        """End of a step (after ``join``): keep the generator's gradients for the next step's push."""
        self.gprev.copy_(engine.grads[:engine.arena.g_end])

    def reset_deferred(self) -> None:
        if self.gprev is not None:
            self.gprev.zero_()

    def flush(self, engine) -> None:
        """Push what ``stash_g`` holds now (end of training / before a checkpoint)."""
        if self.deferred_g and self.gprev is not None:
            self._apply(engine, 0, engine.arena.g_end, False, self.blocks, grads=self.gprev)
            self.gprev.zero_()

    def global_step(self) -> int:
        """Applies of all workers so far: rank 0's counter, read through peer memory from any rank."""
        torch.cuda.synchronize(self.device)
        if self.rank == 0:
            return int(self.shard_counters[1].item())
        return int(self.handles["shard_counters"].get_buffer(0, (2,), torch.int64)[1].item())

    def gather_state(self, engine) -> None:
        """Copies the owners' shards into this rank's arrays through peer memory (not a collective; other
        workers may keep pushing, so the snapshot is as loosely consistent as the mode itself)."""
        A = engine.arena
        total4 = A.padded // 4
        per4 = (((total4 + self.world - 1) // self.world) + 1) & ~1  # even: shard_quads() in comm_sm100.cu
        torch.cuda.synchronize(self.device)
        for name, local in (("master", A.master), ("adam_m", A.m), ("adam_v", A.v)):
            hdl = self.handles[name]
            for r in range(self.world):
                if r == self.rank:
                    continue
                s0, s1 = min(r * per4, total4) * 4, min((r + 1) * per4, total4) * 4
                if s1 > s0:
                    local[s0:s1].copy_(hdl.get_buffer(r, (A.padded,), torch.float32)[s0:s1])
        torch.cuda.synchronize(self.device)

    def close(self) -> None:
        # a worker may only release its shard once nobody pushes to it any more
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.group)
