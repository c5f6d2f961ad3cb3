"""Fused communication over NVLink peer memory (the product path; NCCL is only the baseline).

``FusedComm`` plugs into the engine as ``engine.comm``:

* it hands the engine *symmetric* (peer-mapped) memory for the buffers other ranks must reach -- the
  flat gradient region and the bf16 parameter copy -- allocated and exchanged once with
  ``torch.distributed._symmetric_memory`` (``torch.distributed`` is the plumbing; no collective runs
  through it on the hot path);
* ``launch_bucket`` runs ONE kernel per bucket (D variables, G variables) that does the gradient
  reduce-scatter + average + partitioned TF-form Adam + bf16 parameter all-gather
  (csrc/comm_sm100.cu: ``rs_adam_ag_kernel``); the D bucket is launched on a side stream as soon as
  D's backward is done, so it overlaps G's backward pass;
* f32 master weights and both Adam slots exist only on the owning rank's shard -- the B200 analogue of
  the reference's variables-on-ps placement with real sharding; ``gather_state``
  reassembles them for checkpoints;
* ``allreduce_bn`` averages batch-norm statistics across replicas (BASELINE.json config 3).
"""
from __future__ import annotations

import ctypes
import os
from typing import Dict

import torch
import torch.distributed as dist

from ..ops import _lib

SLOT_G, SLOT_D, SLOT_BN, SLOT_G1 = 0, 1, 2, 3
N_SLOTS = 4
MAX_RANKS = 16
BN_STAGING = 8192  # floats per parity (largest site: 3 slabs x 2 x 1024 channels)


def _ptr_array(ptrs):
    arr = (ctypes.c_void_p * len(ptrs))(*[int(p) for p in ptrs])
    return arr


class FusedComm:
    def __init__(self, device, sync_bn: bool = False, overlap: bool = True, use_multicast: bool = True,
                 blocks: int = 64, gather_in_consumer: bool = False, group=None):
        import torch.distributed._symmetric_memory as symm_mem
        self.symm_mem = symm_mem
        self.device, self.sync_bn, self.overlap = device, sync_bn, overlap
        # the worker ranks only: with ps tasks in the launch (CPU joiners) this is an NCCL sub-group of the world
        self.group = group if group is not None else dist.group.WORLD
        self.world, self.rank = dist.get_world_size(self.group), dist.get_rank(self.group)
        self.chief = dist.get_global_rank(self.group, 0)
        if self.world > MAX_RANKS:
            raise ValueError(f"at most {MAX_RANKS} ranks per NVLink domain")
        if self.group is not dist.group.WORLD and hasattr(symm_mem, "enable_symm_mem_for_group"):
            symm_mem.enable_symm_mem_for_group(self.group.group_name)
        self.handles: Dict[str, object] = {}
        self.tensors: Dict[str, torch.Tensor] = {}
        self.use_multicast = use_multicast
        self.blocks = int(os.environ.get("DCGAN_COMM_BLOCKS", blocks))  # CTAs of an exchange that overlaps compute
        # Measured at 2 GPUs (profiles/mg_ab_n2_r2.log): pushing g_h0_lin/Matrix with the rest of its bucket beats
        # pulling it inside lin_fwd (whose loop then pays NVLink latency at the head of the step): 0.707 vs 0.723 ms,
        # so the consumer-side gather is opt-in (DCGAN_GATHER_IN_CONSUMER=1)
        env = os.environ.get("DCGAN_GATHER_IN_CONSUMER")
        self.gather_in_consumer = gather_in_consumer if env is None else env != "0"
        self.overlap = overlap and os.environ.get("DCGAN_COMM_OVERLAP", "1") != "0"
        self.stream = torch.cuda.Stream(device)
        self._fork_event = torch.cuda.Event()
        self._join_event = torch.cuda.Event()
        self._forked = False
        # control block: signal pads + batch-norm staging, symmetric
        self.sig = self._symmetric("signals", N_SLOTS * 2 * MAX_RANKS, torch.int32)
        self.bn_staging = self._symmetric("bn_staging", 2 * BN_STAGING, torch.float32)
        self.state = torch.zeros(N_SLOTS, 2, dtype=torch.int32, device=device)
        self._sig_ptrs = _ptr_array(self.handles["signals"].signal_ptrs if False else self.handles["signals"].buffer_ptrs)
        self._bn_ptrs = _ptr_array(self.handles["bn_staging"].buffer_ptrs)

    # ---- symmetric allocation -----------------------------------------------------------------------------------
    def _symmetric(self, name: str, numel: int, dtype) -> torch.Tensor:
        t = self.symm_mem.empty(numel, dtype=dtype, device=self.device)
        t.zero_()
        torch.cuda.synchronize(self.device)
        self.handles[name] = self.symm_mem.rendezvous(t, self.group.group_name)
        self.tensors[name] = t
        return t

    def alloc(self, name: str, numel: int, dtype) -> torch.Tensor:
        """Engine allocation hook: buffers that peers read or write live in symmetric memory."""
        if name in ("zero_region", "bf16", "master", "adam_m", "adam_v"):
            # gradients / bf16 parameters are exchanged every step; the sharded f32 state is peer-mapped too so
            # that a checkpoint can read the owners' shards without their participation
            return self._symmetric(name, numel, dtype)
        return torch.zeros(numel, dtype=dtype, device=self.device)

    def attach(self, engine) -> None:
        A = engine.arena
        dist.broadcast(A.master, src=self.chief, group=self.group)  # the chief's initialisation (image_train.py:117,125)
        engine.refresh_compute_weights()
        self._grad_ptrs = _ptr_array(self.handles["zero_region"].buffer_ptrs)
        self._bf16_ptrs = _ptr_array(self.handles["bf16"].buffer_ptrs)
        self._master_ptrs = _ptr_array(self.handles["master"].buffer_ptrs)
        hg, hb, hm = self.handles["zero_region"], self.handles["bf16"], self.handles["master"]
        mc_ok = (self.use_multicast and getattr(hg, "has_multicast_support", False) and hg.multicast_ptr
                 and hb.multicast_ptr and hm.multicast_ptr)
        self.mc_grad = int(hg.multicast_ptr) if mc_ok else None
        self.mc_bf16 = int(hb.multicast_ptr) if mc_ok else None
        self.mc_master = int(hm.multicast_ptr) if mc_ok else None
        # Buckets = ranges of the arena exchanged by one kernel launch each, every one sharded over the ranks.  The
        # generator is split in two: "g1" = g_h1/w .. g_h3/w (85 % of its parameters; their gradients are complete
        # while the tail of the backward pass still runs, so this exchange overlaps it) and "g2" = the rest
        # (g_h0_lin/Matrix, whose gradient is the last kernel of the step, and the small f32 variables).
        split = A.offsets["g_h0_lin/Matrix"][0] if os.environ.get("DCGAN_COMM_SPLIT_G", "1") != "0" else 0
        if 0 < split < A.f32_start["g"] and split % 8 == 0:
            self.buckets = {"g1": (0, split), "g2": (split, A.g_end), "d": (A.g_end, A.padded)}
            self.slots = {"g1": SLOT_G1, "g2": SLOT_G, "d": SLOT_D}
            # f32-consumed variables start here inside each bucket (their master value is gathered too)
            self.f32_from = {"g1": split, "g2": A.f32_start["g"], "d": A.f32_start["d"]}
            self.parts = {"g": ("g1", "g2"), "d": ("d",), "g1": ("g1",), "g2": ("g2",), "both": ("d", "g1", "g2")}
        else:
            self.buckets = {"g": (0, A.g_end), "d": (A.g_end, A.padded)}
            self.slots = {"g": SLOT_G, "d": SLOT_D}
            self.f32_from = {"g": A.f32_start["g"], "d": A.f32_start["d"]}
            self.parts = {"g": ("g",), "d": ("d",), "both": ("d", "g")}
        self.bucket_names = tuple(self.buckets)
        # g_h0_lin/Matrix is gathered by its consumer (lin_fwd_gather), not pushed by the optimizer kernel
        off, shape = A.offsets["g_h0_lin/Matrix"]
        self.pull_range = (off, off + shape[0] * shape[1]) if self.gather_in_consumer else (0, 0)
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.group)

    def set_peer_timeout(self, seconds: float) -> None:
        """Bound of every wait for the peers' signals (about 0.7 us per poll)."""
        polls = int(min(max(seconds, 1.0) / 0.7e-6, 2 ** 32 - 1))
        _lib.load().dcgan_comm_set_peer_spin_limit(ctypes.c_uint(polls))

    def owner_ranks(self, lo: int, hi: int):
        """Ranks whose optimizer shard intersects the arena range [lo, hi)."""
        out = []
        for bucket in self.bucket_names:
            for r in range(self.world):
                s0, s1 = self.shard_range(bucket, r)
                if s0 < hi and lo < s1 and r not in out:
                    out.append(r)
        return sorted(out)

    # ---- the fused optimizer step ---------------------------------------------------------------------------------
    def _launch(self, engine, which: str, blocks: int = 0) -> None:
        A = engine.arena
        lo, hi = self.buckets[which]
        slot = self.slots[which]
        pull_lo, pull_hi = self.pull_range if (lo <= self.pull_range[0] and self.pull_range[1] <= hi) else (0, 0)
        _lib.call("dcgan_comm_rs_adam_ag", self._grad_ptrs, self._bf16_ptrs, self._sig_ptrs, self._master_ptrs,
                  A.master.data_ptr(), A.m.data_ptr(), A.v.data_ptr(), engine.counters.data_ptr(),
                  self.state[slot].data_ptr(), self.rank, self.world, lo, hi, engine.lr, engine.beta1, engine.beta2,
                  engine.eps, slot, self.mc_grad, self.mc_bf16, self.mc_master, self.f32_from[which], pull_lo,
                  pull_hi, blocks or self.blocks)

    def lin_fwd_gather(self, engine, out2d: torch.Tensor) -> None:
        """g_h0_lin forward that pulls its weight slices from the owning ranks (parameter all-gather fused into
        the first consumer).  Falls back to the local shadow when the feature is off."""
        from ..ops import kernels as K
        A = engine.arena
        if not self.gather_in_consumer:
            K.lin_fwd(engine.z, engine.w_native["g_h0_lin/Matrix"], engine.params["g_h0_lin/bias"], out2d)
            return
        off, shape = A.offsets["g_h0_lin/Matrix"]
        lo, hi = next(self.buckets[b] for b in self.bucket_names
                      if self.buckets[b][0] <= off and off + shape[0] * shape[1] <= self.buckets[b][1])
        _lib.call("dcgan_comm_lin_fwd_gather", self._bf16_ptrs, engine.z.data_ptr(),
                  engine.params["g_h0_lin/bias"].data_ptr(), out2d.data_ptr(), engine.B, shape[0], shape[1], off, lo,
                  hi, self.rank, self.world)

    def launch_bucket(self, engine, which: str, side_stream: bool, also_wait=None) -> None:
        if side_stream:
            cur = torch.cuda.current_stream(self.device)
            self._fork_event.record(cur)
            self.stream.wait_event(self._fork_event)
            if also_wait is not None:   # e.g. the engine's side stream that carries the filter gradients
                self.stream.wait_event(also_wait)
            with torch.cuda.stream(self.stream):
                for part in self.parts[which]:
                    self._launch(engine, part)
                self._join_event.record(self.stream)
            self._forked = True
        else:
            # nothing left to overlap with: give the exchange the whole GPU
            for part in self.parts[which]:
                self._launch(engine, part, blocks=int(os.environ.get("DCGAN_COMM_BLOCKS_ALONE", 148)))

    def has_early_generator_bucket(self) -> bool:
        return "g1" in self.buckets

    def join(self) -> None:
        if self._forked:
            torch.cuda.current_stream(self.device).wait_event(self._join_event)
            self._forked = False

    def optimizer_step(self, engine, which: str = "both") -> None:
        for part in self.parts[which]:
            self._launch(engine, part)

    # ---- cross-replica batch norm -----------------------------------------------------------------------------------
    def allreduce_bn(self, buf: torch.Tensor) -> None:
        n = buf.numel()
        assert n <= BN_STAGING
        _lib.call("dcgan_comm_allreduce_small", self._bn_ptrs, self._sig_ptrs, buf.data_ptr(), n, BN_STAGING,
                  self.state[SLOT_BN].data_ptr(), self.rank, self.world, SLOT_BN)

    # ---- sharded state -> full state (checkpoints) ---------------------------------------------------------------------
    def shard_range(self, bucket: str, rank: int):
        lo, hi = self.buckets[bucket]
        total4 = (hi - lo) // 4
        per4 = (((total4 + self.world - 1) // self.world) + 1) & ~1  # even: shard_quads() in comm_sm100.cu
        return lo + min(rank * per4, total4) * 4, lo + min((rank + 1) * per4, total4) * 4

    def gather_state(self, engine) -> None:
        """Fills this rank's master / Adam slots with the owners' shards, read through peer memory.  Not a
        collective: only the caller (the chief, at a step boundary) takes part.  In sync mode the owners cannot
        touch their shards until the caller joins the next step's exchange, so the copy is consistent."""
        A = engine.arena
        torch.cuda.synchronize(self.device)
        for name, local in (("master", A.master), ("adam_m", A.m), ("adam_v", A.v)):
            hdl = self.handles[name]
            for r in range(self.world):
                if r == self.rank:
                    continue
                peer = hdl.get_buffer(r, (A.padded,), torch.float32)
                for bucket in self.bucket_names:
                    s0, s1 = self.shard_range(bucket, r)
                    if s1 > s0:
                        local[s0:s1].copy_(peer[s0:s1])
        torch.cuda.synchronize(self.device)

    def close(self) -> None:
        torch.cuda.synchronize(self.device)
