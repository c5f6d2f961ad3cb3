"""Cluster bootstrap: the reference's ps/worker launch surface on top of ``torch.distributed``.

Reference (image_train.py:52-67): ``ps_hosts`` / ``worker_hosts`` are split on ``,`` into a
``tf.train.ClusterSpec{"ps", "worker"}``; every process starts an in-process gRPC server for its
``(job_name, task_index)``; ``ps`` tasks block in ``server.join()``; workers build the graph with
variables pinned to the ps job.  Here:

* the same four flags produce a ``ClusterSpec``; global rank = ps tasks first, then workers;
* rendezvous is a ``torch.distributed`` process group (gloo on CPU, NCCL on GPUs) whose store lives at
  the first ps host (first worker host when there is no ps job) -- the gRPC star becomes one TCP store;
* CPU mode: ps tasks own parameter shards and serve push/pull (parallel/ps.py);
* GPU mode: parameter shards live in the worker GPUs' NVLink-mapped memory (parallel/fused_comm.py),
  so a ps task has nothing to serve.  The world group (ps + workers) is gloo; every process -- ps tasks too,
  as non-members -- takes part in creating the NCCL sub-group of the worker ranks (``worker_group``) that
  carries the symmetric-memory rendezvous, the initial broadcast and the barriers.  A ps task then waits at
  two world barriers (workers started / workers finished) and exits: ``server.join()`` with a way out.
"""
from __future__ import annotations

import dataclasses
import datetime
import os
from typing import List, Optional


@dataclasses.dataclass
class ClusterSpec:
    ps: List[str]
    worker: List[str]

    @staticmethod
    def from_flags(ps_hosts: str, worker_hosts: str) -> "ClusterSpec":
        """``FLAGS.ps_hosts.split(',')`` (image_train.py:52-53); empty strings give an empty job."""
        ps = [h for h in ps_hosts.split(",") if h]
        worker = [h for h in worker_hosts.split(",") if h]
        return ClusterSpec(ps, worker)

    @property
    def num_ps(self) -> int:
        return len(self.ps)

    @property
    def num_workers(self) -> int:
        return max(len(self.worker), 1)

    @property
    def world_size(self) -> int:
        return self.num_ps + self.num_workers

    def global_rank(self, job_name: str, task_index: int) -> int:
        if job_name == "ps":
            if not 0 <= task_index < self.num_ps:
                raise ValueError(f"ps task_index {task_index} outside ps_hosts ({self.num_ps})")
            return task_index
        if job_name == "worker":
            if not 0 <= task_index < self.num_workers:
                raise ValueError(f"worker task_index {task_index} outside worker_hosts ({self.num_workers})")
            return self.num_ps + task_index
        raise ValueError(f"job_name must be 'ps' or 'worker', got {job_name!r}")

    def worker_rank(self, task_index: int) -> int:
        return self.num_ps + task_index

    def master_endpoint(self):
        host = (self.ps or self.worker or ["127.0.0.1:29500"])[0]
        name, _, port = host.rpartition(":")
        return (name or "127.0.0.1"), int(port)


@dataclasses.dataclass
class Role:
    job_name: str
    task_index: int
    rank: int
    world_size: int
    cluster: ClusterSpec

    @property
    def is_ps(self) -> bool:
        return self.job_name == "ps"

    @property
    def is_chief(self) -> bool:
        """``is_chief = (FLAGS.task_index == 0)`` for workers (image_train.py:123,139)."""
        return self.job_name == "worker" and self.task_index == 0

    @property
    def distributed(self) -> bool:
        return self.world_size > 1


def resolve_role(job_name: str, task_index: int, ps_hosts: str, worker_hosts: str) -> Role:
    """Role of this process.  With no cluster flags at all (single process) it is worker 0 / chief."""
    spec = ClusterSpec.from_flags(ps_hosts, worker_hosts)
    if not job_name:
        # also accept a torchrun launch: every rank is a worker
        if "RANK" in os.environ and "WORLD_SIZE" in os.environ and int(os.environ["WORLD_SIZE"]) > 1:
            w = int(os.environ["WORLD_SIZE"])
            addr = os.environ.get("MASTER_ADDR", "127.0.0.1")
            port = int(os.environ.get("MASTER_PORT", "29500"))
            spec = ClusterSpec([], [f"{addr}:{port + i}" for i in range(w)])
            r = int(os.environ["RANK"])
            return Role("worker", r, r, w, spec)
        job_name = "worker"
    if not spec.worker:
        spec = ClusterSpec(spec.ps, ["127.0.0.1:29500"])
    return Role(job_name, task_index, spec.global_rank(job_name, task_index), spec.world_size, spec)


def init_process_group(role: Role, backend: str, timeout_s: int = 1800, device=None) -> None:
    import torch.distributed as dist
    if not role.distributed or dist.is_initialized():
        return
    host, port = role.cluster.master_endpoint()
    kwargs = {}
    if backend == "nccl" and device is not None:
        kwargs["device_id"] = device
    dist.init_process_group(backend, init_method=f"tcp://{host}:{port}", rank=role.rank,
                            world_size=role.world_size, timeout=datetime.timedelta(seconds=timeout_s), **kwargs)


def worker_ranks(spec: ClusterSpec) -> List[int]:
    """Global ranks of the worker job (ps tasks come first in the rank order)."""
    return [spec.worker_rank(i) for i in range(spec.num_workers)]


def worker_group(role: Role, backend: str):
    """Process group of the worker ranks only.  COLLECTIVE over the whole world: every process of the cluster,
    ps tasks included, must call it (non-members get ``None``).  Without ps tasks the world group is returned."""
    import torch.distributed as dist
    if not role.distributed:
        return None
    if role.cluster.num_ps == 0:
        return dist.group.WORLD
    g = dist.new_group(ranks=worker_ranks(role.cluster), backend=backend)
    return None if role.is_ps else g


def shutdown() -> None:
    import torch.distributed as dist
    if dist.is_initialized():
        dist.destroy_process_group()
