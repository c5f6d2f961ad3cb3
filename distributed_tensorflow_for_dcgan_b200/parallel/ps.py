"""Parameter-server plumbing for the CPU / gloo path (BASELINE.json config 1 and the async mode).

Reference semantics: variables live on the ps job; each worker reads them,
computes D and G gradients on its own batch and the Adam applies run on the ps -- with no
aggregation, barrier or lock (asynchronous, Hogwild-style).  ``global_step`` is a ps variable that
only ``g_optim`` increments, so it counts applies of *all* workers.

What is built here:

* real parameter sharding: variables are spread over the ps tasks by size (the reference pins every
  weight to ``/job:ps/task:0``, distriubted_model.py:70 -- extra ps tasks there hold almost nothing);
* one request/response protocol over gloo point-to-point messages serves both modes:
    - async (``sync=False``): a push is applied the moment it arrives, the reply carries whatever the
      parameters are now -> stale gradients, ``global_step`` advances once per push;
    - sync  (``sync=True``): the shard waits for one push from every worker, applies Adam once on
      the mean gradient and answers all of them -> ``SyncReplicas`` semantics, ``global_step``
      advances once per round;
* Adam slots live with the shard (``<var>/Adam``, ``<var>/Adam_1``), like TF's slot colocation.

Pushes to one shard are applied one at a time (no torn updates); the reference's unlocked races
are not reproduced on purpose.
"""
from __future__ import annotations

import queue
import threading
import time
from typing import Dict, List

import torch
import torch.distributed as dist

from ..models import manifest
from ..optim.adam_tf import adam_tf_step_

OP_PUSH, OP_PULL, OP_SET, OP_GET_SLOTS, OP_SET_SLOTS, OP_DONE = 1, 2, 3, 4, 5, 6
OP_DEAD = -1  # internal: the receiver thread lost the connection to a worker in the middle of a request


def shard_variables(specs, num_ps: int) -> List[List[str]]:
    """Greedy size-balanced assignment of variables to ps tasks (deterministic)."""
    shards: List[List[str]] = [[] for _ in range(max(num_ps, 1))]
    load = [0] * max(num_ps, 1)
    for s in sorted(specs, key=lambda s: (-s.numel, s.name)):
        i = load.index(min(load))
        shards[i].append(s.name)
        load[i] += s.numel
    order = {s.name: k for k, s in enumerate(specs)}
    return [sorted(sh, key=order.get) for sh in shards]


class _Flat:
    """Flat f32 view of an ordered list of named tensors."""

    def __init__(self, names: List[str], shapes: Dict[str, tuple]):
        self.names = names
        self.shapes = shapes
        self.sizes = [int(torch.Size(shapes[n]).numel()) for n in names]
        self.numel = sum(self.sizes)

    def pack(self, tensors: Dict[str, torch.Tensor]) -> torch.Tensor:
        if not self.names:
            return torch.zeros(0)
        return torch.cat([tensors[n].detach().reshape(-1).float().cpu() for n in self.names])

    def unpack_into(self, flat: torch.Tensor, tensors: Dict[str, torch.Tensor]) -> None:
        off = 0
        for n, sz in zip(self.names, self.sizes):
            tensors[n].copy_(flat[off:off + sz].view(self.shapes[n]))
            off += sz


def _send_header(dst: int, op: int, a: int = 0, b: int = 0) -> None:
    dist.send(torch.tensor([op, a, b, 0], dtype=torch.int64), dst)


class PSServer:
    """Runs in a ``--job_name=ps`` process: ``serve()`` is the analogue of ``server.join()``
    (image_train.py:62-63) except that it returns once every worker has said goodbye."""

    def __init__(self, cfg, ps_index: int, num_ps: int, worker_ranks: List[int], sync: bool, lr: float = 2e-4,
                 beta1: float = 0.5, beta2: float = 0.999, eps: float = 1e-8, worker_timeout: float = 0.0):
        specs = manifest.trainable_specs(cfg)
        shapes = {s.name: s.shape for s in specs}
        self.names = shard_variables(specs, num_ps)[ps_index]
        self.flat = _Flat(self.names, shapes)
        self.p = torch.zeros(self.flat.numel)
        self.m = torch.zeros(self.flat.numel)
        self.v = torch.zeros(self.flat.numel)
        self.t = 0
        self.global_step = 0
        self.ps_index, self.workers, self.sync = ps_index, list(worker_ranks), sync
        self.lr, self.beta1, self.beta2, self.eps = lr, beta1, beta2, eps
        self.applies = 0
        # failure detection: the reference's ps blocks in server.join() forever and a dead worker
        # stalls a synchronous round for good.  Here a worker that has been silent for `worker_timeout` seconds
        # while somebody is waiting for it -- a sync round it has not joined, or a shutdown that only misses its
        # goodbye -- is presumed dead and dropped; 0 keeps the wait-forever behaviour.
        self.worker_timeout = float(worker_timeout)
        self.dropped: List[int] = []

    def _receiver(self, inbox: "queue.Queue") -> None:
        """Receiver thread: reads whole requests (header + payload) from any worker into ``inbox``.  gloo cannot
        time out a receive and carry on (a timed-out wait leaves the pair unusable), so the blocking receive
        lives here and the serving loop waits on the queue, where a timeout is harmless."""
        while True:
            header = torch.zeros(4, dtype=torch.int64)
            src = None
            try:
                src = dist.recv(header)
                op = int(header[0])
                payload = []
                if self.flat.numel:
                    n = {OP_PUSH: 1, OP_SET: 1, OP_SET_SLOTS: 2}.get(op, 0)
                    for _ in range(n):
                        buf = torch.zeros(self.flat.numel)
                        dist.recv(buf, src)
                        payload.append(buf)
                inbox.put((src, op, header, payload))
            except RuntimeError as e:
                # a worker died in the middle of a request ("Connection closed by peer"): tell the serving loop
                # which one, then keep receiving from the others
                if src is not None:
                    inbox.put((src, OP_DEAD, header, [str(e)]))
                else:
                    time.sleep(0.2)

    def _apply(self, grad: torch.Tensor) -> None:
        self.t += 1
        adam_tf_step_(self.p, grad, self.m, self.v, self.t, self.lr, self.beta1, self.beta2, self.eps)
        self.global_step += 1  # g_optim's apply (image_train.py:112); every shard keeps the same count
        self.applies += 1

    def _reply(self, dst: int) -> None:
        dist.send(torch.tensor([self.global_step, self.t, 0, 0], dtype=torch.int64), dst)
        if self.flat.numel:
            dist.send(self.p, dst)

    def serve(self) -> int:
        alive = set(self.workers)
        pending: Dict[int, torch.Tensor] = {}
        inbox: "queue.Queue" = queue.Queue()
        threading.Thread(target=self._receiver, args=(inbox,), daemon=True, name="ps-receiver").start()
        last_seen = {w: time.time() for w in alive}
        said_goodbye = False
        while alive:
            try:
                src, op, header, payload = inbox.get(timeout=1.0 if self.worker_timeout > 0 else None)
            except queue.Empty:
                # somebody waits for the silent ones: an open sync round, or a shutdown already under way
                if not (pending or said_goodbye):
                    continue
                now = time.time()
                for w in [w for w in alive if w not in pending and now - last_seen[w] > self.worker_timeout]:
                    print(f"ps {self.ps_index}: worker rank {w} silent for more than {self.worker_timeout:.0f} s, "
                          f"presumed dead and dropped", flush=True)
                    alive.discard(w)
                    self.dropped.append(w)
                if self.sync and pending and set(pending) >= alive and alive:
                    self._finish_round(pending)
                continue
            last_seen[src] = time.time()
            if op == OP_DEAD:
                if src in alive:
                    print(f"ps {self.ps_index}: lost worker rank {src} in the middle of a request, dropped", flush=True)
                    alive.discard(src)
                    pending.pop(src, None)
                    self.dropped.append(src)
                    if self.sync and pending and set(pending) >= alive and alive:
                        self._finish_round(pending)
            elif op == OP_DONE:
                alive.discard(src)
                said_goodbye = True
                if self.sync and pending and set(pending) >= alive and alive:
                    self._finish_round(pending)
            elif op == OP_PUSH:
                grad = payload[0] if payload else torch.zeros(0)
                if self.sync:
                    pending[src] = grad
                    if set(pending) >= alive:
                        self._finish_round(pending)
                else:
                    self._apply(grad)
                    self._reply(src)
            elif op == OP_PULL:
                self._reply(src)
            elif op == OP_SET:
                if payload:
                    self.p.copy_(payload[0])
                self.global_step, self.t = int(header[1]), int(header[2])
            elif op == OP_SET_SLOTS:
                if payload:
                    self.m.copy_(payload[0])
                    self.v.copy_(payload[1])
            elif op == OP_GET_SLOTS:
                if self.flat.numel:
                    dist.send(self.m, src)
                    dist.send(self.v, src)
            else:
                raise RuntimeError(f"ps {self.ps_index}: unknown op {op} from rank {src}")
        return self.applies

    def _finish_round(self, pending: Dict[int, torch.Tensor]) -> None:
        grads = list(pending.values())
        mean = torch.stack(grads).mean(0) if self.flat.numel else torch.zeros(0)
        self._apply(mean)
        for dst in list(pending):
            self._reply(dst)
        pending.clear()


class PSClient:
    """Worker side: pushes gradients for every shard, pulls the (possibly stale) parameters back."""

    def __init__(self, cfg, num_ps: int, params: Dict[str, torch.Tensor]):
        specs = manifest.trainable_specs(cfg)
        shapes = {s.name: s.shape for s in specs}
        self.params = params
        self.shards = [_Flat(names, shapes) for names in shard_variables(specs, num_ps)]
        self.num_ps = num_ps
        self.global_step = 0
        self.t = 0

    def _recv_reply(self, ps: int) -> None:
        hdr = torch.zeros(4, dtype=torch.int64)
        dist.recv(hdr, ps)
        self.global_step, self.t = int(hdr[0]), int(hdr[1])
        fl = self.shards[ps]
        if fl.numel:
            buf = torch.zeros(fl.numel)
            dist.recv(buf, ps)
            with torch.no_grad():
                fl.unpack_into(buf, self.params)

    def push_pull(self, grads: Dict[str, torch.Tensor]) -> int:
        for ps, fl in enumerate(self.shards):
            _send_header(ps, OP_PUSH)
            if fl.numel:
                dist.send(fl.pack(grads), ps)
        for ps in range(self.num_ps):
            self._recv_reply(ps)
        return self.global_step

    def pull(self) -> int:
        for ps in range(self.num_ps):
            _send_header(ps, OP_PULL)
        for ps in range(self.num_ps):
            self._recv_reply(ps)
        return self.global_step

    def set_params(self, global_step: int = 0, t: int = 0) -> None:
        """Chief only: initialise / restore the ps variables (image_train.py:117,125,143)."""
        for ps, fl in enumerate(self.shards):
            _send_header(ps, OP_SET, global_step, t)
            if fl.numel:
                dist.send(fl.pack(self.params), ps)

    def set_slots(self, m: Dict[str, torch.Tensor], v: Dict[str, torch.Tensor]) -> None:
        for ps, fl in enumerate(self.shards):
            _send_header(ps, OP_SET_SLOTS)
            if fl.numel:
                dist.send(fl.pack(m), ps)
                dist.send(fl.pack(v), ps)

    def get_slots(self):
        m = {n: torch.zeros_like(t) for n, t in self.params.items()}
        v = {n: torch.zeros_like(t) for n, t in self.params.items()}
        for ps, fl in enumerate(self.shards):
            _send_header(ps, OP_GET_SLOTS)
            if fl.numel:
                bm, bv = torch.zeros(fl.numel), torch.zeros(fl.numel)
                dist.recv(bm, ps)
                dist.recv(bv, ps)
                fl.unpack_into(bm, m)
                fl.unpack_into(bv, v)
        return m, v

    def done(self) -> None:
        for ps in range(self.num_ps):
            _send_header(ps, OP_DONE)
