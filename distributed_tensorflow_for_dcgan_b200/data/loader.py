"""``distorted_inputs`` -- the reference's input function (image_input.py:98-143) as an iterator.

Semantics kept: every file of ``data_dir`` is listed and read by *every* worker with its own shuffle
(no data sharding); records are ``image_raw`` float64 ``[64,64,3]`` cast to float32;
batches come out of a shuffle queue with ``min_after_dequeue = int(107766 * 0.1) = 10776`` and
``capacity = min_after_dequeue + 3 * batch_size`` fed by 16 threads.  The commented-out
augmentations of the reference (image_input.py:123-132) are not applied.

B200-first parts: the queue and the readers are native threads (csrc/tfrecord_reader.cc) that
dequeue straight into pinned host memory, and ``DevicePrefetcher`` keeps a ring of batches in flight
on a copy stream so the step never waits on the host round trip the reference pays
(``sess.run(images)`` then ``feed_dict``, image_train.py:153,158).
"""
from __future__ import annotations

import random
import threading
from typing import Iterator, Optional

import numpy as np
import torch

from ..config import NUM_EXAMPLES_PER_EPOCH_FOR_TRAIN
from . import tfrecord

MIN_FRACTION_OF_EXAMPLES_IN_QUEUE = 0.1  # image_input.py:134


def queue_sizes(batch_size: int = 64):
    """(min_after_dequeue, capacity) = (10776, 10968) for batch 64 (image_input.py:134-136,82-84)."""
    min_q = int(NUM_EXAMPLES_PER_EPOCH_FOR_TRAIN * MIN_FRACTION_OF_EXAMPLES_IN_QUEUE)
    return min_q, min_q + 3 * batch_size


class PythonShuffleReader:
    """Pure-Python reference implementation of the shuffle queue (oracle for the native reader)."""

    def __init__(self, files, height=64, width=64, depth=3, capacity=10968, min_after_dequeue=10776, seed=0,
                 shuffle=True, epochs=0):
        self.files, self.shape = list(files), (height, width, depth)
        self.capacity, self.min_after, self.shuffle, self.epochs = capacity, min_after_dequeue, shuffle, epochs
        self.rng = random.Random(seed)
        self.buf = []
        self._src = self._records()
        self._done = False

    def _records(self):
        epoch = 0
        while not self.epochs or epoch < self.epochs:
            order = list(self.files)
            if self.shuffle:
                self.rng.shuffle(order)
            for f in order:
                for rec in tfrecord.read_records(f):
                    yield tfrecord.decode_image_example(rec, *self.shape)
            epoch += 1

    def _fill(self):
        while not self._done and len(self.buf) < self.capacity:
            try:
                self.buf.append(next(self._src))
            except StopIteration:
                self._done = True

    def dequeue_many(self, n: int) -> np.ndarray:
        out = []
        for _ in range(n):
            self._fill()
            if not self.buf:
                break
            if not self._done and len(self.buf) <= self.min_after:
                raise RuntimeError("queue underflow")
            idx = self.rng.randrange(len(self.buf)) if self.shuffle else 0
            self.buf[idx], self.buf[-1] = self.buf[-1], self.buf[idx]
            out.append(self.buf.pop() if self.shuffle else self.buf.pop(0))
        return np.stack(out) if out else np.empty((0,) + self.shape, np.float32)


def distorted_inputs(data_dir: str, batch_size: int = 64, image_size: int = 64, depth: int = 3, seed: int = 0,
                     num_threads: int = 16, native: bool = True, min_after_dequeue: Optional[int] = None,
                     capacity: Optional[int] = None, raw_f64: bool = False):
    """Returns a reader whose ``dequeue_many(batch_size)`` yields float32 ``[B, 64, 64, 3]`` batches."""
    files = tfrecord.list_data_files(data_dir)
    if not files:
        raise ValueError("no input files in " + data_dir)
    mq, cap = queue_sizes(batch_size)
    mq = mq if min_after_dequeue is None else min_after_dequeue
    cap = cap if capacity is None else capacity
    print("Filling queue with %d CIFAR images before starting to train. "
          "This will take a few minutes." % mq)  # image_input.py:137-138 (wording kept)
    if native:
        from .native_reader import NativeShuffleReader
        return NativeShuffleReader(files, image_size, image_size, depth, cap, mq, num_threads, seed, True,
                                   raw_f64=raw_f64)
    if raw_f64:
        raise ValueError("raw float64 records need the native reader")
    return PythonShuffleReader(files, image_size, image_size, depth, cap, mq, seed, True)


class SyntheticImages:
    """Synthetic images of the named shape (benchmark default; BASELINE.json measurement rules)."""

    def __init__(self, batch_size=64, image_size=64, depth=3, seed=0, pool=8):
        g = torch.Generator().manual_seed(seed)
        self.pool = [torch.empty(batch_size, image_size, image_size, depth).uniform_(-1, 1, generator=g) for _ in range(pool)]
        self.i = 0

    def dequeue_many(self, n: int) -> np.ndarray:
        t = self.pool[self.i % len(self.pool)]
        self.i += 1
        return t[:n].numpy()


class DevicePrefetcher:
    """Device-resident prefetch ring: a background thread dequeues into pinned host buffers and issues
    the H2D copies on a side stream, ``depth`` batches ahead of the consumer."""

    def __init__(self, reader, batch_size: int, shape=(64, 64, 3), device="cuda", depth: int = 3,
                 gpu_decode: bool = False):
        """``gpu_decode``: the reader hands out the records as stored (float64, ``raw_f64`` readers); they are copied
        to the device unconverted and cast to float32 there by ``decode_f64_f32_kernel`` on the prefetch stream -- the
        reference's ``decode_raw(float64)`` + ``cast(float32)`` (image_input.py:48-51) as device work."""
        self.reader, self.B, self.device, self.depth = reader, batch_size, torch.device(device), depth
        self.cuda = self.device.type == "cuda"
        self.gpu_decode = bool(gpu_decode) and self.cuda
        if gpu_decode and not getattr(reader, "raw_f64", False):
            raise ValueError("gpu_decode needs a reader that yields raw float64 records (raw_f64=True)")
        hdt = torch.float64 if self.gpu_decode else torch.float32
        self.host = [torch.empty((batch_size,) + tuple(shape), dtype=hdt, pin_memory=self.cuda) for _ in range(depth)]
        self.raw = ([torch.empty((batch_size,) + tuple(shape), dtype=torch.float64, device=self.device)
                     for _ in range(depth)] if self.gpu_decode else None)
        self.dev = [torch.empty((batch_size,) + tuple(shape), device=self.device) for _ in range(depth)]
        self.ready = [threading.Event() for _ in range(depth)]
        self.free = [threading.Event() for _ in range(depth)]
        for e in self.free:
            e.set()
        self.events = [None] * depth
        self.stream = torch.cuda.Stream(self.device) if self.cuda else None
        self._stop = False
        self._head = 0
        self._err = None
        self.thread = threading.Thread(target=self._run, daemon=True)
        self.thread.start()

    def _run(self):
        i = 0
        try:
            while not self._stop:
                slot = i % self.depth
                self.free[slot].wait()
                if self._stop:
                    return
                self.free[slot].clear()
                h = self.host[slot]
                if hasattr(self.reader, "dequeue_into"):
                    got = self.reader.dequeue_into(h.data_ptr(), self.B)
                else:
                    arr = self.reader.dequeue_many(self.B)
                    got = arr.shape[0]
                    h[:got].copy_(torch.from_numpy(np.ascontiguousarray(arr)))
                if got < self.B:
                    self._err = StopIteration()
                    self.ready[slot].set()
                    return
                if self.cuda:
                    with torch.cuda.stream(self.stream):
                        if self.gpu_decode:
                            from ..ops import kernels as K
                            self.raw[slot].copy_(h, non_blocking=True)
                            K.decode_f64(self.raw[slot], self.dev[slot])
                        else:
                            self.dev[slot].copy_(h, non_blocking=True)
                        ev = torch.cuda.Event()
                        ev.record(self.stream)
                    self.events[slot] = ev
                else:
                    self.dev[slot].copy_(h)
                self.ready[slot].set()
                i += 1
        except Exception as e:  # surfaced to the consumer
            self._err = e
            for r in self.ready:
                r.set()

    def next(self) -> torch.Tensor:
        slot = self._head % self.depth
        self.ready[slot].wait()
        if self._err is not None:
            raise self._err
        self.ready[slot].clear()
        if self.cuda:
            torch.cuda.current_stream(self.device).wait_event(self.events[slot])
        return self.dev[slot]

    def release(self) -> None:
        """Marks the batch returned by the last ``next()`` as consumed (its slot may be overwritten)."""
        slot = self._head % self.depth
        if self.cuda:
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(self.device))
            ev.synchronize()
        self.free[slot].set()
        self._head += 1

    def close(self):
        self._stop = True
        for e in self.free:
            e.set()
