"""TensorBoard event-file writer (scalars, histograms, images) without TensorFlow.

The reference logs through ``tf.scalar_summary`` / ``tf.histogram_summary`` / ``tf.image_summary``
merged into ``summary_op`` and written by the chief with ``sv.summary_computed`` every
``save_summaries_secs`` (image_train.py:86-89,98-101,114-118,163-178; per-layer activation histogram
and "sparsity" = zero fraction, distriubted_model.py:75-80).  Files are standard
``events.out.tfevents.*`` records (TFRecord framing, ``Event`` protos) in ``checkpoint_dir``.
"""
from __future__ import annotations

import io
import os
import socket
import struct
import time
from typing import Optional

import numpy as np

from . import proto
from .crc32c import masked_crc32c


def _default_buckets():
    # TensorFlow's histogram bucket limits: +-1e-12 * 1.1^k out to 1e20, plus 0 and +-max
    pos, v = [], 1e-12
    while v < 1e20:
        pos.append(v)
        v *= 1.1
    return np.array([-x for x in reversed(pos)] + [0.0] + pos + [1.7976931348623157e308])


_BUCKETS = _default_buckets()


def _scalar_value(tag: str, value: float) -> bytes:
    return proto.f_bytes(1, proto.f_string(1, tag) + proto.f_float(2, float(value)))


def _histogram_value(tag: str, values: np.ndarray) -> bytes:
    v = np.asarray(values, dtype=np.float64).ravel()
    if v.size == 0:
        v = np.zeros(1)
    counts = np.histogram(v, bins=np.concatenate([[-1.7976931348623157e308], _BUCKETS]))[0]
    nz = np.nonzero(counts)[0]
    lo, hi = (nz[0], nz[-1] + 1) if nz.size else (0, 1)
    limits, bucket = _BUCKETS[lo:hi], counts[lo:hi].astype(np.float64)
    histo = (proto.f_double(1, float(v.min())) + proto.f_double(2, float(v.max())) + proto.f_double(3, float(v.size)) +
             proto.f_double(4, float(v.sum())) + proto.f_double(5, float((v * v).sum())) +
             proto.f_packed_doubles(6, limits.tolist()) + proto.f_packed_doubles(7, bucket.tolist()))
    return proto.f_bytes(1, proto.f_string(1, tag) + proto.f_bytes(5, histo))


def _png_bytes(img_u8: np.ndarray) -> bytes:
    from PIL import Image
    buf = io.BytesIO()
    Image.fromarray(img_u8).save(buf, format="PNG")
    return buf.getvalue()


def _image_value(tag: str, img_u8: np.ndarray) -> bytes:
    h, w, c = img_u8.shape
    image = proto.f_varint(1, h) + proto.f_varint(2, w) + proto.f_varint(3, c) + proto.f_bytes(4, _png_bytes(img_u8))
    return proto.f_bytes(1, proto.f_string(1, tag) + proto.f_bytes(4, image))


class Summary:
    """Accumulates values; ``serialize()`` yields one ``Summary`` proto (the reference's summary_str)."""

    def __init__(self):
        self._parts = []

    def scalar(self, tag: str, value: float) -> None:
        self._parts.append(_scalar_value(tag, value))

    def histogram(self, tag: str, values) -> None:
        self._parts.append(_histogram_value(tag, np.asarray(values)))

    def image(self, tag: str, images: np.ndarray, max_images: int = 3) -> None:
        """``tf.image_summary(tag, tensor)``: first ``max_images`` of ``[N,H,W,C]`` floats, each min/max
        rescaled to [0,255]; tags ``tag/image/<i>``."""
        for i, img in enumerate(np.asarray(images)[:max_images]):
            lo, hi = float(img.min()), float(img.max())
            scale = 255.0 / (hi - lo) if hi > lo else 1.0
            self._parts.append(_image_value(f"{tag}/image/{i}", ((img - lo) * scale).clip(0, 255).astype(np.uint8)))

    def serialize(self) -> bytes:
        return b"".join(self._parts)


class SummaryWriter:
    def __init__(self, logdir: str):
        os.makedirs(logdir, exist_ok=True)
        name = f"events.out.tfevents.{int(time.time())}.{socket.gethostname()}.{os.getpid()}"
        self.path = os.path.join(logdir, name)
        self._f = open(self.path, "wb")
        self._write_event(proto.f_double(1, time.time()) + proto.f_string(3, "brain.Event:2"))

    def _write_event(self, event: bytes) -> None:
        header = struct.pack("<Q", len(event))
        self._f.write(header + struct.pack("<I", masked_crc32c(header)) + event +
                      struct.pack("<I", masked_crc32c(event)))
        self._f.flush()

    def add_summary(self, summary: bytes, global_step: int) -> None:
        """``sv.summary_computed(sess, summary_str)`` (image_train.py:174)."""
        self._write_event(proto.f_double(1, time.time()) + proto.f_varint(2, int(global_step)) +
                          proto.f_bytes(5, summary))

    def close(self) -> None:
        self._f.close()


def read_events(path: str):
    """Minimal reader used by the tests: yields (step, {tag: kind}) per event."""
    from ..data.tfrecord import read_records
    for rec in read_records(path):
        step, tags = 0, {}
        for f, w, v in proto.fields(rec):
            if f == 2 and w == 0:
                step = v
            elif f == 5 and w == 2:
                for f2, w2, val in proto.fields(v):
                    if f2 != 1:
                        continue
                    tag, kind, payload = None, None, None
                    for f3, w3, v3 in proto.fields(val):
                        if f3 == 1:
                            tag = bytes(v3).decode()
                        elif f3 == 2:
                            kind, payload = "scalar", struct.unpack("<f", v3)[0]
                        elif f3 == 4:
                            kind = "image"
                        elif f3 == 5:
                            kind = "histo"
                    tags[tag] = (kind, payload)
        yield step, tags
