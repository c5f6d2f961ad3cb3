"""Pure-Python TFRecord reader / writer -- the oracle for the C++ reader (csrc/tfrecord_reader.cc).

Record = one serialized ``tf.train.Example`` with a single bytes feature ``image_raw`` holding
64*64*3 float64 values (98,304 bytes), which the reference decodes with ``tf.decode_raw(..,
tf.float64)``, reshapes to ``[64,64,3]`` and casts to float32 (image_input.py:40-53,118).
"""
from __future__ import annotations

import os
import struct
from typing import Iterator, List

import numpy as np

from ..utils.crc32c import masked_crc32c
from ..utils import proto

IMAGE_FEATURE = "image_raw"


class TFRecordWriter:
    def __init__(self, path: str):
        self._f = open(path, "wb")

    def write(self, record: bytes) -> None:
        header = struct.pack("<Q", len(record))
        self._f.write(header)
        self._f.write(struct.pack("<I", masked_crc32c(header)))
        self._f.write(record)
        self._f.write(struct.pack("<I", masked_crc32c(record)))

    def close(self) -> None:
        self._f.close()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


def read_records(path: str, verify_crc: bool = True) -> Iterator[bytes]:
    with open(path, "rb") as f:
        while True:
            header = f.read(8)
            if not header:
                return
            if len(header) < 8:
                raise IOError(f"{path}: truncated record header")
            (length,) = struct.unpack("<Q", header)
            (hcrc,) = struct.unpack("<I", f.read(4))
            if verify_crc and hcrc != masked_crc32c(header):
                raise IOError(f"{path}: corrupted record length")
            data = f.read(length)
            if len(data) < length:
                raise IOError(f"{path}: truncated record")
            (dcrc,) = struct.unpack("<I", f.read(4))
            if verify_crc and dcrc != masked_crc32c(data):
                raise IOError(f"{path}: corrupted record data")
            yield data


def encode_image_example(image: np.ndarray) -> bytes:
    """image: [H,W,3]; stored as raw float64 like the files the reference consumes."""
    return proto.encode_example_bytes({IMAGE_FEATURE: np.ascontiguousarray(image, dtype=np.float64).tobytes()})


def decode_image_example(record: bytes, height: int = 64, width: int = 64, depth: int = 3) -> np.ndarray:
    """``decode_raw(float64)`` -> reshape ``[h,w,c]`` -> cast float32 (image_input.py:48-53,118)."""
    raw = proto.parse_example_bytes(record, IMAGE_FEATURE)
    return np.frombuffer(raw, dtype=np.float64).reshape(height, width, depth).astype(np.float32)


def write_image_tfrecord(path: str, images: np.ndarray) -> None:
    with TFRecordWriter(path) as w:
        for img in images:
            w.write(encode_image_example(img))


def list_data_files(data_dir: str) -> List[str]:
    """Every file in ``data_dir`` (image_input.py:107), existence-checked (:111-113)."""
    names = [os.path.join(data_dir, f) for f in sorted(os.listdir(data_dir))]
    for f in names:
        if not os.path.exists(f):
            raise ValueError("Failed to find file: " + f)
    return [f for f in names if os.path.isfile(f)]
