"""ctypes binding (and in-tree g++ build) of the C++ input pipeline, csrc/tfrecord_reader.cc."""
from __future__ import annotations

import ctypes
import os
import subprocess
from typing import List, Optional

import numpy as np

_PKG = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_ROOT = os.path.dirname(_PKG)
_SRC = os.path.join(_ROOT, "csrc", "tfrecord_reader.cc")
_LIB_PATH = os.path.join(_PKG, "lib", "libdcgan_io.so")
_lib: Optional[ctypes.CDLL] = None


def build(force: bool = False) -> str:
    os.makedirs(os.path.dirname(_LIB_PATH), exist_ok=True)
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_SRC) > os.path.getmtime(_LIB_PATH):
        cmd = ["g++", "-O3", "-std=c++17", "-fPIC", "-shared", "-pthread", "-msse4.2", "-o", _LIB_PATH, _SRC]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("g++ failed:\n" + r.stderr)
    return _LIB_PATH


def available() -> bool:
    return os.path.exists(_LIB_PATH)


def load() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not available():
            build()
        lib = ctypes.CDLL(_LIB_PATH)
        lib.dcgan_io_crc32c.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.c_uint32]
        lib.dcgan_io_crc32c.restype = ctypes.c_uint32
        lib.dcgan_io_last_error.restype = ctypes.c_char_p
        lib.dcgan_io_reader_create.argtypes = [ctypes.POINTER(ctypes.c_char_p), ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                               ctypes.c_int, ctypes.c_long, ctypes.c_long, ctypes.c_int,
                                               ctypes.c_ulonglong, ctypes.c_int, ctypes.c_int, ctypes.c_int]
        lib.dcgan_io_reader_create.restype = ctypes.c_void_p
        lib.dcgan_io_reader_create2.argtypes = list(lib.dcgan_io_reader_create.argtypes) + [ctypes.c_int]
        lib.dcgan_io_reader_create2.restype = ctypes.c_void_p
        lib.dcgan_io_reader_dequeue.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
        lib.dcgan_io_reader_dequeue.restype = ctypes.c_int
        lib.dcgan_io_reader_size.argtypes = [ctypes.c_void_p]
        lib.dcgan_io_reader_size.restype = ctypes.c_long
        lib.dcgan_io_reader_records_read.argtypes = [ctypes.c_void_p]
        lib.dcgan_io_reader_records_read.restype = ctypes.c_longlong
        lib.dcgan_io_reader_destroy.argtypes = [ctypes.c_void_p]
        lib.dcgan_io_reader_destroy.restype = None
        _lib = lib
    return _lib


def crc32c(data: bytes, crc: int = 0) -> int:
    return load().dcgan_io_crc32c(bytes(data), len(data), crc)


class NativeShuffleReader:
    """Threads + shuffle queue in C++.  ``dequeue_many(n)`` blocks like ``tf.train.shuffle_batch``:
    until more than ``min_after_dequeue`` elements remain (image_input.py:77-84)."""

    def __init__(self, files: List[str], height=64, width=64, depth=3, capacity=10968, min_after_dequeue=10776,
                 num_threads=16, seed=0, shuffle=True, epochs=0, verify_crc=True, raw_f64=False):
        """``raw_f64``: keep the records as stored (float64, image_input.py:48 ``decode_raw(.., tf.float64)``) and
        hand them out unconverted -- the cast to float32 then runs on the GPU (``DevicePrefetcher(gpu_decode=True)``)."""
        lib = load()
        arr = (ctypes.c_char_p * len(files))(*[f.encode() for f in files])
        self._h = lib.dcgan_io_reader_create2(arr, len(files), height, width, depth, capacity, min_after_dequeue,
                                              num_threads, seed, int(shuffle), epochs, int(verify_crc), int(raw_f64))
        if not self._h:
            raise RuntimeError(lib.dcgan_io_last_error().decode())
        self.shape = (height, width, depth)
        self.raw_f64 = bool(raw_f64)
        self.dtype = np.float64 if raw_f64 else np.float32

    def dequeue_many(self, n: int, out: Optional[np.ndarray] = None) -> np.ndarray:
        if out is None:
            out = np.empty((n,) + self.shape, dtype=self.dtype)
        got = load().dcgan_io_reader_dequeue(self._h, n, out.ctypes.data_as(ctypes.c_void_p))
        if got < 0:
            raise IOError(load().dcgan_io_last_error().decode())
        return out[:got]

    def dequeue_into(self, ptr: int, n: int) -> int:
        """Dequeues straight into caller memory (e.g. a pinned torch tensor's data_ptr)."""
        got = load().dcgan_io_reader_dequeue(self._h, n, ctypes.c_void_p(ptr))
        if got < 0:
            raise IOError(load().dcgan_io_last_error().decode())
        return got

    def size(self) -> int:
        return load().dcgan_io_reader_size(self._h)

    def records_read(self) -> int:
        return load().dcgan_io_reader_records_read(self._h)

    def close(self) -> None:
        if self._h:
            load().dcgan_io_reader_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
