"""CRC-32C (Castagnoli) and TensorFlow's masked variant, used by the TFRecord / tfevents framing
(record layout ``[u64 len][u32 masked_crc(len)][data][u32 masked_crc(data)]``; the reference reads
such files through ``tf.TFRecordReader``, image_input.py:40-41).  Table-driven, slice-by-8."""
from __future__ import annotations

import struct

_POLY = 0x82F63B78


def _make_tables():
    t0 = []
    for i in range(256):
        c = i
        for _ in range(8):
            c = (c >> 1) ^ (_POLY if c & 1 else 0)
        t0.append(c)
    tables = [t0]
    for k in range(1, 8):
        prev = tables[k - 1]
        tables.append([(prev[i] >> 8) ^ t0[prev[i] & 0xFF] for i in range(256)])
    return tables


_T = _make_tables()


def crc32c(data: bytes, crc: int = 0) -> int:
    try:  # the native reader exports a fast version
        from ..data import native_reader
        if native_reader.available():
            return native_reader.crc32c(data, crc)
    except Exception:
        pass
    return crc32c_py(data, crc)


def crc32c_py(data: bytes, crc: int = 0) -> int:
    crc ^= 0xFFFFFFFF
    n = len(data)
    i = 0
    T0, T1, T2, T3, T4, T5, T6, T7 = _T
    mv = memoryview(data)
    while n - i >= 8:
        lo, hi = struct.unpack_from("<II", mv, i)
        lo ^= crc
        crc = (T7[lo & 0xFF] ^ T6[(lo >> 8) & 0xFF] ^ T5[(lo >> 16) & 0xFF] ^ T4[lo >> 24] ^
               T3[hi & 0xFF] ^ T2[(hi >> 8) & 0xFF] ^ T1[(hi >> 16) & 0xFF] ^ T0[hi >> 24])
        i += 8
    while i < n:
        crc = (crc >> 8) ^ T0[(crc ^ mv[i]) & 0xFF]
        i += 1
    return crc ^ 0xFFFFFFFF


def masked_crc32c(data: bytes) -> int:
    c = crc32c(data)
    return ((((c >> 15) | (c << 17)) & 0xFFFFFFFF) + 0xA282EAD8) & 0xFFFFFFFF
