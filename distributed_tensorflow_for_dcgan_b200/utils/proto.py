"""A few lines of protobuf wire format: enough to write ``tf.train.Example`` / ``Event`` /
``Summary`` messages and to walk an Example down to ``features.feature["image_raw"].bytes_list
.value[0]`` (image_input.py:42-48) without TensorFlow."""
from __future__ import annotations

import struct
from typing import Dict, Iterator, List, Tuple


def varint(n: int) -> bytes:
    if n < 0:
        n += 1 << 64
    out = bytearray()
    while True:
        b = n & 0x7F
        n >>= 7
        if n:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def key(field: int, wire: int) -> bytes:
    return varint((field << 3) | wire)


def f_bytes(field: int, data: bytes) -> bytes:
    return key(field, 2) + varint(len(data)) + data


def f_string(field: int, s: str) -> bytes:
    return f_bytes(field, s.encode("utf-8"))


def f_varint(field: int, n: int) -> bytes:
    return key(field, 0) + varint(n)


def f_double(field: int, v: float) -> bytes:
    return key(field, 1) + struct.pack("<d", v)


def f_float(field: int, v: float) -> bytes:
    return key(field, 5) + struct.pack("<f", v)


def f_packed_doubles(field: int, vals) -> bytes:
    data = struct.pack(f"<{len(vals)}d", *vals)
    return f_bytes(field, data)


def read_varint(buf, pos: int) -> Tuple[int, int]:
    shift = result = 0
    while True:
        b = buf[pos]
        pos += 1
        result |= (b & 0x7F) << shift
        if not b & 0x80:
            return result, pos
        shift += 7


def fields(buf) -> Iterator[Tuple[int, int, object]]:
    """Yields (field_number, wire_type, value) for one message; length-delimited values are memoryviews."""
    mv = memoryview(buf)
    pos, n = 0, len(mv)
    while pos < n:
        k, pos = read_varint(mv, pos)
        field, wire = k >> 3, k & 7
        if wire == 0:
            v, pos = read_varint(mv, pos)
        elif wire == 1:
            v = bytes(mv[pos:pos + 8]); pos += 8
        elif wire == 2:
            ln, pos = read_varint(mv, pos)
            v = mv[pos:pos + ln]; pos += ln
        elif wire == 5:
            v = bytes(mv[pos:pos + 4]); pos += 4
        else:
            raise ValueError(f"unsupported wire type {wire}")
        yield field, wire, v


# ---- tf.train.Example ------------------------------------------------------------------------------
def encode_example_bytes(features: Dict[str, bytes]) -> bytes:
    """Example{features=1: Features{feature=1: map<string, Feature{bytes_list=1: BytesList{value=1}}>}}"""
    entries = b""
    for name, value in features.items():
        feature = f_bytes(1, f_bytes(1, value))            # Feature.bytes_list -> BytesList.value
        entry = f_string(1, name) + f_bytes(2, feature)    # map entry: key=1, value=2
        entries += f_bytes(1, entry)
    return f_bytes(1, entries)


def parse_example_bytes(record, name: str) -> bytes:
    """Returns ``features.feature[name].bytes_list.value[0]`` of a serialized Example."""
    target = name.encode("utf-8")
    for f, w, features in fields(record):
        if f != 1 or w != 2:
            continue
        for f2, w2, entry in fields(features):
            if f2 != 1 or w2 != 2:
                continue
            k, val = None, None
            for f3, w3, v3 in fields(entry):
                if f3 == 1:
                    k = bytes(v3)
                elif f3 == 2:
                    val = v3
            if k == target and val is not None:
                for f4, w4, bl in fields(val):
                    if f4 == 1 and w4 == 2:  # bytes_list
                        for f5, w5, v5 in fields(bl):
                            if f5 == 1:
                                return bytes(v5)
    raise KeyError(f"feature {name!r} not found in Example")
