"""A minimal protobuf wire-format codec (varint / fixed32 / fixed64 / length-delimited).

The framework speaks four TensorFlow protos — ``tf.train.Example`` (reference mnist_input.py:29-33),
``BundleHeaderProto``/``BundleEntryProto`` (checkpoints, main.py:143) and ``Event``/``Summary`` (main.py:78) —
and hand-encodes them with this module so that no generated code or TensorFlow install is needed.
"""
from __future__ import annotations

import struct
from typing import Iterator

VARINT, FIXED64, BYTES, FIXED32 = 0, 1, 2, 5


def enc_varint(v: int) -> bytes:
    if v < 0:
        v += 1 << 64  # two's-complement, as protobuf encodes negative int64/int32
    out = bytearray()
    while True:
        b = v & 0x7F
        v >>= 7
        if v:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def dec_varint(buf, pos: int) -> tuple[int, int]:
    result = 0
    shift = 0
    while True:
        if pos >= len(buf):
            raise ValueError("truncated varint")
        b = buf[pos]
        pos += 1
        result |= (b & 0x7F) << shift
        if not b & 0x80:
            return result, pos
        shift += 7
        if shift > 70:
            raise ValueError("varint too long")


def to_signed64(v: int) -> int:
    return v - (1 << 64) if v >= 1 << 63 else v


def tag(field: int, wire: int) -> bytes:
    return enc_varint((field << 3) | wire)


def f_varint(field: int, v: int) -> bytes:
    return tag(field, VARINT) + enc_varint(v)


def f_bytes(field: int, b: bytes) -> bytes:
    return tag(field, BYTES) + enc_varint(len(b)) + b


def f_str(field: int, s: str) -> bytes:
    return f_bytes(field, s.encode("utf-8"))


def f_fixed32(field: int, v: int) -> bytes:
    return tag(field, FIXED32) + struct.pack("<I", v & 0xFFFFFFFF)


def f_float(field: int, v: float) -> bytes:
    return tag(field, FIXED32) + struct.pack("<f", v)


def f_double(field: int, v: float) -> bytes:
    return tag(field, FIXED64) + struct.pack("<d", v)


def f_packed_doubles(field: int, vals) -> bytes:
    payload = struct.pack("<%dd" % len(vals), *vals)
    return f_bytes(field, payload)


def parse(buf) -> Iterator[tuple[int, int, object]]:
    """Yield (field_number, wire_type, value); value is int for varint/fixed, memoryview-like bytes for BYTES."""
    pos = 0
    n = len(buf)
    while pos < n:
        key, pos = dec_varint(buf, pos)
        field, wire = key >> 3, key & 7
        if wire == VARINT:
            v, pos = dec_varint(buf, pos)
        elif wire == FIXED64:
            v = struct.unpack_from("<Q", buf, pos)[0]
            pos += 8
        elif wire == FIXED32:
            v = struct.unpack_from("<I", buf, pos)[0]
            pos += 4
        elif wire == BYTES:
            ln, pos = dec_varint(buf, pos)
            if pos + ln > n:
                raise ValueError("truncated length-delimited field")
            v = bytes(buf[pos:pos + ln])
            pos += ln
        else:
            raise ValueError(f"unsupported wire type {wire}")
        yield field, wire, v


def parse_dict(buf) -> dict[int, list]:
    out: dict[int, list] = {}
    for field, _, v in parse(buf):
        out.setdefault(field, []).append(v)
    return out
