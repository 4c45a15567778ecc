"""TFRecord framing and the ``tf.train.Example`` schema of the reference dataset.

Record = ``uint64 length | uint32 masked_crc(length) | data | uint32 masked_crc(data)`` (little endian).
The reference reads such files with ``tf.TFRecordReader`` + ``parse_single_example`` using the features
``image_raw`` (bytes) and ``label`` (int64) — reference mnist_input.py:27-33, inference.py:50-56.
"""
from __future__ import annotations

import os
import struct
from typing import BinaryIO, Iterable, Iterator

from . import proto
from .crc32c import masked_crc32c


class TFRecordCorruption(ValueError):
    pass


class TFRecordWriter:
    def __init__(self, path: str):
        self._f: BinaryIO = open(path, "wb")

    def write(self, record: bytes) -> None:
        header = struct.pack("<Q", len(record))
        self._f.write(header)
        self._f.write(struct.pack("<I", masked_crc32c(header)))
        self._f.write(record)
        self._f.write(struct.pack("<I", masked_crc32c(record)))

    def flush(self) -> None:
        self._f.flush()

    def close(self) -> None:
        self._f.close()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


def frame(record: bytes) -> bytes:
    header = struct.pack("<Q", len(record))
    return header + struct.pack("<I", masked_crc32c(header)) + record + struct.pack("<I", masked_crc32c(record))


def read_records(path: str, verify: bool = True) -> Iterator[bytes]:
    """Iterate over the payloads of a TFRecord file; a truncated tail or CRC mismatch raises."""
    size = os.path.getsize(path)
    with open(path, "rb") as f:
        while True:
            header = f.read(8)
            if not header:
                return
            if len(header) < 8:
                raise TFRecordCorruption(f"{path}: truncated length field")
            crc_len = f.read(4)
            if len(crc_len) < 4:
                raise TFRecordCorruption(f"{path}: truncated length crc")
            if verify and struct.unpack("<I", crc_len)[0] != masked_crc32c(header):
                raise TFRecordCorruption(f"{path}: length crc mismatch")
            (length,) = struct.unpack("<Q", header)
            if length > size - f.tell():
                raise TFRecordCorruption(f"{path}: record length {length} exceeds the file size")
            data = f.read(length)
            crc_data = f.read(4)
            if len(data) < length or len(crc_data) < 4:
                raise TFRecordCorruption(f"{path}: truncated record")
            if verify and struct.unpack("<I", crc_data)[0] != masked_crc32c(data):
                raise TFRecordCorruption(f"{path}: data crc mismatch")
            yield data


# ---------------------------------------------------------------------------------------------------
# tf.train.Example   Example{1: Features}  Features{1: map<string, Feature>}  Feature{1 bytes_list | 2 float_list | 3 int64_list}
# ---------------------------------------------------------------------------------------------------
def _feature(value) -> bytes:
    if isinstance(value, (bytes, bytearray)):
        return proto.f_bytes(1, proto.f_bytes(1, bytes(value)))
    if isinstance(value, str):
        return proto.f_bytes(1, proto.f_bytes(1, value.encode()))
    if isinstance(value, float):
        return proto.f_bytes(2, proto.f_bytes(1, struct.pack("<f", value)))
    if isinstance(value, int):
        return proto.f_bytes(3, proto.f_bytes(1, proto.enc_varint(value)))
    if isinstance(value, (list, tuple)):
        if all(isinstance(v, int) for v in value):
            return proto.f_bytes(3, proto.f_bytes(1, b"".join(proto.enc_varint(v) for v in value)))
        if all(isinstance(v, float) for v in value):
            return proto.f_bytes(2, proto.f_bytes(1, struct.pack("<%df" % len(value), *value)))
        if all(isinstance(v, (bytes, bytearray)) for v in value):
            return proto.f_bytes(1, b"".join(proto.f_bytes(1, bytes(v)) for v in value))
    raise TypeError(f"unsupported feature value {type(value)}")


def encode_example(features: dict) -> bytes:
    """Serialise a ``tf.train.Example``; map entries are emitted in sorted key order (deterministic bytes)."""
    entries = b""
    for key in sorted(features):
        entry = proto.f_str(1, key) + proto.f_bytes(2, _feature(features[key]))
        entries += proto.f_bytes(1, entry)
    return proto.f_bytes(1, entries)


def decode_example(buf: bytes) -> dict:
    """Parse a ``tf.train.Example`` into {name: bytes | [bytes] | int | [int] | float | [float]}."""
    out: dict = {}
    for field, _, features in proto.parse(buf):
        if field != 1:
            continue
        for f2, _, entry in proto.parse(features):
            if f2 != 1:
                continue
            kv = proto.parse_dict(entry)
            key = kv[1][0].decode()
            feat = kv.get(2, [b""])[0]
            vals: list = []
            for kind, _, payload in proto.parse(feat):
                if kind == 1:
                    vals = [v for f, _, v in proto.parse(payload) if f == 1]
                elif kind == 2:
                    for f, wire, v in proto.parse(payload):
                        if f != 1:
                            continue
                        if wire == proto.BYTES:
                            vals += list(struct.unpack("<%df" % (len(v) // 4), v))
                        else:
                            vals.append(struct.unpack("<f", struct.pack("<I", v))[0])
                elif kind == 3:
                    for f, wire, v in proto.parse(payload):
                        if f != 1:
                            continue
                        if wire == proto.BYTES:
                            pos = 0
                            while pos < len(v):
                                x, pos = proto.dec_varint(v, pos)
                                vals.append(proto.to_signed64(x))
                        else:
                            vals.append(proto.to_signed64(v))
            out[key] = vals[0] if len(vals) == 1 else vals
    return out


def mnist_example(image_u8: bytes, label: int) -> bytes:
    """The reference schema: ``image_raw`` = H*W*C uint8 row-major, ``label`` = int64."""
    return encode_example({"image_raw": bytes(image_u8), "label": int(label)})


def write_mnist_tfrecord(path: str, images: Iterable[bytes], labels: Iterable[int]) -> int:
    n = 0
    with TFRecordWriter(path) as w:
        for img, lab in zip(images, labels):
            w.write(mnist_example(img, lab))
            n += 1
    return n
