"""TensorBoard event files (``events.out.tfevents.*``): TFRecord-framed ``Event`` protos.

Replaces ``tf.summary.FileWriter`` (reference main.py:78, :133-135, :156).  Scalars use ``simple_value``; histograms
use ``HistogramProto`` with TensorFlow's default exponential bucket boundaries (1e-12 * 1.1^k, mirrored, plus 0).
"""
from __future__ import annotations

import os
import socket
import struct
import time

import numpy as np

from . import proto
from .tfrecord import TFRecordWriter


def _default_bucket_limits() -> np.ndarray:
    pos = []
    v = 1.0e-12
    while v < 1.0e20:
        pos.append(v)
        v *= 1.1
    pos.append(np.finfo(np.float64).max)
    neg = [-x for x in reversed(pos)]
    return np.array(neg + [0.0] + pos, dtype=np.float64)


_BUCKET_LIMITS = _default_bucket_limits()


def histogram_proto(values) -> bytes:
    """Serialise ``HistogramProto`` the way tensorflow::histogram::Histogram::EncodeToProto does (empty runs collapsed)."""
    v = np.asarray(values, dtype=np.float64).reshape(-1)
    if v.size == 0:
        return b""
    idx = np.searchsorted(_BUCKET_LIMITS, v, side="right")
    idx = np.minimum(idx, len(_BUCKET_LIMITS) - 1)
    counts = np.bincount(idx, minlength=len(_BUCKET_LIMITS)).astype(np.float64)
    limits, buckets = [], []
    i, n = 0, len(counts)
    while i < n:
        end, count = _BUCKET_LIMITS[i], counts[i]
        i += 1
        if count <= 0:
            while i < n and counts[i] <= 0:
                end, count = _BUCKET_LIMITS[i], counts[i]
                i += 1
        limits.append(float(end))
        buckets.append(float(count))
    out = proto.f_double(1, float(v.min())) + proto.f_double(2, float(v.max())) + proto.f_double(3, float(v.size))
    out += proto.f_double(4, float(v.sum())) + proto.f_double(5, float((v * v).sum()))
    out += proto.f_packed_doubles(6, limits) + proto.f_packed_doubles(7, buckets)
    return out


def scalar_value(tag: str, value: float) -> bytes:
    return proto.f_bytes(1, proto.f_str(1, tag) + proto.f_float(2, float(value)))


def histogram_value(tag: str, values) -> bytes:
    return proto.f_bytes(1, proto.f_str(1, tag) + proto.f_bytes(5, histogram_proto(values)))


def event(wall_time: float, step: int, *, file_version: str | None = None, summary: bytes | None = None) -> bytes:
    e = proto.f_double(1, wall_time)
    if step:
        e += proto.f_varint(2, step)
    if file_version is not None:
        e += proto.f_str(3, file_version)
    if summary is not None:
        e += proto.f_bytes(5, summary)
    return e


class EventFileWriter:
    """``tf.summary.FileWriter`` stand-in: ``add_summary(serialised Summary, step)``, ``flush()``, ``close()``."""

    def __init__(self, logdir: str, filename_suffix: str = ""):
        os.makedirs(logdir, exist_ok=True)
        name = f"events.out.tfevents.{int(time.time())}.{socket.gethostname()}{filename_suffix}"
        self.path = os.path.join(logdir, name)
        self._w = TFRecordWriter(self.path)
        self._w.write(event(time.time(), 0, file_version="brain.Event:2"))
        self._w.flush()

    def add_summary(self, summary: bytes, global_step: int) -> None:
        self._w.write(event(time.time(), int(global_step), summary=summary))

    def flush(self) -> None:
        self._w.flush()

    def close(self) -> None:
        self._w.close()


def read_events(path: str) -> list[dict]:
    """Decode an event file into dicts: {wall_time, step, file_version?, scalars{tag: v}, histograms{tag: {...}}}."""
    from .tfrecord import read_records

    out = []
    for rec in read_records(path):
        d = proto.parse_dict(rec)
        ev: dict = {"wall_time": struct.unpack("<d", struct.pack("<Q", d[1][0]))[0] if 1 in d else 0.0,
                    "step": proto.to_signed64(d[2][0]) if 2 in d else 0, "scalars": {}, "histograms": {}}
        if 3 in d:
            ev["file_version"] = d[3][0].decode()
        for summ in d.get(5, []):
            for f, _, val in proto.parse(summ):
                if f != 1:
                    continue
                vd = proto.parse_dict(val)
                tag = vd[1][0].decode()
                if 2 in vd:
                    ev["scalars"][tag] = struct.unpack("<f", struct.pack("<I", vd[2][0]))[0]
                if 5 in vd:
                    hd = proto.parse_dict(vd[5][0])
                    dbl = lambda k: struct.unpack("<d", struct.pack("<Q", hd[k][0]))[0] if k in hd else 0.0
                    unpack = lambda k: list(struct.unpack("<%dd" % (len(hd[k][0]) // 8), hd[k][0])) if k in hd else []
                    ev["histograms"][tag] = {"min": dbl(1), "max": dbl(2), "num": dbl(3), "sum": dbl(4), "sum_squares": dbl(5),
                                             "bucket_limit": unpack(6), "bucket": unpack(7)}
        out.append(ev)
    return out
