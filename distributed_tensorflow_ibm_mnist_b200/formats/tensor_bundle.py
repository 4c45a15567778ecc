"""TensorFlow checkpoint V2 ("tensor bundle") reader and writer in pure Python.

``tf.train.Saver`` — used implicitly by the reference through ``MonitoredTrainingSession(checkpoint_dir=...)``
(reference main.py:140-146) and explicitly by ``saver.restore`` (inference.py:86-91) — stores a checkpoint as

* ``<prefix>.index``                     an immutable LevelDB-format table: key ``""`` -> BundleHeaderProto,
                                          key ``<variable name>`` -> BundleEntryProto (dtype, shape, shard, offset,
                                          size, masked crc32c),
* ``<prefix>.data-0000k-of-0000N``       raw little-endian tensor bytes, back to back.

This module reproduces that format bit-exactly (block layout, restart points, shortest-separator index keys,
block trailers, footer), so checkpoints written here restore in TensorFlow and vice versa.  The reader also
accepts snappy-compressed blocks (TF never writes them for bundles, LevelDB tools may).
"""
from __future__ import annotations

import os
import struct
from dataclasses import dataclass
from typing import Iterable, Mapping

import numpy as np

from . import proto
from .crc32c import crc32c, mask, masked_crc32c

TABLE_MAGIC = 0xDB4775248B80FB57
FOOTER_LEN = 48
BLOCK_TRAILER_LEN = 5
BLOCK_SIZE = 262144          # tensorflow/core/lib/io/table_options.h default
RESTART_INTERVAL = 16

# tensorflow DataType enum values
DT_FLOAT, DT_DOUBLE, DT_INT32, DT_UINT8, DT_INT16, DT_INT8, DT_STRING, DT_INT64, DT_BOOL = 1, 2, 3, 4, 5, 6, 7, 9, 10
DT_BFLOAT16, DT_UINT16, DT_HALF, DT_UINT32, DT_UINT64 = 14, 17, 19, 22, 23

_NP_TO_DT = {
    np.dtype("<f4"): DT_FLOAT, np.dtype("<f8"): DT_DOUBLE, np.dtype("<i4"): DT_INT32, np.dtype("u1"): DT_UINT8,
    np.dtype("<i2"): DT_INT16, np.dtype("i1"): DT_INT8, np.dtype("<i8"): DT_INT64, np.dtype("bool"): DT_BOOL,
    np.dtype("<u2"): DT_UINT16, np.dtype("<f2"): DT_HALF, np.dtype("<u4"): DT_UINT32, np.dtype("<u8"): DT_UINT64,
}
_DT_TO_NP = {v: k for k, v in _NP_TO_DT.items()}
_DT_TO_NP[DT_BFLOAT16] = np.dtype("<u2")  # surfaced as raw uint16 bit patterns


class BundleError(ValueError):
    pass


# ---------------------------------------------------------------------------------------------------
# snappy (decompression only)
# ---------------------------------------------------------------------------------------------------
def snappy_decompress(buf: bytes) -> bytes:
    n, pos = proto.dec_varint(buf, 0)
    out = bytearray()
    ln = len(buf)
    while pos < ln:
        t = buf[pos]
        pos += 1
        kind = t & 3
        if kind == 0:  # literal
            size = t >> 2
            if size >= 60:
                nb = size - 59
                size = int.from_bytes(buf[pos:pos + nb], "little")
                pos += nb
            size += 1
            out += buf[pos:pos + size]
            pos += size
            continue
        if kind == 1:
            length = ((t >> 2) & 7) + 4
            offset = ((t >> 5) << 8) | buf[pos]
            pos += 1
        elif kind == 2:
            length = (t >> 2) + 1
            offset = int.from_bytes(buf[pos:pos + 2], "little")
            pos += 2
        else:
            length = (t >> 2) + 1
            offset = int.from_bytes(buf[pos:pos + 4], "little")
            pos += 4
        if offset == 0 or offset > len(out):
            raise BundleError("corrupt snappy stream")
        start = len(out) - offset
        for i in range(length):  # may overlap its own output
            out.append(out[start + i])
    if len(out) != n:
        raise BundleError("snappy length mismatch")
    return bytes(out)


# ---------------------------------------------------------------------------------------------------
# LevelDB-format table
# ---------------------------------------------------------------------------------------------------
class _BlockBuilder:
    def __init__(self, restart_interval: int):
        self.interval = restart_interval
        self.buf = bytearray()
        self.restarts = [0]
        self.counter = 0
        self.last_key = b""

    def empty(self) -> bool:
        return not self.buf

    def size_estimate(self) -> int:
        return len(self.buf) + 4 * len(self.restarts) + 4

    def add(self, key: bytes, value: bytes) -> None:
        shared = 0
        if self.counter < self.interval:
            lim = min(len(self.last_key), len(key))
            while shared < lim and self.last_key[shared] == key[shared]:
                shared += 1
        else:
            self.restarts.append(len(self.buf))
            self.counter = 0
        self.buf += proto.enc_varint(shared) + proto.enc_varint(len(key) - shared) + proto.enc_varint(len(value))
        self.buf += key[shared:] + value
        self.last_key = key
        self.counter += 1

    def finish(self) -> bytes:
        return bytes(self.buf) + b"".join(struct.pack("<I", r) for r in self.restarts) + struct.pack("<I", len(self.restarts))


def _shortest_separator(start: bytes, limit: bytes) -> bytes:
    n = min(len(start), len(limit))
    i = 0
    while i < n and start[i] == limit[i]:
        i += 1
    if i < n:
        b = start[i]
        if b < 0xFF and b + 1 < limit[i]:
            return start[:i] + bytes([b + 1])
    return start


def _short_successor(key: bytes) -> bytes:
    for i, b in enumerate(key):
        if b != 0xFF:
            return key[:i] + bytes([b + 1])
    return key


def _block_handle(offset: int, size: int) -> bytes:
    return proto.enc_varint(offset) + proto.enc_varint(size)


def build_table(items: Iterable[tuple[bytes, bytes]], block_size: int = BLOCK_SIZE) -> bytes:
    """Serialise sorted (key, value) pairs into a LevelDB table image (no compression, no filter block)."""
    out = bytearray()
    data = _BlockBuilder(RESTART_INTERVAL)
    index = _BlockBuilder(1)
    pending: tuple[int, int] | None = None
    last_key = b""

    def write_block(contents: bytes) -> tuple[int, int]:
        off = len(out)
        out.extend(contents)
        crc = crc32c(b"\x00", crc32c(contents))
        out.extend(b"\x00" + struct.pack("<I", mask(crc)))
        return off, len(contents)

    def flush():
        nonlocal data, pending
        if data.empty():
            return
        pending = write_block(data.finish())
        data = _BlockBuilder(RESTART_INTERVAL)

    first = True
    for key, value in items:
        if not first and key <= last_key:
            raise BundleError("table keys must be strictly increasing")
        first = False
        if pending is not None:
            index.add(_shortest_separator(last_key, key), _block_handle(*pending))
            pending = None
        data.add(key, value)
        last_key = key
        if data.size_estimate() >= block_size:
            flush()
    flush()
    meta_handle = write_block(_BlockBuilder(RESTART_INTERVAL).finish())
    if pending is not None:
        index.add(_short_successor(last_key), _block_handle(*pending))
    index_handle = write_block(index.finish())
    footer = _block_handle(*meta_handle) + _block_handle(*index_handle)
    footer += b"\x00" * (40 - len(footer)) + struct.pack("<Q", TABLE_MAGIC)
    out.extend(footer)
    return bytes(out)


def _read_block(buf: bytes, offset: int, size: int, verify: bool) -> bytes:
    if offset + size + BLOCK_TRAILER_LEN > len(buf):
        raise BundleError("block handle out of range")
    contents = buf[offset:offset + size]
    btype = buf[offset + size]
    if verify:
        (stored,) = struct.unpack_from("<I", buf, offset + size + 1)
        if stored != mask(crc32c(bytes([btype]), crc32c(contents))):
            raise BundleError("table block checksum mismatch")
    if btype == 0:
        return contents
    if btype == 1:
        return snappy_decompress(contents)
    raise BundleError(f"unknown block compression type {btype}")


def _iter_block(block: bytes):
    if len(block) < 4:
        raise BundleError("block too small")
    (num_restarts,) = struct.unpack_from("<I", block, len(block) - 4)
    limit = len(block) - 4 - 4 * num_restarts
    pos = 0
    key = b""
    while pos < limit:
        shared, pos = proto.dec_varint(block, pos)
        non_shared, pos = proto.dec_varint(block, pos)
        vlen, pos = proto.dec_varint(block, pos)
        key = key[:shared] + block[pos:pos + non_shared]
        pos += non_shared
        yield key, block[pos:pos + vlen]
        pos += vlen


def read_table(buf: bytes, verify: bool = True) -> list[tuple[bytes, bytes]]:
    if len(buf) < FOOTER_LEN:
        raise BundleError("file too short to be a table")
    footer = buf[-FOOTER_LEN:]
    if struct.unpack("<Q", footer[40:])[0] != TABLE_MAGIC:
        raise BundleError("bad table magic (not a TensorFlow checkpoint index)")
    _, p = proto.dec_varint(footer, 0)
    _, p = proto.dec_varint(footer, p)
    ioff, p = proto.dec_varint(footer, p)
    isize, p = proto.dec_varint(footer, p)
    items: list[tuple[bytes, bytes]] = []
    for _, handle in _iter_block(_read_block(buf, ioff, isize, verify)):
        off, q = proto.dec_varint(handle, 0)
        size, q = proto.dec_varint(handle, q)
        items.extend(_iter_block(_read_block(buf, off, size, verify)))
    return items


# ---------------------------------------------------------------------------------------------------
# Bundle protos
# ---------------------------------------------------------------------------------------------------
def _encode_header(num_shards: int) -> bytes:
    # BundleHeaderProto{1: num_shards, 2: endianness (LITTLE = 0, omitted), 3: VersionDef{1: producer = 1}}
    return proto.f_varint(1, num_shards) + proto.f_bytes(3, proto.f_varint(1, 1))


def _encode_shape(shape: tuple[int, ...]) -> bytes:
    return b"".join(proto.f_bytes(2, proto.f_varint(1, int(d))) for d in shape)


@dataclass
class BundleEntry:
    dtype: int
    shape: tuple[int, ...]
    shard_id: int
    offset: int
    size: int
    crc32c: int  # masked

    def encode(self) -> bytes:
        b = proto.f_varint(1, self.dtype) + proto.f_bytes(2, _encode_shape(self.shape))
        if self.shard_id:
            b += proto.f_varint(3, self.shard_id)
        if self.offset:
            b += proto.f_varint(4, self.offset)
        if self.size:
            b += proto.f_varint(5, self.size)
        if self.crc32c:
            b += proto.f_fixed32(6, self.crc32c)
        return b

    @staticmethod
    def decode(buf: bytes) -> "BundleEntry":
        d = proto.parse_dict(buf)
        shape: list[int] = []
        for sp in d.get(2, []):
            for f, _, dim in proto.parse(sp):
                if f == 2:
                    dd = proto.parse_dict(dim)
                    shape.append(proto.to_signed64(dd.get(1, [0])[0]))
        if 7 in d:
            raise BundleError("sliced (partitioned) tensors are not supported")
        return BundleEntry(d.get(1, [0])[0], tuple(shape), d.get(3, [0])[0], d.get(4, [0])[0], d.get(5, [0])[0],
                           d.get(6, [0])[0])


def data_filename(prefix: str, shard: int, num_shards: int) -> str:
    return f"{prefix}.data-{shard:05d}-of-{num_shards:05d}"


def index_filename(prefix: str) -> str:
    return prefix + ".index"


def _as_le_array(value) -> np.ndarray:
    arr = np.asarray(value)
    if arr.dtype.byteorder == ">":
        arr = arr.astype(arr.dtype.newbyteorder("<"))
    if arr.dtype not in _NP_TO_DT:
        raise BundleError(f"dtype {arr.dtype} has no TensorFlow checkpoint representation here")
    return np.ascontiguousarray(arr).reshape(arr.shape)   # ascontiguousarray promotes 0-d to 1-d


def write_bundle(prefix: str, tensors: Mapping[str, np.ndarray], shard_of: Mapping[str, int] | None = None,
                 num_shards: int = 1) -> list[str]:
    """Write ``tensors`` as checkpoint ``prefix``.

    ``shard_of`` maps variable name -> data shard (TensorFlow's default Saver writes one shard per device that
    owns variables, i.e. per parameter-server task: reference main.py:80-82 with several ``--ps_hosts``).
    Tensors are laid out in sorted-name order inside each shard, exactly as ``SaveV2`` does.
    Files are written to temporaries and renamed, so a crash never leaves a half-written checkpoint visible.
    """
    if num_shards < 1:
        raise BundleError("num_shards must be >= 1")
    os.makedirs(os.path.dirname(os.path.abspath(prefix)), exist_ok=True)
    names = sorted(tensors)
    shard_bufs = [bytearray() for _ in range(num_shards)]
    entries: list[tuple[bytes, bytes]] = [(b"", _encode_header(num_shards))]
    for name in names:
        if name == "":
            raise BundleError("empty tensor name is reserved for the bundle header")
        arr = _as_le_array(tensors[name])
        shard = (shard_of or {}).get(name, 0)
        if not 0 <= shard < num_shards:
            raise BundleError(f"shard {shard} of {name!r} outside [0, {num_shards})")
        raw = arr.tobytes()
        e = BundleEntry(_NP_TO_DT[arr.dtype], tuple(arr.shape), shard, len(shard_bufs[shard]), len(raw), masked_crc32c(raw))
        shard_bufs[shard] += raw
        entries.append((name.encode("utf-8"), e.encode()))
    written = []
    for k, buf in enumerate(shard_bufs):
        path = data_filename(prefix, k, num_shards)
        _atomic_write(path, bytes(buf))
        written.append(path)
    _atomic_write(index_filename(prefix), build_table(entries))
    written.append(index_filename(prefix))
    return written


def _atomic_write(path: str, data: bytes) -> None:
    tmp = f"{path}.tmp{os.getpid()}"
    with open(tmp, "wb") as f:
        f.write(data)
        f.flush()
        os.fsync(f.fileno())
    os.replace(tmp, path)


class BundleReader:
    """Random access to a V2 checkpoint: ``keys()``, ``entry(name)``, ``get(name) -> np.ndarray`` (crc-checked)."""

    def __init__(self, prefix: str, verify: bool = True):
        self.prefix = prefix
        self.verify = verify
        path = index_filename(prefix)
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        with open(path, "rb") as f:
            items = read_table(f.read(), verify)
        if not items or items[0][0] != b"":
            raise BundleError("checkpoint index lacks the bundle header entry")
        hdr = proto.parse_dict(items[0][1])
        self.num_shards = hdr.get(1, [0])[0]
        if hdr.get(2, [0])[0] != 0:
            raise BundleError("big-endian checkpoints are not supported")
        self.entries: dict[str, BundleEntry] = {k.decode("utf-8"): BundleEntry.decode(v) for k, v in items[1:]}
        self._shards: dict[int, np.memmap | bytes] = {}

    def keys(self) -> list[str]:
        return sorted(self.entries)

    def __contains__(self, name: str) -> bool:
        return name in self.entries

    def entry(self, name: str) -> BundleEntry:
        return self.entries[name]

    def _shard(self, k: int):
        if k not in self._shards:
            path = data_filename(self.prefix, k, self.num_shards)
            self._shards[k] = np.memmap(path, dtype=np.uint8, mode="r") if os.path.getsize(path) else b""
        return self._shards[k]

    def get(self, name: str) -> np.ndarray:
        if name not in self.entries:
            raise KeyError(f"{name!r} not in checkpoint {self.prefix} (has: {', '.join(self.keys()[:8])} ...)")
        e = self.entries[name]
        if e.dtype == DT_STRING:
            raise BundleError("string tensors are not supported")
        if e.dtype not in _DT_TO_NP:
            raise BundleError(f"unsupported checkpoint dtype enum {e.dtype}")
        raw = bytes(self._shard(e.shard_id)[e.offset:e.offset + e.size])
        if len(raw) != e.size:
            raise BundleError(f"{name}: data shard truncated")
        if self.verify and masked_crc32c(raw) != e.crc32c:
            raise BundleError(f"{name}: tensor checksum mismatch")
        dt = _DT_TO_NP[e.dtype]
        count = int(np.prod(e.shape)) if e.shape else 1
        if count * dt.itemsize != e.size:
            raise BundleError(f"{name}: size {e.size} does not match shape {e.shape} of {dt}")
        return np.frombuffer(raw, dtype=dt).reshape(e.shape).copy()

    def read_all(self) -> dict[str, np.ndarray]:
        return {k: self.get(k) for k in self.keys()}


def checkpoint_exists(prefix: str) -> bool:
    return os.path.exists(index_filename(prefix))
