"""On-disk formats: CRC-32C, TFRecord + tf.train.Example, tensor-bundle checkpoints, event files.

Oracles: RFC 3720 check values, tensorboard's own record reader / protobuf classes (an independent implementation
installed on this box), hand-built golden bytes, and write -> read -> write byte identity.
"""
import os
import struct

import numpy as np
import pytest

from distributed_tensorflow_ibm_mnist_b200.formats import checkpoint_state as cs
from distributed_tensorflow_ibm_mnist_b200.formats import crc32c, events, proto, tensor_bundle as tb, tfrecord


# ---- crc32c -------------------------------------------------------------------------------------------
def test_crc32c_known_vectors():
    assert crc32c.crc32c(b"123456789") == 0xE3069283
    assert crc32c.crc32c(b"") == 0
    assert crc32c.crc32c(bytes(32)) == 0x8A9136AA           # RFC 3720 B.4: 32 bytes of zeros
    assert crc32c.crc32c(bytes([0xFF] * 32)) == 0x62A8AB43  # 32 bytes of 0xFF
    assert crc32c.crc32c(bytes(range(32))) == 0x46DD794E    # ascending
    big = bytes(range(256)) * 41
    assert crc32c.crc32c(big) == crc32c._crc32c_py(big)     # native (when built) == table
    assert crc32c.crc32c(big[100:], crc32c.crc32c(big[:100])) == crc32c.crc32c(big)   # incremental
    for n in (767, 768, 769, 775, 1535, 1536, 1543, 2304, 2311, 5000):                # around the native path's 3 x 256-byte blocks
        for off in (0, 1, 7):
            assert crc32c.crc32c(big[off:off + n], 0x1234) == crc32c._crc32c_py(big[off:off + n], 0x1234), (n, off)


def test_crc_mask_roundtrip_and_tensorboard_agreement():
    for v in (0, 1, 0xDEADBEEF, 0xFFFFFFFF):
        assert crc32c.unmask(crc32c.mask(v)) == v
    tbmod = pytest.importorskip("tensorboard.compat.tensorflow_stub.pywrap_tensorflow")
    data = b"hello tfrecord" * 7
    assert crc32c.masked_crc32c(data) == tbmod.masked_crc32c(data)


# ---- protobuf wire codec ---------------------------------------------------------------------------------
def test_varint_roundtrip():
    for v in (0, 1, 127, 128, 300, 2 ** 32, 2 ** 63 - 1, -1, -12345):
        enc = proto.enc_varint(v)
        dec, pos = proto.dec_varint(enc, 0)
        assert pos == len(enc)
        assert proto.to_signed64(dec) == v
    assert proto.enc_varint(300) == b"\xac\x02"
    with pytest.raises(ValueError):
        proto.dec_varint(b"\x80", 0)


# ---- TFRecord / Example ------------------------------------------------------------------------------------
def test_tfrecord_golden_bytes(tmp_path):
    p = tmp_path / "one.tfrecord"
    with tfrecord.TFRecordWriter(str(p)) as w:
        w.write(b"abc")
    raw = p.read_bytes()
    assert raw[:8] == struct.pack("<Q", 3)
    assert raw[12:15] == b"abc"
    assert len(raw) == 8 + 4 + 3 + 4
    assert struct.unpack("<I", raw[8:12])[0] == crc32c.masked_crc32c(raw[:8])
    assert struct.unpack("<I", raw[15:19])[0] == crc32c.masked_crc32c(b"abc")


def test_tfrecord_roundtrip_and_corruption(tmp_path):
    p = str(tmp_path / "d.tfrecord")
    recs = [os.urandom(n) for n in (0, 1, 17, 2352, 70000)]
    with tfrecord.TFRecordWriter(p) as w:
        for r in recs:
            w.write(r)
    assert list(tfrecord.read_records(p)) == recs
    raw = bytearray(open(p, "rb").read())
    raw[40] ^= 0x55
    open(p, "wb").write(bytes(raw))
    with pytest.raises(tfrecord.TFRecordCorruption):
        list(tfrecord.read_records(p))
    open(p, "wb").write(bytes(raw[:-3]))
    with pytest.raises(tfrecord.TFRecordCorruption):
        list(tfrecord.read_records(p, verify=False))


def test_tfrecord_readable_by_tensorboard(tmp_path):
    tbmod = pytest.importorskip("tensorboard.compat.tensorflow_stub.pywrap_tensorflow")
    p = str(tmp_path / "x.tfrecord")
    recs = [b"first", b"second record", bytes(1000)]
    with tfrecord.TFRecordWriter(p) as w:
        for r in recs:
            w.write(r)
    rd = tbmod.PyRecordReader_New(p)
    got = []
    while True:
        try:
            rd.GetNext()
        except Exception:
            break
        got.append(rd.record())
    assert got == recs


def test_example_schema_golden_and_roundtrip():
    img = bytes(range(12))
    ex = tfrecord.mnist_example(img, 7)
    # Example{1: Features{1: [entry(image_raw -> bytes_list), entry(label -> int64_list)]}} hand-assembled
    f_img = b"\x0a" + bytes([2 + len(img)]) + b"\x0a" + bytes([len(img)]) + img          # Feature{bytes_list{value}}
    e_img = b"\x0a\x09image_raw" + b"\x12" + bytes([len(f_img)]) + f_img
    f_lab = b"\x1a\x03\x0a\x01\x07"                                                       # Feature{int64_list{packed 7}}
    e_lab = b"\x0a\x05label" + b"\x12" + bytes([len(f_lab)]) + f_lab
    feats = b"\x0a" + bytes([len(e_img)]) + e_img + b"\x0a" + bytes([len(e_lab)]) + e_lab
    assert ex == b"\x0a" + bytes([len(feats)]) + feats
    assert tfrecord.decode_example(ex) == {"image_raw": img, "label": 7}
    d = tfrecord.decode_example(tfrecord.encode_example({"a": [1, -2, 3], "b": [0.5, 1.5], "c": b"xyz", "d": 2.0}))
    assert d["a"] == [1, -2, 3] and d["b"] == [0.5, 1.5] and d["c"] == b"xyz" and d["d"] == 2.0


# ---- tensor bundle ---------------------------------------------------------------------------------------
def _sample_tensors():
    rng = np.random.default_rng(0)
    return {
        "global_step": np.array(1234, dtype=np.int64),
        "conv1/weights": rng.standard_normal((5, 5, 3, 32)).astype(np.float32),
        "conv1/biases": np.zeros(32, np.float32),
        "local3/weights": rng.standard_normal((64, 48)).astype(np.float32),
        "beta1_power": np.array(0.9, np.float32),
        "counts": np.arange(7, dtype=np.int32),
    }


def test_bundle_roundtrip_bit_exact_and_byte_identical(tmp_path):
    t = _sample_tensors()
    p1, p2 = str(tmp_path / "a" / "model.ckpt-1234"), str(tmp_path / "b" / "model.ckpt-1234")
    tb.write_bundle(p1, t)
    rd = tb.BundleReader(p1)
    assert rd.keys() == sorted(t)
    got = rd.read_all()
    for k, v in t.items():
        assert got[k].dtype == v.dtype and got[k].shape == v.shape
        assert got[k].tobytes() == v.tobytes()
    assert rd.entry("global_step").shape == () and rd.entry("global_step").dtype == tb.DT_INT64
    tb.write_bundle(p2, got)
    assert open(p1 + ".index", "rb").read() == open(p2 + ".index", "rb").read()
    assert open(tb.data_filename(p1, 0, 1), "rb").read() == open(tb.data_filename(p2, 0, 1), "rb").read()


def test_bundle_index_layout_golden(tmp_path):
    p = str(tmp_path / "g")
    x = np.array([1.0, 2.0], np.float32)
    tb.write_bundle(p, {"v": x})
    raw = open(p + ".index", "rb").read()
    assert struct.unpack("<Q", raw[-8:])[0] == 0xDB4775248B80FB57
    header = b"\x08\x01\x1a\x02\x08\x01"                       # num_shards=1, version{producer=1}
    entry = (b"\x08\x01" + b"\x12\x04\x12\x02\x08\x02" + b"\x28\x08" +
             b"\x35" + struct.pack("<I", crc32c.masked_crc32c(x.tobytes())))   # dtype, shape[2], size=8, crc
    block = (b"\x00\x00" + bytes([len(header)]) + header +
             b"\x00\x01" + bytes([len(entry)]) + b"v" + entry +
             struct.pack("<II", 0, 1))
    assert raw[:len(block)] == block
    assert raw[len(block)] == 0                                 # no compression
    assert struct.unpack("<I", raw[len(block) + 1:len(block) + 5])[0] == crc32c.mask(crc32c.crc32c(b"\x00", crc32c.crc32c(block)))
    assert open(tb.data_filename(p, 0, 1), "rb").read() == x.tobytes()
    items = tb.read_table(raw)
    assert [k for k, _ in items] == [b"", b"v"]


def test_bundle_sharded_layout(tmp_path):
    t = _sample_tensors()
    p = str(tmp_path / "model.ckpt-5")
    shard_of = {k: i % 3 for i, k in enumerate(sorted(t))}
    files = tb.write_bundle(p, t, shard_of, 3)
    assert sorted(os.path.basename(f) for f in files) == ["model.ckpt-5.data-00000-of-00003", "model.ckpt-5.data-00001-of-00003",
                                                         "model.ckpt-5.data-00002-of-00003", "model.ckpt-5.index"]
    rd = tb.BundleReader(p)
    assert rd.num_shards == 3
    for k, v in t.items():
        assert rd.entry(k).shard_id == shard_of[k]
        assert np.array_equal(rd.get(k), v)


def test_bundle_many_entries_multi_block_and_separators(tmp_path):
    # > 256 KiB of index entries forces several data blocks; keys share long prefixes (restart points, separators)
    t = {f"layer_{i:05d}/some/long/variable/name/weights": np.full(3, i, np.float32) for i in range(8000)}
    p = str(tmp_path / "big")
    tb.write_bundle(p, t)
    raw = open(p + ".index", "rb").read()
    assert len(raw) > tb.BLOCK_SIZE
    rd = tb.BundleReader(p)
    assert len(rd.keys()) == 8000
    assert rd.get("layer_07999/some/long/variable/name/weights")[0] == 7999
    assert tb._shortest_separator(b"abcd", b"abzz") == b"abd"
    assert tb._shortest_separator(b"abc", b"abd") == b"abc"
    assert tb._short_successor(b"\xff\xffa") == b"\xff\xffb"


def test_bundle_detects_corruption(tmp_path):
    p = str(tmp_path / "c")
    tb.write_bundle(p, _sample_tensors())
    data = tb.data_filename(p, 0, 1)
    raw = bytearray(open(data, "rb").read())
    raw[10] ^= 1
    open(data, "wb").write(bytes(raw))
    rd = tb.BundleReader(p)
    with pytest.raises(tb.BundleError):
        rd.read_all()
    idx = bytearray(open(p + ".index", "rb").read())
    idx[5] ^= 1
    open(p + ".index", "wb").write(bytes(idx))
    with pytest.raises(tb.BundleError):
        tb.BundleReader(p)
    with pytest.raises(tb.BundleError):
        tb.read_table(b"x" * 100)


def test_snappy_decompress():
    # literal "abcd", then copy (offset 4, length 8) -> "abcd" * 3
    comp = bytes([12]) + bytes([(4 - 1) << 2]) + b"abcd" + bytes([((8 - 4) << 2) | 1, 4])
    assert tb.snappy_decompress(comp) == b"abcd" * 3
    with pytest.raises(tb.BundleError):
        tb.snappy_decompress(bytes([5, 0x01, 0x09]))


def test_checkpoint_state_file(tmp_path):
    d = str(tmp_path)
    assert cs.read_state(d) is None and cs.latest_checkpoint(d) is None
    cs.write_state(d, cs.CheckpointState("model.ckpt-20", ["model.ckpt-10", "model.ckpt-20"]))
    txt = open(os.path.join(d, "checkpoint")).read()
    assert txt == 'model_checkpoint_path: "model.ckpt-20"\nall_model_checkpoint_paths: "model.ckpt-10"\nall_model_checkpoint_paths: "model.ckpt-20"\n'
    st = cs.read_state(d)
    assert st.model_checkpoint_path == "model.ckpt-20" and st.all_model_checkpoint_paths == ["model.ckpt-10", "model.ckpt-20"]
    assert cs.latest_checkpoint(d) is None                       # no .index file yet
    tb.write_bundle(os.path.join(d, "model.ckpt-20"), {"v": np.zeros(1, np.float32)})
    assert cs.latest_checkpoint(d) == os.path.join(d, "model.ckpt-20")


# ---- event files ------------------------------------------------------------------------------------------
def test_event_file_parses_with_tensorboard_protos(tmp_path):
    event_pb2 = pytest.importorskip("tensorboard.compat.proto.event_pb2")
    w = events.EventFileWriter(str(tmp_path))
    vals = np.concatenate([np.random.default_rng(0).standard_normal(1000) * 0.1, [0.0, -3.0, 5.5]])
    w.add_summary(events.scalar_value("train loss", 1.25) + events.histogram_value("conv1/weight", vals), 100)
    w.flush()
    w.close()
    evs = []
    for rec in tfrecord.read_records(w.path):
        e = event_pb2.Event()
        e.ParseFromString(rec)
        evs.append(e)
    assert evs[0].file_version == "brain.Event:2"
    assert evs[1].step == 100
    by_tag = {v.tag: v for v in evs[1].summary.value}
    assert abs(by_tag["train loss"].simple_value - 1.25) < 1e-7
    h = by_tag["conv1/weight"].histo
    assert h.num == len(vals) and abs(h.sum - vals.sum()) < 1e-9 and h.min == -3.0 and h.max == 5.5
    assert abs(sum(h.bucket) - len(vals)) < 1e-9 and len(h.bucket) == len(h.bucket_limit)
    assert list(h.bucket_limit) == sorted(h.bucket_limit)
    mine = events.read_events(w.path)
    assert mine[1]["step"] == 100 and abs(mine[1]["scalars"]["train loss"] - 1.25) < 1e-7
    assert mine[1]["histograms"]["conv1/weight"]["num"] == len(vals)


def test_histogram_bucket_rule():
    lim = events._BUCKET_LIMITS
    assert lim[len(lim) // 2] == 0.0 and np.all(np.diff(lim) > 0)
    d = proto.parse_dict(events.histogram_proto([0.5, 0.5, 2.0]))
    limits = struct.unpack("<%dd" % (len(d[6][0]) // 8), d[6][0])
    counts = struct.unpack("<%dd" % (len(d[7][0]) // 8), d[7][0])
    # every value lies in (previous limit, limit]; empty runs are collapsed into single zero-count entries
    assert sum(counts) == 3
    nz = [(l, c) for l, c in zip(limits, counts) if c > 0]
    assert len(nz) == 2 and nz[0][1] == 2 and nz[0][0] >= 0.5 and nz[1][0] >= 2.0
