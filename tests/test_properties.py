"""Property-based tests of the on-disk formats (hypothesis): whatever goes in comes back bit for bit, every reader rejects
every single-byte corruption it is meant to detect, and the native C++ record parser agrees with the Python one."""
import os
import string

import numpy as np
import pytest
from hypothesis import HealthCheck, given, settings
from hypothesis import strategies as st

from distributed_tensorflow_ibm_mnist_b200.formats import proto, tensor_bundle as tb, tfrecord

SET = dict(max_examples=40, deadline=None, suppress_health_check=[HealthCheck.function_scoped_fixture, HealthCheck.too_slow])


@settings(**SET)
@given(st.integers(min_value=0, max_value=2 ** 64 - 1))
def test_varint_roundtrip_any_u64(v):
    enc = proto.enc_varint(v)
    got, pos = proto.dec_varint(enc, 0)
    assert got == v and pos == len(enc) and len(enc) <= 10


@settings(**SET)
@given(st.lists(st.binary(min_size=0, max_size=300), min_size=0, max_size=12))
def test_tfrecord_roundtrip_any_payloads(tmp_path_factory, payloads):
    p = str(tmp_path_factory.mktemp("rec") / "f.tfrecord")
    with tfrecord.TFRecordWriter(p) as w:
        for b in payloads:
            w.write(b)
    assert list(tfrecord.read_records(p)) == payloads


@settings(**SET)
@given(st.binary(min_size=1, max_size=200), st.data())
def test_tfrecord_detects_any_flipped_byte(tmp_path_factory, payload, data):
    p = str(tmp_path_factory.mktemp("rec") / "f.tfrecord")
    with tfrecord.TFRecordWriter(p) as w:
        w.write(payload)
    raw = bytearray(open(p, "rb").read())
    i = data.draw(st.integers(min_value=0, max_value=len(raw) - 1))
    raw[i] ^= data.draw(st.integers(min_value=1, max_value=255))
    open(p, "wb").write(bytes(raw))
    with pytest.raises((tfrecord.TFRecordCorruption, ValueError, EOFError)):
        list(tfrecord.read_records(p))


_names = st.text(alphabet=string.ascii_letters + string.digits + "_/", min_size=1, max_size=40).filter(lambda s: not s.startswith("/"))
_dtypes = st.sampled_from(["<f4", "<f8", "<i4", "<i8", "u1", "<i2", "bool", "<u2", "<f2"])


@st.composite
def _tensors(draw):
    out = {}
    for name in draw(st.lists(_names, min_size=1, max_size=8, unique=True)):
        dt = np.dtype(draw(_dtypes))
        shape = tuple(draw(st.lists(st.integers(min_value=0, max_value=5), min_size=0, max_size=3)))
        n = int(np.prod(shape)) if shape else 1
        raw = draw(st.binary(min_size=n * dt.itemsize, max_size=n * dt.itemsize))
        arr = np.frombuffer(raw, dtype=dt).reshape(shape).copy()
        if dt == np.dtype("bool"):
            arr = (np.frombuffer(raw, dtype="u1").reshape(shape) & 1).astype(bool)
        out[name] = arr
    return out


@settings(**SET)
@given(_tensors(), st.integers(min_value=1, max_value=3))
def test_bundle_roundtrip_any_tensors(tmp_path_factory, tensors, shards):
    prefix = str(tmp_path_factory.mktemp("ckpt") / "model.ckpt-1")
    names = sorted(tensors)
    tb.write_bundle(prefix, tensors, {n: i % shards for i, n in enumerate(names)}, shards)
    rd = tb.BundleReader(prefix)
    assert rd.keys() == names and rd.num_shards == shards
    for n in names:
        got = rd.get(n)
        assert got.dtype == tensors[n].dtype and got.shape == tensors[n].shape and got.tobytes() == tensors[n].tobytes()


@settings(**SET)
@given(st.binary(min_size=0, max_size=3000), st.integers(min_value=-2 ** 63, max_value=2 ** 63 - 1))
def test_example_roundtrip_and_native_parser_agrees(tmp_path_factory, image, label):
    rec = tfrecord.mnist_example(image, label)
    ex = tfrecord.decode_example(rec)
    assert ex["image_raw"] == image and int(ex["label"]) == label
    # the C++ loader must extract the same bytes and label (it checks the image size, so feed it the right one)
    from distributed_tensorflow_ibm_mnist_b200.data import readers

    if not image or readers._native_loader_lib() is None:
        return
    p = str(tmp_path_factory.mktemp("rec") / "one.tfrecord")
    tfrecord.write_mnist_tfrecord(p, [image], [label])
    ds = readers.TFRecordBatches([p], 1, len(image), shuffle=False, native=True)
    x, y = ds.next()
    ds.close()
    assert x.numpy().tobytes() == image and int(y[0]) == int(np.int64(label).astype(np.int32))


@settings(**SET)
@given(st.lists(st.tuples(st.binary(min_size=24, max_size=24),
                          st.one_of(st.integers(0, 9), st.integers(0, 127), st.integers(128, 70000), st.integers(-5, -1))),
                min_size=2, max_size=40))
def test_native_loader_fast_path_agrees_with_parser_on_any_label_mix(tmp_path_factory, records):
    """Same-sized images whose labels need one, two, three or ten varint bytes, in any order: the C++ reader's remembered
    record shape applies only where the bytes outside pixels and label value are identical; everything else must take the
    general parser - each record is compared with what the Python codec decodes."""
    from distributed_tensorflow_ibm_mnist_b200.data import readers

    if readers._native_loader_lib() is None:
        return
    p = str(tmp_path_factory.mktemp("rec") / "mix.tfrecord")
    tfrecord.write_mnist_tfrecord(p, [r[0] for r in records], [r[1] for r in records])
    ds = readers.TFRecordBatches([p], len(records), 24, shuffle=False, native=True, prefetch_batches=0)
    for _ in range(2):                                   # the second pass starts with a remembered shape
        x, y = ds.next()
        for i, (img, lab) in enumerate(records):
            assert x[i].numpy().tobytes() == img and int(y[i]) == int(np.int64(lab).astype(np.int32)), (i, lab)
    ds.close()
