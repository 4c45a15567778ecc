"""Real-data input paths: TFRecord training stream with a shuffle buffer, 1-epoch readers for inference.

* :class:`TFRecordBatches` — the training input of the reference: ``inputs(eval_data, batch_size)``
  (mnist_input.py:44-73): file list -> existence check (ValueError) -> shuffled file order, infinite epochs ->
  records parsed with the ``image_raw``/``label`` schema (:27-34) -> shuffle buffer with
  ``min_after_dequeue = 1000`` and ``capacity = 1000 + 3 * batch`` (:67-71) filled by background threads.
* :func:`iter_image_dir` / :func:`iter_tfrecord_examples` — the two single-epoch, unshuffled readers of
  inference.py:38-65 (``*.jpg|jpeg|png`` decoded to one channel, label -1; or the validation TFRecords).
"""
from __future__ import annotations

import glob
import os
import queue
import random
import threading

import numpy as np
import torch

from ..formats.tfrecord import decode_example, read_records

MIN_AFTER_DEQUEUE = 1000   # mnist_input.py:71


def check_files(filenames: list[str]) -> None:
    for f in filenames:
        if not os.path.exists(f):
            raise ValueError("Failed to find file: " + f)   # mnist_input.py:53-55


def parse_record(record: bytes, image_bytes: int | None = None) -> tuple[np.ndarray, int]:
    """``read_mnist``: uint8 pixels (still un-normalised) and int32 label — mnist_input.py:27-38."""
    ex = decode_example(record)
    raw = ex["image_raw"]
    img = np.frombuffer(raw, dtype=np.uint8)
    if image_bytes is not None and img.size != image_bytes:
        raise ValueError(f"record holds {img.size} image bytes, expected {image_bytes} "
                         "(28x28x3 = 2352 for reference-style RGB records, 784 for grey-scale)")
    return img, int(ex["label"])


def _native_loader_lib():
    """libdtm_b200.so if it is built (the loader is host code: no GPU needed), else None."""
    try:
        from ..ops import _lib
        if not _lib.available():
            return None
        import ctypes
        from ctypes import c_char_p, c_int, c_uint64, c_void_p, POINTER
        _lib.register("dtm_loader_create", [POINTER(c_char_p), c_int, c_int, c_int, c_int, c_uint64, c_int, c_int], c_void_p)
        _lib.register("dtm_loader_next", [c_void_p, c_void_p, c_void_p], c_int)
        _lib.register("dtm_loader_ring_start", [c_void_p, POINTER(c_void_p), c_int, c_uint64], c_int)
        _lib.register("dtm_loader_ring_acquire", [c_void_p], c_int)
        _lib.register("dtm_loader_ring_release", [c_void_p, c_int], c_int)
        _lib.register("dtm_loader_records_read", [c_void_p], c_uint64)
        _lib.register("dtm_loader_destroy", [c_void_p], None)
        return _lib.load()
    except Exception:  # noqa: BLE001 - no library, wrong ABI, ...: the Python pipeline below does the same job
        return None


class NativeRecordBatches:
    """The same stream as :class:`TFRecordBatches`, produced by the C++ pipeline (csrc/io/record_loader.cpp): reader
    threads, CRC checks, Example parsing and the shuffle queue run outside the interpreter, and a batch is written straight
    into one packed buffer (pixels | padding | labels) that ``NativeEngine.stage_packed`` ships with a single H2D copy.

    ``prefetch_batches`` >= 2 (default 4) turns on the loader's batch ring: one more C++ thread assembles whole batches
    into that many (pinned) buffers ahead of the caller, and ``next_packed`` only swaps buffers - the trainer's thread no
    longer spends the ~75 us a 512-image dequeue costs.  A returned buffer stays untouched for the next
    ``prefetch_batches - 2`` calls; a caller that copies it asynchronously calls :meth:`mark_in_flight` with the copy's
    stream, and the buffer then goes back to the ring as soon as that copy has finished (never before).
    ``prefetch_batches=0``: every call dequeues into a fresh buffer of its own."""

    def __init__(self, lib, filenames: list[str], batch_size: int, image_bytes: int, *, num_threads: int = 2, seed: int = 0,
                 shuffle: bool = True, min_after_dequeue: int = MIN_AFTER_DEQUEUE, pin: bool = False,
                 prefetch_batches: int = 4):
        import collections
        import ctypes

        self._lib = lib
        self.filenames, self.B, self.image_bytes = list(filenames), batch_size, image_bytes
        self.capacity = min_after_dequeue + 3 * batch_size
        self.shuffle = shuffle
        self.pin = pin and torch.cuda.is_available()
        nb = batch_size * image_bytes
        self.label_offset = nb + (-nb) % 16                       # NativeEngine's packed slot layout
        self.packed_bytes = self.label_offset + 4 * batch_size
        arr = (ctypes.c_char_p * len(filenames))(*[os.fsencode(f) for f in filenames])
        self._h = lib.dtm_loader_create(arr, len(filenames), batch_size, image_bytes, num_threads, seed, int(shuffle),
                                        min_after_dequeue)
        if not self._h:
            raise ValueError(lib.dtm_last_error().decode())
        self._ring: list[torch.Tensor] = []
        self._out = collections.deque()                           # [buffer index, CUDA event or None], oldest first
        if prefetch_batches >= 2:
            self._ring = [torch.empty(self.packed_bytes, dtype=torch.uint8, pin_memory=self.pin) for _ in range(prefetch_batches)]
            ptrs = (ctypes.c_void_p * prefetch_batches)(*[t.data_ptr() for t in self._ring])
            if lib.dtm_loader_ring_start(self._h, ptrs, prefetch_batches, self.label_offset) != 0:
                raise ValueError(lib.dtm_last_error().decode())

    def next_packed(self) -> torch.Tensor:
        if not self._ring:
            buf = torch.empty(self.packed_bytes, dtype=torch.uint8, pin_memory=self.pin)
            base = buf.data_ptr()
            if self._lib.dtm_loader_next(self._h, base, base + self.label_offset) != 0:
                raise ValueError(self._lib.dtm_last_error().decode())
            return buf
        # hand back what the caller is done with: buffers whose marked copy has finished, and - so that the ring's thread
        # always has a buffer to fill - the oldest ones beyond prefetch_batches - 2 (waiting for their copy if marked)
        while self._out:
            idx, ev = self._out[0]
            done = ev is not None and ev.query()
            if not done and len(self._out) < len(self._ring) - 1:
                break
            if ev is not None and not done:
                ev.synchronize()
            self._out.popleft()
            self._lib.dtm_loader_ring_release(self._h, idx)
        idx = self._lib.dtm_loader_ring_acquire(self._h)
        if idx < 0:
            raise ValueError(self._lib.dtm_last_error().decode())
        self._out.append([idx, None])
        return self._ring[idx]

    def mark_in_flight(self, stream) -> None:
        """The buffer returned last is being copied on ``stream``: keep it out of the ring until that copy is done."""
        if self._ring and self._out and torch.cuda.is_available():
            ev = torch.cuda.Event()
            ev.record(stream)
            self._out[-1][1] = ev

    def next(self) -> tuple[torch.Tensor, torch.Tensor]:
        buf = self.next_packed()
        if self._ring:
            buf = buf.clone()                     # callers of next() keep their batches; only next_packed() lends ring buffers
        return buf[:self.B * self.image_bytes].view(self.B, self.image_bytes), buf[self.label_offset:].view(torch.int32)

    @property
    def records_read(self) -> int:
        return int(self._lib.dtm_loader_records_read(self._h)) if self._h else 0

    def close(self) -> None:
        if self._h:
            self._lib.dtm_loader_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass


class TFRecordBatches:
    """Infinite stream of shuffled (uint8 [B, pixels], int32 [B]) batches read by background threads.

    ``TFRecordBatches(...)`` returns the native C++ pipeline (:class:`NativeRecordBatches`) when libdtm_b200.so is built;
    ``native=False`` (or ``DTM_NATIVE_LOADER=0``) selects this pure-Python implementation, which is also the test oracle."""

    def __new__(cls, filenames, batch_size, image_bytes, *, native: bool | None = None, **kw):
        if native is None:
            native = os.environ.get("DTM_NATIVE_LOADER", "1") != "0"
        if native and cls is TFRecordBatches:
            if not filenames:
                raise ValueError("no input files")
            check_files(list(filenames))
            lib = _native_loader_lib()
            if lib is not None:
                return NativeRecordBatches(lib, list(filenames), batch_size, image_bytes, **kw)
        return super().__new__(cls)

    def __init__(self, filenames: list[str], batch_size: int, image_bytes: int, *, num_threads: int = 2, seed: int = 0,
                 shuffle: bool = True, min_after_dequeue: int = MIN_AFTER_DEQUEUE, pin: bool = False,
                 native: bool | None = None):
        if not filenames:
            raise ValueError("no input files")
        check_files(filenames)
        self.filenames, self.B, self.image_bytes = list(filenames), batch_size, image_bytes
        self.capacity = min_after_dequeue + 3 * batch_size       # mnist_input.py:70
        self.min_after = min_after_dequeue if shuffle else 0
        self.shuffle = shuffle
        self.pin = pin and torch.cuda.is_available()
        self._rng = random.Random(seed)
        self._q: queue.Queue = queue.Queue(maxsize=self.capacity)
        self._buf: list[tuple[np.ndarray, int]] = []
        self._stop = threading.Event()
        self._err: list[BaseException] = []
        if not shuffle:
            num_threads = 1          # several parsers would interleave the file order
        self._file_seed = seed + 1
        self._read_lock = threading.Lock()
        self._stream = self._raw_records()
        self._threads = [threading.Thread(target=self._produce, args=(seed + 1 + i,), daemon=True) for i in range(num_threads)]
        for t in self._threads:
            t.start()

    def _raw_records(self):
        """ONE record stream for all producer threads, like the reference's single TFRecordReader behind its filename queue
        (mnist_input.py:57-71): every record is delivered once per epoch, whichever thread asks."""
        rng = random.Random(self._file_seed)
        while True:
            files = list(self.filenames)
            if self.shuffle:
                rng.shuffle(files)
            n = 0
            for f in files:
                for rec in read_records(f):
                    n += 1
                    yield rec
            if n == 0:
                raise ValueError("input files contain no records")

    def _produce(self, seed: int) -> None:
        try:
            while not self._stop.is_set():
                with self._read_lock:
                    rec = next(self._stream)
                item = parse_record(rec, self.image_bytes)
                while not self._stop.is_set():
                    try:
                        self._q.put(item, timeout=0.1)
                        break
                    except queue.Full:
                        continue
        except BaseException as e:  # surfaced to the consumer
            self._err.append(e)

    def _get(self):
        while True:
            if self._err:
                raise self._err[0]
            try:
                return self._q.get(timeout=0.2)
            except queue.Empty:
                continue

    def next(self) -> tuple[torch.Tensor, torch.Tensor]:
        # keep at least min_after_dequeue elements in the buffer while drawing uniformly from it (RandomShuffleQueue)
        while len(self._buf) < self.min_after + self.B:
            self._buf.append(self._get())
        images = torch.empty(self.B, self.image_bytes, dtype=torch.uint8)
        labels = torch.empty(self.B, dtype=torch.int32)
        for i in range(self.B):
            j = self._rng.randrange(len(self._buf)) if self.shuffle else 0
            img, lab = self._buf[j]
            if self.shuffle:
                self._buf[j] = self._buf[-1]
                self._buf.pop()
            else:
                self._buf.pop(0)
            images[i] = torch.from_numpy(img.copy())
            labels[i] = lab
        if self.pin:
            images, labels = images.pin_memory(), labels.pin_memory()
        return images, labels

    def close(self) -> None:
        self._stop.set()

    def __del__(self):
        self._stop.set()


# ---- single-epoch readers for inference ------------------------------------------------------------------
IMAGE_EXTS = (".jpg", ".jpeg", ".png")


def list_image_files(input_dir: str) -> list[str]:
    """``glob(input_dir + '/*.*')`` filtered to .jpg/.jpeg/.png — inference.py:39-42 (sorted for determinism)."""
    d = os.path.expanduser(input_dir)
    return sorted(f for f in glob.iglob(os.path.join(d, "*.*")) if f.endswith(IMAGE_EXTS))


def load_image_gray(path: str) -> np.ndarray:
    """``tf.image.decode_image(value, channels=1)``: uint8 [H, W]."""
    from PIL import Image

    with Image.open(path) as im:
        return np.asarray(im.convert("L"), dtype=np.uint8)


def crop_or_pad(img: np.ndarray, size: int) -> np.ndarray:
    """``tf.image.resize_image_with_crop_or_pad`` (centre crop / zero pad) to size x size — inference.py:73."""
    h, w = img.shape[:2]
    if h > size:
        t = (h - size) // 2
        img = img[t:t + size]
    if w > size:
        l = (w - size) // 2
        img = img[:, l:l + size]
    h, w = img.shape[:2]
    if h < size or w < size:
        out = np.zeros((size, size) + img.shape[2:], dtype=img.dtype)
        t, l = (size - h) // 2, (size - w) // 2
        out[t:t + h, l:l + w] = img
        img = out
    return img


def per_image_standardization(img: np.ndarray) -> np.ndarray:
    """``tf.image.per_image_standardization``: (x - mean) / max(stddev, 1/sqrt(N)) — inference.py:74."""
    x = img.astype(np.float32)
    n = x.size
    std = max(float(x.std()), 1.0 / np.sqrt(n))
    return (x - x.mean()) / std


def iter_image_dir(input_dir: str):
    """Yields (filename, uint8 [H, W] image, label=-1) — inference.py:38-47."""
    for f in list_image_files(input_dir):
        yield f, load_image_gray(f), -1


def iter_tfrecord_examples(filenames: list[str]):
    """Yields (key, uint8 flat image, label) over the files once, in order — inference.py:49-65."""
    check_files(filenames)
    for f in filenames:
        for i, rec in enumerate(read_records(f)):
            img, lab = parse_record(rec)
            yield f"{f}:{i}", img, lab
