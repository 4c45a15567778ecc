"""Synthetic MNIST-shaped data and the pinned-host, double-buffered device feeder.

Replaces the reference's queue-runner input pipeline (mnist_input.py:44-73: filename queue -> TFRecordReader ->
2 preprocessing threads -> RandomShuffleQueue) for the configurations that use synthetic data (no dataset can be
downloaded here).  Batches are raw uint8 pixels + int32 labels, exactly the payload of the TFRecord schema
(``image_raw``/``label``); the ``x/255 - 0.5`` normalisation of mnist_input.py:39 happens inside conv1's load.
"""
from __future__ import annotations

import numpy as np
import torch


class SyntheticMnist:
    """A pool of ``pool_batches`` pre-generated batches in (optionally pinned) host memory.

    ``separable=True`` draws each class from its own fixed random template plus noise, so that a model can actually
    learn (used by convergence tests); otherwise pixels and labels are i.i.d. uniform (throughput benchmarks).
    """

    def __init__(self, batch_size: int, image_pixels: int, *, pool_batches: int = 8, seed: int = 0, separable: bool = False,
                 pin: bool = True):
        self.batch_size, self.image_pixels, self.pool_batches = batch_size, image_pixels, pool_batches
        rng = np.random.default_rng(seed)
        labels = rng.integers(0, 10, size=(pool_batches, batch_size), dtype=np.int32)
        if separable:
            templates = rng.integers(0, 256, size=(10, image_pixels)).astype(np.float32)
            noise = rng.normal(0.0, 48.0, size=(pool_batches, batch_size, image_pixels)).astype(np.float32)
            images = np.clip(templates[labels] * 0.6 + 50.0 + noise, 0, 255).astype(np.uint8)
        else:
            images = rng.integers(0, 256, size=(pool_batches, batch_size, image_pixels), dtype=np.uint8)
        # one packed buffer per batch (pixels | padding to 16 bytes | int32 labels): the feeder moves a batch with a single
        # H2D copy; ``images`` / ``labels`` are views into it
        nb = batch_size * image_pixels
        pad = (-nb) % 16
        self.label_offset = nb + pad
        packed = np.zeros((pool_batches, nb + pad + 4 * batch_size), dtype=np.uint8)
        packed[:, :nb] = images.reshape(pool_batches, nb)
        packed[:, nb + pad:] = labels.view(np.uint8).reshape(pool_batches, 4 * batch_size)
        self.packed = torch.from_numpy(packed)
        if pin and torch.cuda.is_available():
            self.packed = self.packed.pin_memory()
        self.images = self.packed[:, :nb].view(pool_batches, batch_size, image_pixels)
        self.labels = self.packed[:, nb + pad:].view(torch.int32)
        self._cursor = 0

    def next(self) -> tuple[torch.Tensor, torch.Tensor]:
        i = self._cursor % self.pool_batches
        self._cursor += 1
        return self.images[i], self.labels[i]

    def next_packed(self) -> torch.Tensor:
        """The next batch as one contiguous buffer (see ``NativeEngine.stage_packed``)."""
        i = self._cursor % self.pool_batches
        self._cursor += 1
        return self.packed[i]

    def __iter__(self):
        while True:
            yield self.next()

    @property
    def bytes_per_batch(self) -> int:
        return self.batch_size * self.image_pixels + self.batch_size * 4


class DeviceFeeder:
    """Double-buffered host->device feeder for :class:`NativeEngine`.

    While the compute stream runs step *i* out of device slot ``i % 2``, a dedicated copy stream moves batch *i+1*
    from pinned host memory into the other slot; two events per slot order "copy finished -> compute may read" and
    "compute finished -> copy may overwrite".  No host synchronisation is involved.
    """

    def __init__(self, engine, source):
        self.engine, self.source = engine, source
        self.copy_stream = torch.cuda.Stream(device=engine.device)
        self.ready = [torch.cuda.Event() for _ in range(2)]
        self.free = [torch.cuda.Event() for _ in range(2)]
        self._step = 0
        self._primed = False
        self._used: set[int] = set()
        self.h2d_bytes_per_step = engine.B * engine.spec.image_pixels + engine.B * 4
        # packed sources (pixels + labels in one pinned buffer of the engine's slot layout) need one copy per batch
        self._packed = (hasattr(source, "next_packed") and hasattr(engine, "stage_packed")
                        and getattr(source, "label_offset", -1) == getattr(engine, "packed_label_offset", -2))

    def _issue_copy(self, slot: int) -> None:
        if slot in self._used:
            self.copy_stream.wait_event(self.free[slot])
        if self._packed:
            self.engine.stage_packed(slot, self.source.next_packed(), stream=self.copy_stream)
            if hasattr(self.source, "mark_in_flight"):           # ring buffers go back to the loader when the copy is done
                self.source.mark_in_flight(self.copy_stream)
        else:
            images, labels = self.source.next()
            self.engine.stage_batch(slot, images, labels, stream=self.copy_stream)
        self.ready[slot].record(self.copy_stream)

    def step(self) -> None:
        """Train on the next batch (enqueues copy of the batch after it; returns immediately)."""
        eng = self.engine
        slot = self._step % 2
        if not self._primed:
            self._issue_copy(slot)
            self._primed = True
        main = torch.cuda.current_stream()
        main.wait_event(self.ready[slot])
        eng.train_step(slot)
        self.free[slot].record(main)
        self._used.add(slot)
        self._step += 1
        self._issue_copy(self._step % 2)   # prefetch the next batch while this step computes

    def invalidate(self, slot: int) -> None:
        """The slot was overwritten behind the feeder's back (e.g. by an evaluation batch): refill it."""
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream())
        self.copy_stream.wait_event(ev)
        images, labels = self.source.next()
        self.engine.stage_batch(slot, images, labels, stream=self.copy_stream)
        self.ready[slot].record(self.copy_stream)
