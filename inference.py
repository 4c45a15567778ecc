#!/usr/bin/env python
"""Batch prediction entry point — same name and flags as the reference's inference.py (:14-30).

    python inference.py --model=<checkpoint dir or prefix> --input_dir=<images> --output_dir=out --output_file=res.txt
    python inference.py --model=<ckpt> --validate --val_data=a.tfrecord,b.tfrecord --output_dir=out --output_file=res.json

Pipeline (reference inference.py:67-110): read images (``*.jpg|jpeg|png`` decoded to one channel, label -1) or the
validation TFRecords once, in order -> centre crop / pad to 28x28 -> normalise -> batches of 128 with a smaller
final batch -> model -> softmax -> result file via ``inference_helper.writeClassificationResult``.

Two reference quirks are handled explicitly:
  B1  the reference feeds 1-channel images into a 3-channel graph (a reshape error at run time); here the grey image
      is replicated to the checkpoint's channel count.
  B2  training normalises x/255-0.5 while the reference's inference uses per-image standardisation; the default here
      is the *training* normalisation, ``--per_image_standardization`` restores the reference behaviour.
"""
from __future__ import annotations

import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

from distributed_tensorflow_ibm_mnist_b200.config import flags  # noqa: E402
from distributed_tensorflow_ibm_mnist_b200.config import parameter_mgr as tf_parameter_mgr  # noqa: E402
from distributed_tensorflow_ibm_mnist_b200.utils import inference_helper  # noqa: E402

FLAGS = flags.FLAGS

flags.DEFINE_string("input_dir", "", "Directory holding the images to predict")
flags.DEFINE_string("output_dir", "", "Directory where to put the inference result")
flags.DEFINE_string("output_file", "", "File name of the inference result")
flags.DEFINE_string("model", "", "The pre-trained model: a checkpoint directory (or checkpoint prefix)")
flags.DEFINE_string("label_file", "", "Labels of the image classes (one per line)")
flags.DEFINE_float("prob_thresh", 0.5, "The prediction probability threshold to display")
flags.DEFINE_bool("validate", False, "Evaluate the model with the validation dataset")
# ---- additions ---------------------------------------------------------------------------------------------
flags.DEFINE_string("val_data", "", "Comma-separated validation TFRecord files (reference: tf_parameter_mgr.getValData())")
flags.DEFINE_string("params_file", "", "JSON hyper-parameter file")
flags.DEFINE_string("preset", "", "Model preset of the checkpoint; '' = infer from its variables")
flags.DEFINE_bool("per_image_standardization", False, "Normalise like the reference's inference.py:74 instead of like training")
flags.DEFINE_bool("use_ema", False, "Predict with the ExponentialMovingAverage shadow weights")
flags.DEFINE_string("engine", "auto", "auto | native | torch")
flags.DEFINE_integer("decode_threads", 1, "Threads that decode / crop / standardise images ahead of the model (order is kept). "
                                          "Worth it for large JPEGs only: 28x28 PNGs are interpreter bound (7.7 k images/s with "
                                          "one thread, 1.8 k with four on an 8-core box)")

BATCH_SIZE = 128   # inference.py:34
IMAGE_SIZE = 28


def _infer_spec(prefix: str):
    from distributed_tensorflow_ibm_mnist_b200.formats.tensor_bundle import BundleReader
    from distributed_tensorflow_ibm_mnist_b200.models import presets

    rd = BundleReader(prefix)
    channels = rd.entry("conv1/weights").shape[2]
    fc = rd.entry("local3/weights").shape[1]
    if FLAGS.preset:
        return presets.make_spec(FLAGS.preset, channels, fc), rd
    if "local4/weights" in rd:
        return presets.make_spec("reference", channels, fc), rd
    return presets.make_spec("wide" if fc == 8192 else "baseline", channels, fc), rd


def _prepare(img_u8: np.ndarray, channels: int, standardize: bool):
    """[H, W] or flat uint8 -> flat [28*28*channels]; uint8 (training normalisation is applied by the engine) or float32."""
    from distributed_tensorflow_ibm_mnist_b200.data import readers

    if img_u8.ndim == 1:
        c = img_u8.size // (IMAGE_SIZE * IMAGE_SIZE)
        img = img_u8.reshape(IMAGE_SIZE, IMAGE_SIZE, c)
        img = img[..., 0] if c == 1 else img
    else:
        img = img_u8
    if img.ndim == 2:
        img = readers.crop_or_pad(img, IMAGE_SIZE)
        if standardize:
            img = readers.per_image_standardization(img)
        img = np.repeat(img[..., None], channels, axis=2)          # quirk B1: replicate grey to the model's channels
    else:
        if img.shape[2] != channels:
            img = np.repeat(img.mean(axis=2, keepdims=True), channels, axis=2).astype(img.dtype)
        if standardize:
            img = readers.per_image_standardization(img)
    return np.ascontiguousarray(img).reshape(-1)


def _ordered_parallel_map(fn, items, threads: int, depth: int = 512):
    """``map(fn, items)`` with ``threads`` workers and at most ``depth`` items in flight; results come back in input order
    (file names, predictions and labels stay aligned).  What ``tf.train.batch``'s enqueue threads do for the reference's
    input graph (inference.py:76-78): PNG decoding and the numpy preprocessing release the GIL."""
    if threads <= 1:
        for it in items:
            yield fn(it)
        return
    import collections
    from concurrent.futures import ThreadPoolExecutor

    with ThreadPoolExecutor(max_workers=threads) as pool:
        pending = collections.deque()
        for it in items:
            pending.append(pool.submit(fn, it))
            if len(pending) >= depth:
                yield pending.popleft().result()
        while pending:
            yield pending.popleft().result()


def predict():
    import torch

    from distributed_tensorflow_ibm_mnist_b200.data import readers
    from distributed_tensorflow_ibm_mnist_b200.train import checkpointing

    if FLAGS.params_file:
        tf_parameter_mgr.load_file(FLAGS.params_file)
    if FLAGS.val_data:
        tf_parameter_mgr.configure(val_data=FLAGS.val_data)
    # (key, loader, label): the loader runs in the decode threads - file decoding for image directories (inference.py:38-47),
    # nothing for TFRecord examples, whose pixels are already in memory (inference.py:49-65)
    if FLAGS.validate:
        stream = ((key, (lambda im=img: im), label) for key, img, label in readers.iter_tfrecord_examples(tf_parameter_mgr.getValData()))
    else:
        stream = ((f, (lambda path=f: readers.load_image_gray(path)), -1) for f in readers.list_image_files(FLAGS.input_dir))

    prediction = np.zeros((0, 10), np.float32)
    imagenames: list[str] = []
    true_label: list[int] = []

    prefix = checkpointing.resolve_checkpoint(FLAGS.model) if FLAGS.model else None
    if prefix is None:
        # the reference silently predicts nothing without a checkpoint (inference.py:90); say so instead
        print(f"no checkpoint found under {FLAGS.model!r}; nothing predicted", file=sys.stderr)
    else:
        spec, rd = _infer_spec(prefix)
        use_cuda = torch.cuda.is_available() and FLAGS.engine != "torch"
        if use_cuda:
            from distributed_tensorflow_ibm_mnist_b200.train.native_engine import NativeEngine

            engine = NativeEngine(spec, BATCH_SIZE, use_graph=False)
        else:
            from distributed_tensorflow_ibm_mnist_b200.train.torch_engine import TorchEngine

            engine = TorchEngine(spec, BATCH_SIZE)
        checkpointing.restore_checkpoint(prefix, engine, spec, "sgd", model_only=True)
        if FLAGS.use_ema:
            ema = {v.name: torch.from_numpy(rd.get(v.name + checkpointing.EMA_SUFFIX)) for v in spec.variables
                   if v.name + checkpointing.EMA_SUFFIX in rd}
            if len(ema) == len(spec.variables):
                engine.load_params(ema)
        start = time.time()
        batch_imgs, dtype = [], (np.float32 if FLAGS.per_image_standardization else np.uint8)
        chunks = []

        def flush():
            nonlocal batch_imgs
            if not batch_imgs:
                return
            x = torch.from_numpy(np.stack(batch_imgs).astype(dtype))
            p = engine.predict(x.to(engine.device) if use_cuda else x)
            chunks.append(p.cpu().numpy().astype(np.float32))
            batch_imgs = []

        def work(item):
            key, load, label = item
            return key, _prepare(load(), spec.channels, FLAGS.per_image_standardization), label

        for key, flat, label in _ordered_parallel_map(work, stream, FLAGS.decode_threads):
            batch_imgs.append(flat)
            imagenames.append(key)
            true_label.append(label)
            if len(batch_imgs) == BATCH_SIZE:
                flush()
        flush()   # allow_smaller_final_batch=True (inference.py:78)
        if chunks:
            prediction = np.concatenate(chunks, 0)
        print("Predictions are Finished in %.2f s." % (time.time() - start))

    out_path = os.path.join(FLAGS.output_dir, FLAGS.output_file)
    names = np.array(imagenames, dtype=object)
    if FLAGS.validate:
        inference_helper.writeClassificationResult(out_path, names, prediction, ground_truth=np.array(true_label, dtype=object))
    else:
        inference_helper.writeClassificationResult(out_path, names, prediction, prob_thresh=FLAGS.prob_thresh,
                                                   label_file=FLAGS.label_file)
    return 0


def main(argv=None):
    return predict()


if __name__ == "__main__":
    flags.run(main)
