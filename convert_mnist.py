#!/usr/bin/env python
"""Dataset preparation — same name as the reference's convert_mnist.py, with the argv handling it left commented out.

    python convert_mnist.py idx2png      <mnist_idx_dir> <png_root>        # reference behaviour (convert_mnist.py:15-57)
    python convert_mnist.py png2tfrecord <png_root>/training train.tfrecord [--channels 3] [--shuffle]
    python convert_mnist.py <mnist_idx_dir> <png_root>                     # the reference's disabled two-argument usage

idx2png   parses the four MNIST IDX files and writes ``<png_root>/<training|testing>/<label>/<index>.png`` (grey-scale).
png2tfrecord replaces the IBM DLI dataset importer the reference relied on (README.md:12-14): it packs such a PNG tree
          into TFRecords with the ``image_raw`` / ``label`` schema that the trainer reads (mnist_input.py:29-33);
          ``--channels 3`` reproduces the importer's RGB records (2352 bytes, mnist_input.py:13-15), 1 gives 784-byte ones.
"""
from __future__ import annotations

import os
import struct
import sys
from array import array
from os import path

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

_FILES = {"training": ("train-images.idx3-ubyte", "train-labels.idx1-ubyte"),
          "testing": ("t10k-images.idx3-ubyte", "t10k-labels.idx1-ubyte")}


def read(dataset="training", path="."):
    """Parse one MNIST split: returns (labels: array('b'), pixels: array('B'), size, rows, cols)."""
    if dataset not in _FILES:
        raise ValueError("dataset must be 'testing' or 'training'")
    img_name, lbl_name = _FILES[dataset]
    with open(os.path.join(path, lbl_name), "rb") as f:
        magic, size = struct.unpack(">II", f.read(8))
        if magic != 2049:
            raise ValueError(f"{lbl_name}: bad magic {magic} (expected 2049)")
        lbl = array("b", f.read())
    with open(os.path.join(path, img_name), "rb") as f:
        magic, size_i, rows, cols = struct.unpack(">IIII", f.read(16))
        if magic != 2051:
            raise ValueError(f"{img_name}: bad magic {magic} (expected 2051)")
        img = array("B", f.read())
    if size_i != size or len(lbl) != size or len(img) != size * rows * cols:
        raise ValueError(f"inconsistent MNIST files: {size} labels, {size_i} images, {len(img)} pixel bytes")
    return lbl, img, size, rows, cols


def write_dataset(labels, data, size, rows, cols, output_dir, verbose=True):
    """Write every image as ``<output_dir>/<label>/<index>.png`` (8-bit grey-scale)."""
    from PIL import Image

    output_dirs = [path.join(output_dir, str(i)) for i in range(10)]
    for d in output_dirs:
        os.makedirs(d, exist_ok=True)
    stride = rows * cols
    for i, label in enumerate(labels):
        output_filename = path.join(output_dirs[label], str(i) + ".png")
        if verbose:
            print("writing " + output_filename)
        Image.frombytes("L", (cols, rows), bytes(data[i * stride:(i + 1) * stride])).save(output_filename)


def idx2png(input_path, output_path, verbose=True):
    for dataset in ["training", "testing"]:
        labels, data, size, rows, cols = read(dataset, input_path)
        write_dataset(labels, data, size, rows, cols, path.join(output_path, dataset), verbose=verbose)


def png2tfrecord(png_root, out_file, channels=1, shuffle=False, seed=0):
    """Pack ``<png_root>/<label>/*.png`` into one TFRecord file; returns the number of records."""
    import random

    import numpy as np
    from PIL import Image

    from distributed_tensorflow_ibm_mnist_b200.formats.tfrecord import TFRecordWriter, mnist_example

    items = []
    for label in range(10):
        d = path.join(png_root, str(label))
        if not path.isdir(d):
            continue
        names = [n for n in os.listdir(d) if n.lower().endswith(".png")]
        names.sort(key=lambda n: (len(n), n))
        items += [(path.join(d, n), label) for n in names]
    if not items:
        raise ValueError(f"no <label>/*.png files under {png_root}")
    if shuffle:
        random.Random(seed).shuffle(items)
    os.makedirs(path.dirname(path.abspath(out_file)), exist_ok=True)
    with TFRecordWriter(out_file) as w:
        for fn, label in items:
            with Image.open(fn) as im:
                g = np.asarray(im.convert("L"), dtype=np.uint8)
            px = np.repeat(g[..., None], channels, axis=2) if channels > 1 else g
            w.write(mnist_example(np.ascontiguousarray(px).tobytes(), label))
    return len(items)


def main(argv):
    args, channels, shuffle, quiet = [], 1, False, False
    it = iter(argv[1:])
    for a in it:
        if a.startswith("--channels"):      # --channels=3, --channels 3, or bare --channels (= 3)
            if "=" in a:
                channels = int(a.split("=", 1)[1])
            else:
                nxt = next(it, None)
                if nxt is not None and nxt.isdigit():
                    channels = int(nxt)
                else:
                    channels = 3
                    if nxt is not None:
                        args.append(nxt)
        elif a == "--shuffle":
            shuffle = True
        elif a == "--quiet":
            quiet = True
        elif a.startswith("--"):
            print(f"unknown option {a}", file=sys.stderr)
            return 2
        else:
            args.append(a)
    if channels not in (1, 3):
        print("--channels must be 1 or 3", file=sys.stderr)
        return 2
    if len(args) == 3 and args[0] == "idx2png":
        idx2png(args[1], args[2], verbose=not quiet)
    elif len(args) == 3 and args[0] == "png2tfrecord":
        n = png2tfrecord(args[1], args[2], channels=channels, shuffle=shuffle)
        print(f"wrote {n} records to {args[2]}")
    elif len(args) == 2:
        idx2png(args[0], args[1], verbose=not quiet)
    else:
        print("usage: {0} idx2png <input_path> <output_path>\n       {0} png2tfrecord <png_root> <out.tfrecord> [--channels=3] [--shuffle]\n"
              "       {0} <input_path> <output_path>".format(argv[0]), file=sys.stderr)
        return 2
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv))
