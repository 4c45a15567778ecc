"""``inference_helper`` stand-in: the result writer called at the end of the reference's inference.py (:103-110).

The IBM platform module is not part of the reference repository; only the two call shapes are known:
``writeClassificationResult(path, names, probs, ground_truth=labels)`` (validation) and
``writeClassificationResult(path, names, probs, prob_thresh=t, label_file=f)`` (prediction).  The result is a JSON
document: per sample the classes whose probability reaches the threshold (mapped through the label file when given),
or, in validation mode, the prediction next to the ground truth plus the overall accuracy.
"""
from __future__ import annotations

import json
import os

import numpy as np


def _read_labels(label_file: str | None, n: int) -> list[str]:
    if label_file and os.path.exists(label_file):
        with open(label_file, "r", encoding="utf-8") as f:
            labels = [line.strip() for line in f if line.strip()]
        if len(labels) >= n:
            return labels[:n]
    return [str(i) for i in range(n)]


def writeClassificationResult(path, names, probs, ground_truth=None, prob_thresh=None, label_file=None):  # noqa: N802
    probs = np.asarray(probs, dtype=np.float32).reshape(len(names), -1) if len(names) else np.zeros((0, 10), np.float32)
    ncls = probs.shape[1] if probs.size else 10
    labels = _read_labels(label_file, ncls)
    doc: dict = {"type": "classification", "num_samples": int(len(names)), "results": []}
    if ground_truth is not None:
        correct = 0
        for name, p, gt in zip(names, probs, ground_truth):
            top = int(np.argmax(p))
            ok = int(gt) >= 0 and top == int(gt)
            correct += int(ok)
            doc["results"].append({"sampleId": str(name), "label": labels[top], "prob": float(p[top]),
                                   "groundTruth": labels[int(gt)] if 0 <= int(gt) < ncls else str(gt), "correct": bool(ok)})
        doc["accuracy"] = correct / len(names) if len(names) else None
    else:
        thresh = 0.0 if prob_thresh is None else float(prob_thresh)
        doc["prob_thresh"] = thresh
        for name, p in zip(names, probs):
            order = np.argsort(-p)
            preds = [{"label": labels[int(c)], "prob": float(p[int(c)])} for c in order if p[int(c)] >= thresh]
            doc["results"].append({"sampleId": str(name), "predictions": preds})
    d = os.path.dirname(os.path.abspath(path))
    os.makedirs(d, exist_ok=True)
    with open(path, "w", encoding="utf-8") as f:
        json.dump(doc, f, indent=1)
    return doc
