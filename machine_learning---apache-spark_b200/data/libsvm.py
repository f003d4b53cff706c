"""libsvm text format reader/writer (``label idx:val idx:val ...``, 1-based indices) — what
``spark.read.format("libsvm").load(...)`` parses for the reference (distributed_multilayer_perceptron.py:62-63)."""
from __future__ import annotations

import os

import numpy as np


def load_libsvm(path: str, num_features: int | None = None):
    """Returns (X float64 [n, d] dense, y float64 [n])."""
    labels, rows, maxidx = [], [], 0
    with open(path) as fh:
        for line in fh:
            line = line.split("#", 1)[0].strip()
            if not line:
                continue
            parts = line.split()
            labels.append(float(parts[0]))
            feats = []
            for tok in parts[1:]:
                i, v = tok.split(":", 1)
                i = int(i)
                feats.append((i, float(v)))
                maxidx = max(maxidx, i)
            rows.append(feats)
    d = int(num_features) if num_features else maxidx
    x = np.zeros((len(rows), d), dtype=np.float64)
    for r, feats in enumerate(rows):
        for i, v in feats:
            x[r, i - 1] = v
    return x, np.asarray(labels, dtype=np.float64)


def save_libsvm(path: str, x, y) -> None:
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    with open(path, "w") as fh:
        for row, lab in zip(np.asarray(x), np.asarray(y)):
            feats = " ".join(f"{i + 1}:{v:.6g}" for i, v in enumerate(row) if v != 0)
            fh.write(f"{lab:g} {feats}\n")


def sample_multiclass_path() -> str:
    """Path of Spark's ``data/mllib/sample_multiclass_classification_data.txt`` (150 rows × 4 features × 3 classes)
    when ``$SPARK_HOME`` has it; otherwise a synthetic stand-in of the same shape and value range (three separable
    4-d clusters scaled to [-1, 1]) is generated once under ``$B200ML_DATA`` (default ./data)."""
    home = os.environ.get("SPARK_HOME", "")
    real = os.path.join(home, "data", "mllib", "sample_multiclass_classification_data.txt")
    if home and os.path.isfile(real):
        return real
    root = os.environ.get("B200ML_DATA", os.path.join(os.getcwd(), "data"))
    path = os.path.join(root, "mllib", "sample_multiclass_classification_data.txt")
    if not os.path.isfile(path):
        rng = np.random.default_rng(150)
        centers = np.array([[-0.55, 0.25, -0.85, -0.90], [0.10, -0.35, 0.20, 0.10], [0.45, -0.20, 0.65, 0.70]])
        x = np.concatenate([c + 0.16 * rng.standard_normal((50, 4)) for c in centers]).clip(-1, 1)
        y = np.repeat(np.arange(3.0), 50)
        perm = rng.permutation(150)
        save_libsvm(path, np.round(x[perm], 6), y[perm])
    return path
