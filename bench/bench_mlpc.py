"""Row B5: fit time of the MLlib-style MultilayerPerceptronClassifier (full-batch L-BFGS) on a synthetic 784-dim,
10-class problem, one GPU worker vs one CPU worker.  The reference arm is Spark MLlib on a JVM (pyspark is not in this
image), so only this framework is timed; prints one JSON line.

    python bench/bench_mlpc.py [--rows 6000] [--iters 30] [--device cuda|cpu]
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200ml.spark import ml  # noqa: E402
from b200ml.spark.dataframe import DataFrame  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=6000)
    ap.add_argument("--iters", type=int, default=30)
    ap.add_argument("--device", default="cuda" if torch.cuda.is_available() else "cpu")
    a = ap.parse_args()
    rng = np.random.default_rng(0)
    centers = rng.normal(size=(10, 784))
    y = rng.integers(0, 10, size=a.rows)
    X = centers[y] + 2.0 * rng.normal(size=(a.rows, 784))
    df = DataFrame({"label": y.astype(np.float64)}, {"features": X})
    layers = [784, 512, 256, 10]
    clf = ml.MultilayerPerceptronClassifier(layers=layers, maxIter=a.iters, blockSize=128, seed=1, numWorkers=1,
                                            device=a.device, tol=0.0)
    clf.fit(df)  # warm-up: library load, kernels, allocator
    if a.device.startswith("cuda"):
        torch.cuda.synchronize()
    t0 = time.perf_counter()
    model = clf.fit(df)
    if a.device.startswith("cuda"):
        torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    acc = float((model.transform(df)._array("prediction") == y).mean())
    res = getattr(model, "summary", None)
    print(json.dumps({"bench": "mlpc_fit", "device": a.device, "rows": a.rows, "layers": layers, "max_iter": a.iters,
                      "fit_seconds": dt, "train_accuracy": acc,
                      "weights": ml.num_weights(layers)}))


if __name__ == "__main__":
    main()
