"""Row B5: fit time of the MLlib-style MultilayerPerceptronClassifier (full-batch L-BFGS) on a synthetic 784-dim,
10-class problem, one GPU worker vs one CPU worker.  The reference arm is Spark MLlib on a JVM (pyspark is not in this
image), so only this framework is timed; prints one JSON line.

    python bench/bench_mlpc.py [--rows 6000] [--iters 30] [--device cuda|cpu] [--workers 1,2,4,8]

``fit_seconds`` is the optimizer loop as timed inside rank 0 (after one untimed function evaluation), i.e. without the
worker start-up of the launcher; with several workers the rows are sharded and every function evaluation ends in one
float64 all-reduce over NVLink peer memory.
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
    ap.add_argument("--workers", default="1")
    ap.add_argument("--reproducible", action="store_true",
                    help="canonical chunks + fixed tree: bit-identical objective history for 1/2/4/8 workers")
    a = ap.parse_args()
    rng = np.random.default_rng(0)
    centers = rng.normal(size=(10, 784))
    y = rng.integers(0, 10, size=a.rows)
    X = centers[y] + 2.0 * rng.normal(size=(a.rows, 784))
    df = DataFrame({"label": y.astype(np.float64)}, {"features": X})
    layers = [784, 512, 256, 10]
    for workers in [int(v) for v in a.workers.split(",")]:
        if workers > 1 and a.device != "cpu" and (not torch.cuda.is_available() or torch.cuda.device_count() < workers):
            continue
        clf = ml.MultilayerPerceptronClassifier(layers=layers, maxIter=a.iters, blockSize=128, seed=1, numWorkers=workers,
                                                device=a.device, tol=0.0, reproducible=a.reproducible)
        if workers == 1:
            clf.fit(df)  # warm-up: library load, kernels, allocator
        model = clf.fit(df)
        acc = float((model.transform(df)._array("prediction") == y).mean())
        sm = model.summary
        print(json.dumps({"bench": "mlpc_fit", "device": a.device, "workers": workers, "rows": a.rows, "layers": layers,
                          "max_iter": a.iters, "fit_seconds": sm.fitSeconds, "function_evaluations": sm.functionEvaluations,
                          "final_objective": sm.objectiveHistory[-1], "train_accuracy": acc,
                          "reproducible": a.reproducible, "objective_history_hash": hash(tuple(sm.objectiveHistory)),
                          "weights": ml.num_weights(layers)}), flush=True)

if __name__ == "__main__":
    main()
