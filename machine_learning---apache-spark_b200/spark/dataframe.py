"""A column-store mini DataFrame with the handful of operations the reference scripts use:
``spark.read.format("libsvm").load(path)``, ``randomSplit``, ``select``, ``toPandas``, ``count``, ``collect``
(distributed_multilayer_perceptron.py:62-66; mllib_multilayer_perceptron_classifier.py:22-29,45-46).

Feature columns are kept as one dense float64 matrix (``_matrix``) so estimators can move them to the GPU in one copy;
``toPandas()`` materialises SparseVector objects per row like Spark does for libsvm input.
"""
from __future__ import annotations

import numpy as np

from .linalg import DenseVector, SparseVector


class Row(dict):
    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e


class DataFrame:
    def __init__(self, columns: dict, vector_cols: dict | None = None, sparse_source: bool = False):
        """columns: name -> 1-D numpy array (scalars) ; vector_cols: name -> 2-D numpy array (one vector per row)."""
        self._cols = dict(columns)
        self._vec = dict(vector_cols or {})
        self._sparse = sparse_source
        lens = {len(v) for v in self._cols.values()} | {len(v) for v in self._vec.values()}
        assert len(lens) <= 1, "ragged columns"
        self._n = lens.pop() if lens else 0

    @property
    def columns(self):
        return list(self._cols) + list(self._vec)

    def count(self):
        return self._n

    def _take(self, idx):
        return DataFrame({k: v[idx] for k, v in self._cols.items()}, {k: v[idx] for k, v in self._vec.items()},
                         self._sparse)

    def select(self, *cols):
        cols = [c for group in cols for c in (group if isinstance(group, (list, tuple)) else [group])]
        return DataFrame({c: self._cols[c] for c in cols if c in self._cols},
                         {c: self._vec[c] for c in cols if c in self._vec}, self._sparse)

    def withColumn(self, name, values):
        values = np.asarray(values)
        cols, vec = dict(self._cols), dict(self._vec)
        (vec if values.ndim == 2 else cols)[name] = values
        return DataFrame(cols, vec, self._sparse)

    def randomSplit(self, weights, seed=None):
        """Bernoulli cell sampling like Spark: each row draws one uniform number and lands in the split whose
        cumulative-weight interval contains it (so split sizes are random, as in mllib_…classifier.py:27)."""
        w = np.asarray(weights, dtype=np.float64)
        w = w / w.sum()
        edges = np.concatenate([[0.0], np.cumsum(w)])
        rng = np.random.default_rng(seed)
        u = rng.random(self._n)
        return [self._take(np.nonzero((u >= edges[i]) & (u < edges[i + 1]))[0]) for i in range(len(w))]

    def _vector_objects(self, name):
        m = self._vec[name]
        if self._sparse:
            out = []
            for r in m:
                nz = np.nonzero(r)[0]
                out.append(SparseVector(m.shape[1], nz, r[nz]))
            return out
        return [DenseVector(r) for r in m]

    def toPandas(self):
        import pandas as pd

        data = {k: v for k, v in self._cols.items()}
        for k in self._vec:
            data[k] = self._vector_objects(k)
        return pd.DataFrame(data, columns=self.columns)

    def collect(self):
        vecs = {k: self._vector_objects(k) for k in self._vec}
        return [Row({**{k: v[i].item() if hasattr(v[i], "item") else v[i] for k, v in self._cols.items()},
                     **{k: vecs[k][i] for k in vecs}}) for i in range(self._n)]

    def show(self, n=20):
        for r in self.collect()[:n]:
            print(dict(r))

    # estimator-facing accessors
    def _matrix(self, name) -> np.ndarray:
        return self._vec[name]

    def _array(self, name) -> np.ndarray:
        return self._cols[name]


class DataFrameReader:
    def __init__(self, session):
        self._fmt = None
        self._opts: dict = {}

    def format(self, fmt):
        self._fmt = fmt
        return self

    def option(self, k, v):
        self._opts[k] = v
        return self

    def load(self, path):
        assert self._fmt == "libsvm", f"only the libsvm source is implemented (got {self._fmt!r})"
        from ..data.libsvm import load_libsvm

        x, y = load_libsvm(path, num_features=self._opts.get("numFeatures"))
        return DataFrame({"label": y}, {"features": x}, sparse_source=True)
