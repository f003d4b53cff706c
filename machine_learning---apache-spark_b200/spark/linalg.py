"""Dense / sparse vector types of pyspark.ml.linalg as far as the reference touches them (``x.toArray()``,
distributed_multilayer_perceptron.py:69)."""
from __future__ import annotations

import numpy as np


class DenseVector:
    def __init__(self, values):
        self.values = np.asarray(values, dtype=np.float64)

    def toArray(self):
        return self.values

    def __len__(self):
        return len(self.values)

    def __getitem__(self, i):
        return self.values[i]

    def __repr__(self):
        return f"DenseVector({self.values.tolist()})"


class SparseVector:
    def __init__(self, size, indices, values):
        self.size = int(size)
        self.indices = np.asarray(indices, dtype=np.int32)
        self.values = np.asarray(values, dtype=np.float64)

    def toArray(self):
        out = np.zeros(self.size, dtype=np.float64)
        out[self.indices] = self.values
        return out

    def __len__(self):
        return self.size

    def __repr__(self):
        return f"SparseVector({self.size}, {self.indices.tolist()}, {self.values.tolist()})"


class Vectors:
    @staticmethod
    def dense(*v):
        return DenseVector(v[0] if len(v) == 1 and hasattr(v[0], "__len__") else v)

    @staticmethod
    def sparse(size, indices, values):
        return SparseVector(size, indices, values)
