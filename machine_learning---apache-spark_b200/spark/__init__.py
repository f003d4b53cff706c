"""Spark-facing API: the names the reference imports from pyspark, re-implemented for a single B200 node.

    from b200ml.spark import SparkConf, SparkSession, TorchDistributor
    from b200ml.spark import MultilayerPerceptronClassifier, MulticlassClassificationEvaluator

``install_pyspark_shim()`` additionally registers these under the ``pyspark.*`` module paths the reference uses
(pyspark, pyspark.sql, pyspark.ml.torch.distributor, pyspark.ml.classification, pyspark.ml.evaluation,
pyspark.ml.linalg) when no real pyspark is installed, so reference-style import lines work unchanged.
"""
from __future__ import annotations

import sys
import types

from .conf import SparkConf  # noqa: F401
from .dataframe import DataFrame, Row  # noqa: F401
from .distributor import TorchDistributor  # noqa: F401
from .linalg import DenseVector, SparseVector, Vectors  # noqa: F401
from .ml import (MulticlassClassificationEvaluator, MultilayerPerceptronClassificationModel,  # noqa: F401
                 MultilayerPerceptronClassifier)
from .session import SparkContext, SparkSession  # noqa: F401


def install_pyspark_shim(force: bool = False) -> bool:
    """Register b200ml.spark under the pyspark module names.  Returns True when the shim is (now) active."""
    if not force:
        try:
            import pyspark  # noqa: F401

            return getattr(sys.modules["pyspark"], "__b200ml_shim__", False)
        except ImportError:
            pass

    def mod(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        m.__b200ml_shim__ = True
        sys.modules[name] = m
        return m

    root = mod("pyspark", SparkConf=SparkConf, SparkContext=SparkContext)
    root.sql = mod("pyspark.sql", SparkSession=SparkSession, DataFrame=DataFrame, Row=Row)
    ml = mod("pyspark.ml")
    root.ml = ml
    ml.torch = mod("pyspark.ml.torch")
    ml.torch.distributor = mod("pyspark.ml.torch.distributor", TorchDistributor=TorchDistributor)
    ml.classification = mod("pyspark.ml.classification", MultilayerPerceptronClassifier=MultilayerPerceptronClassifier,
                            MultilayerPerceptronClassificationModel=MultilayerPerceptronClassificationModel)
    ml.evaluation = mod("pyspark.ml.evaluation", MulticlassClassificationEvaluator=MulticlassClassificationEvaluator)
    ml.linalg = mod("pyspark.ml.linalg", DenseVector=DenseVector, SparseVector=SparseVector, Vectors=Vectors)
    return True
