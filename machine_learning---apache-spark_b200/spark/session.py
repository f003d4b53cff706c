"""SparkSession stand-in: ``SparkSession.builder.appName(..).config(..).getOrCreate()``, ``spark.sparkContext.getConf()``,
``spark.read.format("libsvm").load(..)``, ``spark.stop()`` — everything the reference scripts call
(distributed_cnn.py:41-43,232; mllib_multilayer_perceptron_classifier.py:12-23).  The session only carries
configuration and the DataFrame reader: on B200 the "cluster" is the GPUs of this node and the driver's sole job is to
launch and place one rank per GPU (spark/distributor.py).  When a real pyspark is importable and ``B200ML_USE_PYSPARK=1``
is set, ``getOrCreate`` returns the real session instead.
"""
from __future__ import annotations

import os

from .conf import SparkConf
from .dataframe import DataFrame, DataFrameReader


class SparkContext:
    def __init__(self, conf: SparkConf):
        self._conf = conf
        self.appName = conf.get("spark.app.name", "b200ml")
        self.master = conf.get("spark.master", "local[*]")

    def getConf(self) -> SparkConf:
        return self._conf

    def setLogLevel(self, level):
        self._log_level = level

    @property
    def defaultParallelism(self):
        return int(self._conf.get("spark.executor.instances"))


class SparkSession:
    _active = None

    class Builder:
        def __init__(self):
            self._conf = SparkConf()

        def appName(self, name):
            self._conf.setAppName(name)
            return self

        def master(self, m):
            self._conf.setMaster(m)
            return self

        def config(self, key=None, value=None, conf: SparkConf | None = None, map=None):  # noqa: A002
            if conf is not None:
                for k, v in conf.getAll():
                    self._conf.set(k, v)
            if map:
                for k, v in map.items():
                    self._conf.set(k, v)
            if key is not None:
                self._conf.set(key, value)
            return self

        def enableHiveSupport(self):
            return self

        def getOrCreate(self):
            if os.environ.get("B200ML_USE_PYSPARK") == "1":
                try:
                    from pyspark.sql import SparkSession as Real  # type: ignore

                    b = Real.builder
                    for k, v in self._conf.getAll():
                        b = b.config(k, v)
                    return b.getOrCreate()
                except ImportError:
                    pass
            if SparkSession._active is None:
                SparkSession._active = SparkSession(self._conf)
            else:
                for k, v in self._conf.getAll():
                    SparkSession._active.sparkContext.getConf().set(k, v)
            return SparkSession._active

    class _BuilderAccessor:
        def __get__(self, obj, owner):
            return SparkSession.Builder()

    builder = _BuilderAccessor()

    def __init__(self, conf: SparkConf):
        self.sparkContext = SparkContext(conf)
        self.conf = conf
        self._stopped = False

    @property
    def read(self) -> DataFrameReader:
        return DataFrameReader(self)

    def createDataFrame(self, rows, schema=None) -> DataFrame:
        import numpy as np

        rows = list(rows)
        names = schema if schema else [f"_{i + 1}" for i in range(len(rows[0]))]
        cols, vecs = {}, {}
        for j, n in enumerate(names):
            vals = [r[j] for r in rows]
            if hasattr(vals[0], "toArray"):
                vecs[n] = np.stack([v.toArray() for v in vals])
            else:
                cols[n] = np.asarray(vals)
        return DataFrame(cols, vecs)

    def stop(self):
        self._stopped = True
        SparkSession._active = None
