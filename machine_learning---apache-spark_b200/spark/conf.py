"""SparkConf stand-in (reference: ``SparkConf()`` + ``spark.sparkContext.getConf().get('spark.executor.instances')``,
distributed_cnn.py:41-43).  Values come, in increasing priority, from a spark-defaults.conf file
(``$SPARK_CONF_DIR`` or ``$SPARK_HOME/conf``), ``SPARK_CONF_<key with dots as underscores>`` environment variables,
``B200ML_SPARK_CONF="k=v,k=v"``, and ``--conf k=v`` pairs on the command line (what ``spark-submit --conf`` passes).
``spark.executor.instances`` defaults to the number of visible GPUs (one rank per GPU), else 1.
"""
from __future__ import annotations

import os
import sys


def _default_instances() -> str:
    try:
        import torch

        if torch.cuda.is_available():
            return str(torch.cuda.device_count())
    except Exception:  # noqa: BLE001
        pass
    return "1"


def _ambient() -> dict:
    out: dict[str, str] = {}
    for d in (os.environ.get("SPARK_CONF_DIR"), os.path.join(os.environ.get("SPARK_HOME", ""), "conf")):
        if d and os.path.isfile(os.path.join(d, "spark-defaults.conf")):
            with open(os.path.join(d, "spark-defaults.conf")) as fh:
                for line in fh:
                    line = line.strip()
                    if line and not line.startswith("#"):
                        parts = line.split(None, 1)
                        if len(parts) == 2:
                            out[parts[0]] = parts[1].strip()
            break
    for k, v in os.environ.items():
        if k.startswith("SPARK_CONF_"):
            out[k[len("SPARK_CONF_"):].replace("_", ".")] = v
    for pair in filter(None, os.environ.get("B200ML_SPARK_CONF", "").split(",")):
        if "=" in pair:
            k, v = pair.split("=", 1)
            out[k.strip()] = v.strip()
    argv = sys.argv[1:]
    for i, a in enumerate(argv):
        if a == "--conf" and i + 1 < len(argv) and "=" in argv[i + 1]:
            k, v = argv[i + 1].split("=", 1)
            out[k] = v
        elif a.startswith("--conf=") and "=" in a[7:]:
            k, v = a[7:].split("=", 1)
            out[k] = v
    return out


class SparkConf:
    def __init__(self, loadDefaults: bool = True):
        self._conf: dict[str, str] = {}
        if loadDefaults:
            self._conf.update(_ambient())

    def set(self, key, value):
        self._conf[key] = str(value)
        return self

    def setIfMissing(self, key, value):
        self._conf.setdefault(key, str(value))
        return self

    def setAppName(self, value):
        return self.set("spark.app.name", value)

    def setMaster(self, value):
        return self.set("spark.master", value)

    def setAll(self, pairs):
        for k, v in pairs:
            self.set(k, v)
        return self

    def get(self, key, defaultValue=None):
        if key in self._conf:
            return self._conf[key]
        if key == "spark.executor.instances" and defaultValue is None:
            return _default_instances()
        return defaultValue

    def getAll(self):
        return list(self._conf.items())

    def contains(self, key):
        return key in self._conf

    def toDebugString(self):
        return "\n".join(f"{k}={v}" for k, v in self._conf.items())
