"""Every reference entry point exists under the same file name and runs to completion (quick mode, CPU)."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SINGLE = ["pytorch_multilayer_perceptron.py", "pytorch_cnn.py", "pytorch_lstm.py", "pytorch_machine_translator.py",
          "mllib_multilayer_perceptron_classifier.py"]
DIST = ["distributed_multilayer_perceptron.py", "distributed_cnn.py", "distributed_lstm.py"]
EXPECT = {"pytorch_multilayer_perceptron.py": "PyTorch Training Time", "pytorch_cnn.py": "Training time is",
          "pytorch_lstm.py": "Training time is", "pytorch_machine_translator.py": "Total Training time is",
          "mllib_multilayer_perceptron_classifier.py": "Test set accuracy",
          "distributed_multilayer_perceptron.py": "returned state_dict keys: ['layer_1.weight'",
          "distributed_cnn.py": "returned state_dict keys: ['block_1.0.weight'",
          "distributed_lstm.py": "returned state_dict keys: ['embedding.weight'"}


def _run(script, extra, tmp_path, device="cpu"):
    env = dict(os.environ, B200ML_QUICK="1", B200ML_DATA=str(tmp_path), B200ML_EXECUTORS="2", OMP_NUM_THREADS="2")
    if device == "cpu":
        env["B200ML_DEVICE"] = "cpu"
    else:
        env.pop("B200ML_DEVICE", None)
    r = subprocess.run([sys.executable, os.path.join(ROOT, script), *extra], capture_output=True, text=True, env=env,
                       timeout=900, cwd=str(tmp_path))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert EXPECT[script] in r.stdout, r.stdout[-2000:]


def test_all_reference_filenames_present():
    for f in SINGLE + DIST + ["transformer.py"]:
        assert os.path.isfile(os.path.join(ROOT, f)), f


@pytest.mark.parametrize("script", SINGLE)
def test_single_process_entrypoints(script, tmp_path):
    _run(script, [], tmp_path)


@pytest.mark.parametrize("script", DIST)
def test_spark_launched_entrypoints(script, tmp_path):
    _run(script, ["--conf", "spark.executor.instances=2"], tmp_path)


@pytest.mark.gpu
@pytest.mark.parametrize("script", SINGLE)
def test_single_process_entrypoints_on_gpu(script, tmp_path):
    """The same scripts on the CUDA device: every model family goes through the native kernels end to end."""
    _run(script, [], tmp_path, device="cuda")


@pytest.mark.gpu
@pytest.mark.parametrize("script", DIST)
def test_spark_launched_entrypoints_on_gpu(script, tmp_path):
    """One Spark-launched worker per GPU (two when the box has them, else one)."""
    import torch
    n = 2 if torch.cuda.device_count() >= 2 else 1
    _run(script, ["--conf", f"spark.executor.instances={n}"], tmp_path, device="cuda")
