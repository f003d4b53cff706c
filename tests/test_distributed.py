"""Multi-rank behaviour of the DP engine: gloo/CPU here (world 2), NCCL + fused P2P kernels on >= 2 GPUs."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _launch(cases, nproc, port, env_extra=None, timeout=600):
    env = dict(os.environ)
    env.update(env_extra or {})
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "dist_worker.py"), *cases]
    r = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=timeout, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-6000:] + "\n" + "\n".join(l for l in r.stderr.splitlines() if "rank" in l or "Error" in l)[-8000:]
    return r.stdout


def test_dp_engine_gloo_world2():
    out = _launch(["dp_equals_big_batch"], 2, 29611, {"B200ML_TEST_CPU": "1", "OMP_NUM_THREADS": "2"})
    assert out.count("dp_equals_big_batch ok") == 2


def test_checkpoint_adam_gloo_world2():
    out = _launch(["checkpoint_adam"], 2, 29614, {"B200ML_TEST_CPU": "1", "OMP_NUM_THREADS": "2"})
    assert out.count("checkpoint_adam ok") == 2


@pytest.mark.gpu
@pytest.mark.multigpu
def test_fused_allreduce_two_gpus():
    out = _launch(["allreduce_raw", "allreduce_nvls", "fused_matches_nccl", "dp_equals_big_batch"], 2, 29612)
    assert out.count("allreduce_raw ok") == 2 and out.count("fused_matches_nccl ok") == 2


@pytest.mark.gpu
@pytest.mark.multigpu
def test_gemm_scatter_two_gpus():
    """Weight-gradient GEMMs mirror their outputs into the slice owners' staging buffers (b200ml_scatter) and the
    all-reduce skips its scatter phase (algo 5): same weights as NCCL all-reduce + optimizer."""
    out = _launch(["fused_matches_nccl", "dp_equals_big_batch"], 2, 29613, {"B200ML_GEMM_SCATTER": "1"})
    assert out.count("fused_matches_nccl ok") == 2 and out.count("dp_equals_big_batch ok") == 2
    assert "gemm-scatter" in out
