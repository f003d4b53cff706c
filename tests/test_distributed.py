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


def test_layout_mismatch_is_caught_gloo_world2():
    out = _launch(["layout_mismatch"], 2, 29615, {"B200ML_TEST_CPU": "1", "OMP_NUM_THREADS": "2"})
    assert out.count("layout_mismatch ok") == 2


def test_mlpc_gloo_world2():
    out = _launch(["mlpc"], 2, 29616, {"B200ML_TEST_CPU": "1", "OMP_NUM_THREADS": "2"})
    assert out.count("mlpc ok") == 2


def test_checkpoint_adam_gloo_world2():
    out = _launch(["checkpoint_adam"], 2, 29614, {"B200ML_TEST_CPU": "1", "OMP_NUM_THREADS": "2"})
    assert out.count("checkpoint_adam ok") == 2


def test_reference_style_ddp_api_gloo_world2():
    out = _launch(["compat_ddp"], 2, 29617, {"B200ML_TEST_CPU": "1", "OMP_NUM_THREADS": "2"})
    assert out.count("compat_ddp ok") == 2


def _ngpu():
    import torch
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


@pytest.mark.gpu
@pytest.mark.multigpu
@pytest.mark.parametrize("world", [2, 4, 8])
def test_fused_allreduce_multi_gpu(world):
    """Every all-reduce transport, the fused engine against NCCL, DP == big batch, and Adam checkpoint/resume, at 2, 4
    and 8 ranks (whatever the box has)."""
    if _ngpu() < world:
        pytest.skip(f"needs {world} GPUs")
    out = _launch(["allreduce_raw", "allreduce_nvls", "fused_matches_nccl", "dp_equals_big_batch", "checkpoint_adam",
                   "layout_mismatch", "mlpc"], world, 29612 + world, timeout=900)
    for case in ("allreduce_raw", "fused_matches_nccl", "dp_equals_big_batch", "checkpoint_adam", "layout_mismatch",
                 "mlpc"):
        assert out.count(f"{case} ok") == world, case


@pytest.mark.gpu
@pytest.mark.multigpu
@pytest.mark.parametrize("algo", [2, 5])
def test_engine_on_other_transports_two_gpus(algo):
    """The engine's default two-shot transport is the bulk-async ring; the SM-issued pull (2) and the ring with
    per-thread stores (5) must give the same training result."""
    out = _launch(["fused_matches_nccl", "dp_equals_big_batch"], 2, 29640 + algo, {"B200ML_AR_ALGO": str(algo)})
    assert out.count("fused_matches_nccl ok") == 2 and out.count("dp_equals_big_batch ok") == 2
