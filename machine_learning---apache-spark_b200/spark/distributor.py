"""TorchDistributor stand-in (reference: ``TorchDistributor(num_processes, local_mode, use_gpu).run(train_func)``,
distributed_cnn.py:227-231) — the L0 "cluster launch / placement" layer, B200 edition.

``run(fn, *args)`` pickles ``fn`` by value (cloudpickle: module-level globals the closure captures travel with it, as
with Spark — SURVEY §9 Q5), starts ``num_processes`` worker processes on this node, one per GPU when ``use_gpu``, wires
the torch.distributed env contract (MASTER_ADDR=127.0.0.1, MASTER_PORT, WORLD_SIZE, RANK, LOCAL_RANK, OMP_NUM_THREADS)
plus CUDA device placement and NUMA-friendly CPU affinity, streams every rank's output to the driver with a rank prefix,
fails fast (first non-zero exit kills the job and surfaces that rank's traceback — barrier-stage semantics), and returns
rank 0's return value.  ``run("script.py", *args)`` launches a file the same way.

With a real Spark cluster (pyspark importable and ``B200ML_USE_PYSPARK=1``) the call is delegated to
``pyspark.ml.torch.distributor.TorchDistributor`` so multi-node placement stays Spark's job.
"""
from __future__ import annotations

import os
import pickle
import socket
import subprocess
import sys
import tempfile
import threading

_RUNNER = r"""
import os, sys, pickle, traceback
sys.path[:0] = {paths!r}
try:
    import cloudpickle
    with open({payload!r}, "rb") as fh:
        fn, args, kwargs = cloudpickle.load(fh)
    if os.environ.get("B200ML_USE_GPU") == "1":
        import torch
        if torch.cuda.is_available():
            torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")) % max(torch.cuda.device_count(), 1))
    out = fn(*args, **kwargs)
    if os.environ.get("RANK") == "0":
        with open({result!r}, "wb") as fh:
            cloudpickle.dump(out, fh)
except BaseException:
    traceback.print_exc()
    sys.exit(17)
"""


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


class TorchDistributor:
    def __init__(self, num_processes: int = 1, local_mode: bool = True, use_gpu: bool = True, **_ignored):
        self.num_processes = int(num_processes)
        self.local_mode = local_mode
        self.use_gpu = use_gpu
        if self.num_processes < 1:
            raise ValueError("num_processes has to be a positive integer")

    # ------------------------------------------------------------------------------------------------------------
    def _env_for(self, rank: int, port: int) -> dict:
        env = dict(os.environ)
        env.update({"MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port), "WORLD_SIZE": str(self.num_processes),
                    "RANK": str(rank), "LOCAL_RANK": str(rank), "LOCAL_WORLD_SIZE": str(self.num_processes),
                    "NODE_RANK": "0", "B200ML_USE_GPU": "1" if self.use_gpu else "0"})
        env.setdefault("OMP_NUM_THREADS", str(max(1, (os.cpu_count() or 1) // self.num_processes)))
        if not self.use_gpu:
            env["CUDA_VISIBLE_DEVICES"] = ""
        return env

    @staticmethod
    def _pin(rank: int, world: int):
        """Give each rank a contiguous slice of the CPUs (GPUs of one NUMA domain sit next to their cores)."""
        try:
            cpus = sorted(os.sched_getaffinity(0))
            per = max(1, len(cpus) // world)
            mine = cpus[rank * per:(rank + 1) * per] or cpus
            return lambda: os.sched_setaffinity(0, mine)
        except (AttributeError, OSError):
            return None

    def run(self, train_object, *args, **kwargs):
        if os.environ.get("B200ML_USE_PYSPARK") == "1":
            try:
                from pyspark.ml.torch.distributor import TorchDistributor as Real  # type: ignore

                return Real(self.num_processes, self.local_mode, self.use_gpu).run(train_object, *args, **kwargs)
            except ImportError:
                pass
        tmp = tempfile.mkdtemp(prefix="b200ml_dist_")
        result = os.path.join(tmp, "result.pkl")
        if isinstance(train_object, str):
            cmd_tail = [train_object, *map(str, args)]
            runner = None
        else:
            import cloudpickle

            payload = os.path.join(tmp, "payload.pkl")
            with open(payload, "wb") as fh:
                cloudpickle.dump((train_object, args, kwargs), fh)
            runner = os.path.join(tmp, "runner.py")
            paths = [p for p in sys.path if p]
            with open(runner, "w") as fh:
                fh.write(_RUNNER.format(paths=paths, payload=payload, result=result))
            cmd_tail = [runner]
        port = _free_port()
        procs, pumps, tails = [], [], {}
        lock = threading.Lock()

        def pump(rank, stream):
            tail = tails.setdefault(rank, [])
            for line in stream:
                with lock:
                    sys.stdout.write(f"[rank {rank}] {line}" if self.num_processes > 1 else line)
                    sys.stdout.flush()
                tail.append(line)
                del tail[:-60]

        for rank in range(self.num_processes):
            p = subprocess.Popen([sys.executable, "-u", *cmd_tail], env=self._env_for(rank, port),
                                 stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                                 preexec_fn=self._pin(rank, self.num_processes))
            procs.append(p)
            t = threading.Thread(target=pump, args=(rank, p.stdout), daemon=True)
            t.start()
            pumps.append(t)
        failed = None
        alive = set(range(self.num_processes))
        while alive and failed is None:
            for r in list(alive):
                try:
                    rc = procs[r].wait(timeout=0.2)
                except subprocess.TimeoutExpired:
                    continue
                alive.discard(r)
                if rc != 0:
                    failed = (r, rc)
                    break
        if failed is not None:  # fail fast: tear the job down, surface the failing rank
            for r in alive:
                procs[r].kill()
            for t in pumps:
                t.join(timeout=2)
            r, rc = failed
            raise RuntimeError(f"TorchDistributor: rank {r} exited with code {rc}; last output:\n" +
                               "".join(tails.get(r, [])[-40:]))
        for t in pumps:
            t.join(timeout=5)
        if runner is not None and os.path.exists(result):
            import cloudpickle

            with open(result, "rb") as fh:
                return cloudpickle.load(fh)
        return None
