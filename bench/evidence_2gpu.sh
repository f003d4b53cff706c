#!/bin/bash
# round-end 2-GPU evidence run: distributed tests, headline bench, timeline, MLPC with two workers
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_distributed.py -m gpu -x -q > gpurun_out/pytest_dist_2.log 2>&1; tail -3 gpurun_out/pytest_dist_2.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 > gpurun_out/bench_2.json 2> gpurun_out/bench_2.err; tail -c 1200 gpurun_out/bench_2.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench/timeline.py --gpus 2 --out gpurun_out/timeline_n2.json > gpurun_out/timeline_n2.txt 2>&1
timeout 300 python bench/bench_mlpc.py --workers 1,2 > gpurun_out/mlpc_2.log 2>&1; tail -4 gpurun_out/mlpc_2.log
