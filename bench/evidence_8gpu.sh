#!/bin/bash
# round-end 8-GPU evidence run: headline bench (both issue orders), timeline
N=8
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N > gpurun_out/bench_8.json 2> gpurun_out/bench_8.err; tail -c 1400 gpurun_out/bench_8.json
echo "== enc_first 0"
B200ML_ENC_FIRST=0 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --no-extras 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print(d['ms_per_step'], d['value'], d['e2e']['ms_per_step'], d['final_loss'])"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 bench/timeline.py --gpus $N --out gpurun_out/timeline_n8.json > gpurun_out/timeline_n8.txt 2>&1
grep -v "Warning\|_warn_once" gpurun_out/timeline_n8.txt | grep -i "step span\|allreduce\|embedding_bwd" | cut -c1-120
