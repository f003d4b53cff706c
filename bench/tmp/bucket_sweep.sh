#!/bin/bash
N=${1:-2}
for mb in 25 8 12; do
  echo "== bucket_mb $mb"
  B200ML_BUCKET_MB=$mb timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$((mb % 10)) bench.py --gpus $N --no-extras 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print(d['ms_per_step'], d['value'], d['e2e']['ms_per_step'], d['final_loss'])"
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 bench/timeline.py --gpus $N --out gpurun_out/timeline_n${N}b.json > gpurun_out/timeline_n${N}b.txt 2>&1
grep -v "Warning\|_warn_once" gpurun_out/timeline_n${N}b.txt | grep -i "step span\|allreduce\|embedding_bwd" | cut -c1-120
