#!/bin/bash
N=${1:-2}
for ef in auto 0; do
  echo "== N=$N enc_first $ef"
  B200ML_ENC_FIRST=$ef timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --no-extras 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print(d['ms_per_step'], d['value'], d['e2e']['ms_per_step'], d['final_loss'])"
done
B200ML_ENC_FIRST=auto timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 bench/timeline.py --gpus $N --out gpurun_out/timeline_n${N}.json > gpurun_out/timeline_n${N}.txt 2>&1
grep -v "Warning\|_warn_once" gpurun_out/timeline_n${N}.txt | grep -i "step span\|allreduce\|embedding_bwd" | cut -c1-120
timeout 200 python bench.py --no-extras 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('N=1', d['ms_per_step'], d['value'], d['e2e']['ms_per_step'], d['final_loss'])"
