#!/bin/bash
# round-end 1-GPU evidence run: tests, headline bench, kernel tables, timelines, ncu step captures of the two small models
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_1.log 2>&1; tail -3 gpurun_out/pytest_gpu_1.log
timeout 600 python bench.py > gpurun_out/bench_1.json 2> gpurun_out/bench_1.err; tail -c 1500 gpurun_out/bench_1.json
timeout 200 python bench/kernel_table.py --model lstm > gpurun_out/ktable_lstm.txt 2>&1
timeout 200 python bench/kernel_table.py --model cnn > gpurun_out/ktable_cnn.txt 2>&1
timeout 200 python bench/kernel_table.py > gpurun_out/ktable_transformer.txt 2>&1
timeout 300 python bench/timeline.py --out gpurun_out/timeline_n1.json > gpurun_out/timeline_n1.txt 2>&1
for m in cnn lstm; do
  timeout 300 ncu --set full --clock-control none --profile-from-start off -o /tmp/step_$m python bench/ncu_step.py --model $m > gpurun_out/ncu_$m.log 2>&1
  ncu -i /tmp/step_$m.ncu-rep --page raw --csv > gpurun_out/step_${m}_raw.csv 2>/dev/null
done
ls -la gpurun_out | head -40
