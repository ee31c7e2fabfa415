#!/bin/bash
# r02o (1 GPU): loss-head and segment-layer weight gradients on the side stream too; warm-cache launch list of the training step
out=gpurun_out; tag=r02o
timeout 600 python -m pytest tests/test_gpu_train.py tests/test_gpu_train_api.py tests/test_gpu_wgrad.py -m gpu -q > $out/${tag}_pytest_train.log 2>&1; tail -8 $out/${tag}_pytest_train.log | cut -c1-400
for v in 0 1 all 0 1 all; do
  if [ $v = all ]; then unset XV_TRAIN_WGRAD_STREAM; else export XV_TRAIN_WGRAD_STREAM=$v; fi
  python bench.py --workload train --steps 300 --warmup 30 --no-cpu-baseline > $out/${tag}_train_wg$v.json 2>/dev/null
  python -c "import json; d=json.load(open('$out/${tag}_train_wg$v.json')); print('wgrad stream $v: ms_per_step', d['ms_per_step'], 'e2e', d['e2e']['value'])"
done
unset XV_TRAIN_WGRAD_STREAM
XV_TRAIN_WGRAD_STREAM=0 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -c 900 --csv --log-file $out/${tag}_train_launches_warm_raw.csv \
  python bench.py --workload train --steps 5 --warmup 3 --no-cpu-baseline > $out/${tag}_ncu_train.log 2>&1
python tools/summarize_ncu.py launches $out/${tag}_train_launches_warm_raw.csv > $out/${tag}_train_launches_warm.csv; head -30 $out/${tag}_train_launches_warm.csv
