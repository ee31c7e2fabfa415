#!/bin/bash
# r02s (1 GPU): bulk-staged row-streaming kernels of the training step (xv_train_stream.cuh): parity tests, A/B of the step time
out=gpurun_out; tag=r02s
timeout 600 python -m pytest tests/test_gpu_train.py tests/test_gpu_train_api.py -m gpu -q > $out/${tag}_pytest_train.log 2>&1; tail -8 $out/${tag}_pytest_train.log | cut -c1-600
for v in 0 1 0 1; do
  XV_TRAIN_BULK=$v python bench.py --workload train --steps 300 --warmup 30 --no-cpu-baseline > $out/${tag}_train_bulk$v.json 2> $out/${tag}_train_bulk$v.err; tail -2 $out/${tag}_train_bulk$v.err
  python -c "import json; d=json.load(open('$out/${tag}_train_bulk$v.json')); print('bulk $v: ms_per_step', d['ms_per_step'], 'e2e', d['e2e']['value'], d['loss_first_last'])"
done
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file $out/${tag}_train_launches_raw.csv \
  python bench.py --workload train --steps 5 --warmup 3 --no-cpu-baseline > $out/${tag}_ncu_train.log 2>&1
python tools/summarize_ncu.py launches $out/${tag}_train_launches_raw.csv > $out/${tag}_train_launches.csv; head -30 $out/${tag}_train_launches.csv
