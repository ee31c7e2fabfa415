#!/bin/bash
# r02y (1 GPU): vector reductions in the skinny GEMMs, gradient zeros on the side stream
out=gpurun_out; tag=r02y
timeout 600 python -m pytest tests/test_gpu_train.py tests/test_gpu_train_api.py -m gpu -q > $out/${tag}_pytest_train.log 2>&1; tail -5 $out/${tag}_pytest_train.log | cut -c1-600
for v in 1 2; do
  python bench.py --workload train --steps 300 --warmup 30 --no-cpu-baseline > $out/${tag}_train_$v.json 2> $out/${tag}_train_$v.err; tail -2 $out/${tag}_train_$v.err
  python -c "import json; d=json.load(open('$out/${tag}_train_$v.json')); print('run $v: ms_per_step', d['ms_per_step'], 'e2e', d['e2e']['value'], d['loss_first_last'])"
done
XV_TRAIN_WGRAD_STREAM=0 XV_TRAIN_TIMELINE=100 python bench.py --workload train --steps 150 --warmup 30 --no-cpu-baseline > $out/${tag}_b.json 2> $out/${tag}_timeline_one_stream.txt; grep "timeline" $out/${tag}_timeline_one_stream.txt | cut -c1-3000
