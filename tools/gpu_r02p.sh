#!/bin/bash
# r02p (1 GPU): pipelined training steps (xv_train_submit / xv_train_collect), resident-batch `value`, col_rnorm beside the forward pass
out=gpurun_out; tag=r02p
timeout 600 python -m pytest tests/test_gpu_train.py tests/test_gpu_train_api.py tests/test_gpu_wgrad.py -m gpu -q > $out/${tag}_pytest_train.log 2>&1; tail -12 $out/${tag}_pytest_train.log | cut -c1-600
for v in 0 all; do
  if [ $v = all ]; then unset XV_TRAIN_WGRAD_STREAM; else export XV_TRAIN_WGRAD_STREAM=$v; fi
  python bench.py --workload train --steps 300 --warmup 30 --no-cpu-baseline > $out/${tag}_train_wg$v.json 2> $out/${tag}_train_wg$v.err; tail -2 $out/${tag}_train_wg$v.err
  python -c "import json; d=json.load(open('$out/${tag}_train_wg$v.json')); print('wgrad stream $v: ms_per_step', d['ms_per_step'], 'value', d['value'], 'e2e', d['e2e'])"
done
