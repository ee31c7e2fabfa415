#!/bin/bash
# r02x (1 GPU): row-streaming kernels with 3 thinner blocks per SM (2 rows in flight per thread) vs 2 fat ones
out=gpurun_out; tag=r02x
timeout 600 env XV_TRAIN_ROW_BLOCKS=3 python -m pytest tests/test_gpu_train.py -m gpu -q > $out/${tag}_pytest_train.log 2>&1; tail -3 $out/${tag}_pytest_train.log | cut -c1-600
for v in 2 3 2 3; do
  XV_TRAIN_ROW_BLOCKS=$v python bench.py --workload train --steps 300 --warmup 30 --no-cpu-baseline > $out/${tag}_train_rb$v.json 2> $out/${tag}_train_rb$v.err; tail -2 $out/${tag}_train_rb$v.err
  python -c "import json; d=json.load(open('$out/${tag}_train_rb$v.json')); print('row blocks $v: ms_per_step', d['ms_per_step'], 'e2e', d['e2e']['value'], d['loss_first_last'])"
done
