#!/bin/bash
# r03c-d (1 GPU): round-aware wgrad split (r03c); folded batch-norm constants in the row kernels (r03d)
out=gpurun_out; tag=r03d
timeout 900 python -m pytest tests/test_gpu_wgrad.py tests/test_gpu_train.py -m gpu -q > $out/${tag}_pytest.log 2>&1; tail -4 $out/${tag}_pytest.log | cut -c1-600
for v in 1 2; do
python bench.py --workload train --steps 300 --warmup 30 --no-cpu-baseline > $out/${tag}_train_$v.json 2>/dev/null
python -c "import json; d=json.load(open('$out/${tag}_train_$v.json')); print('run $v: ms_per_step', d['ms_per_step'], 'e2e', d['e2e']['value'], d['loss_first_last'])"
done
XV_TRAIN_WGRAD_STREAM=0 XV_TRAIN_TIMELINE=100 python bench.py --workload train --steps 150 --warmup 30 --no-cpu-baseline > $out/${tag}_b.json 2> $out/${tag}_timeline_one_stream.txt; grep "timeline. total" $out/${tag}_timeline_one_stream.txt | cut -c1-3000
