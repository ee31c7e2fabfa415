#!/bin/bash
# r02z (2 GPUs): early update of bucket 0, data-gradient recast on the side stream: training tests incl. data parallel, A/B, N=2 bench
out=gpurun_out; tag=r02z
timeout 900 python -m pytest tests/test_gpu_train.py tests/test_gpu_train_api.py tests/test_gpu_train_dp.py -m gpu -q > $out/${tag}_pytest_train.log 2>&1; tail -5 $out/${tag}_pytest_train.log | cut -c1-600
for v in 0 1 0 1; do
  XV_TRAIN_EARLY_UPDATE=$v python bench.py --workload train --steps 300 --warmup 30 --no-cpu-baseline > $out/${tag}_train_eu$v.json 2> $out/${tag}_train_eu$v.err; tail -2 $out/${tag}_train_eu$v.err
  python -c "import json; d=json.load(open('$out/${tag}_train_eu$v.json')); print('early update $v: ms_per_step', d['ms_per_step'], 'e2e', d['e2e']['value'], d['loss_first_last'])"
done
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --workload train --steps 200 --warmup 20 > $out/${tag}_train_n2.json 2>/dev/null
python -c "import json; d=json.load(open('$out/${tag}_train_n2.json')); print('N=2 weak ms_per_step', d['ms_per_step'], d.get('allreduce'))"
