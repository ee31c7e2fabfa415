#!/bin/bash
# r03a (8 GPUs): four gradient buckets (tdnn1 alone in the last one), every bucket updated behind its all-reduce
out=gpurun_out; tag=r03a
nvidia-smi -L | wc -l
timeout 600 python -m pytest tests/test_gpu_train_dp.py tests/test_gpu_train_api.py -m gpu -q > $out/${tag}_pytest_dp.log 2>&1; tail -4 $out/${tag}_pytest_dp.log | cut -c1-400
for n in 8 2; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --workload train --steps 300 --warmup 30 > $out/${tag}_train_n$n.json 2>/dev/null
python -c "import json; d=json.load(open('$out/${tag}_train_n$n.json')); print('N=$n weak ms_per_step', d['ms_per_step'], d.get('allreduce'))"
done
XV_TRAIN_EARLY_UPDATE=0 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29528 bench.py --gpus 8 --workload train --steps 300 --warmup 30 > $out/${tag}_train_n8_late.json 2>/dev/null
python -c "import json; d=json.load(open('$out/${tag}_train_n8_late.json')); print('N=8 weak, one update at the end: ms_per_step', d['ms_per_step'])"
python bench.py --workload train --steps 300 --warmup 30 --no-cpu-baseline > $out/${tag}_train_n1.json 2>/dev/null
python -c "import json; d=json.load(open('$out/${tag}_train_n1.json')); print('N=1 ms_per_step', d['ms_per_step'], 'e2e', d['e2e']['value'])"
