#!/bin/bash
# r02d (2 GPUs): all GPU tests incl. the 2-rank ones, training step timing, default bench at N=2 (train_dp leg)
out=gpurun_out; tag=r02d
nvidia-smi -L
timeout 900 python -m pytest tests -m gpu -q > $out/${tag}_pytest.log 2>&1; tail -30 $out/${tag}_pytest.log | cut -c1-600
python bench.py --workload train --steps 300 --warmup 20 --no-cpu-baseline > $out/${tag}_train.json 2> $out/${tag}_train.err; tail -3 $out/${tag}_train.err
( time python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 > $out/${tag}_bench_n2.json 2> $out/${tag}_bench_n2.err ) 2>&1 | tail -3; tail -8 $out/${tag}_bench_n2.err | cut -c1-400
python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/r02d_train.json')); print('train ms_per_step', round(d['ms_per_step'],4), d['loss_first_last'])
except Exception as e: print('train ERR', e)
try:
    d=json.load(open('gpurun_out/r02d_bench_n2.json'))
    print('N=2 headline', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'])
    print('LEG train_dp', json.dumps(d.get('legs',{}).get('train_dp'))[:3000])
except Exception as e: print('bench ERR', e)
PY
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file $out/${tag}_train_launches_raw.csv \
  python bench.py --workload train --steps 5 --warmup 3 > $out/${tag}_ncu_train.log 2>&1
python tools/summarize_ncu.py launches $out/${tag}_train_launches_raw.csv > $out/${tag}_train_launches.csv; head -30 $out/${tag}_train_launches.csv
