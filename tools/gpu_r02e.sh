#!/bin/bash
# r02e: training kernels round 3 + Python training boundary + the default bench line with every leg
out=gpurun_out; tag=r02e
timeout 900 python -m pytest tests -m gpu -q > $out/${tag}_pytest.log 2>&1; tail -30 $out/${tag}_pytest.log | cut -c1-400
python bench.py --workload train --steps 300 --warmup 20 --no-cpu-baseline > $out/${tag}_train.json 2> $out/${tag}_train.err; tail -3 $out/${tag}_train.err
( time python bench.py > $out/${tag}_bench.json 2> $out/${tag}_bench.err ) 2>&1 | tail -3; tail -5 $out/${tag}_bench.err
python - <<'PY'
import json
for f in ('gpurun_out/r02e_train.json',):
    try:
        d=json.load(open(f)); print(f, 'ms_per_step', round(d['ms_per_step'],4), 'launches/step', d['gpu_launches']/d['steps'], d['loss_first_last'])
    except Exception as e: print(f, 'ERR', e)
try:
    d=json.load(open('gpurun_out/r02e_bench.json'))
    print('headline', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], 'frac', d['roofline']['frac'])
    for k,v in d.get('legs',{}).items():
        s=json.dumps(v); print('LEG', k, s[:900])
except Exception as e: print('bench ERR', e)
PY
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file $out/${tag}_train_launches_raw.csv \
  python bench.py --workload train --steps 5 --warmup 3 > $out/${tag}_ncu_train.log 2>&1
python tools/summarize_ncu.py launches $out/${tag}_train_launches_raw.csv > $out/${tag}_train_launches.csv; head -40 $out/${tag}_train_launches.csv
