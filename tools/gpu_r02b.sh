#!/bin/bash
# r02b: training step after the kernel geometry fixes + Python training boundary tests
out=gpurun_out; tag=r02b
timeout 900 python -m pytest tests -m gpu -q -x > $out/${tag}_pytest.log 2>&1; tail -25 $out/${tag}_pytest.log
for fs in 1 0; do
  XV_TRAIN_FUSED_STATS=$fs timeout 300 python -m pytest tests/test_gpu_train.py -m gpu -q -s -k "bf16_matches" 2>&1 | grep -E "bf16 gradient noise|passed|failed|AssertionError" | cut -c1-700 | sed "s/^/fused_stats=$fs: /"
done
python bench.py --workload train --steps 300 --warmup 20 --no-cpu-baseline > $out/${tag}_train.json 2> $out/${tag}_train.err
XV_TRAIN_FUSED_STATS=0 python bench.py --workload train --steps 300 --warmup 20 > $out/${tag}_train_nofuse.json 2>/dev/null
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02b_train*.json')):
    try:
        d=json.load(open(f)); print(f, 'ms_per_step', round(d['ms_per_step'],4), 'launches/step', d['gpu_launches']/d['steps'], d['loss_first_last'])
    except Exception as e: print(f, 'ERR', e)
PY
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file $out/${tag}_train_launches_raw.csv \
  python bench.py --workload train --steps 5 --warmup 3 > $out/${tag}_ncu_train.log 2>&1
python tools/summarize_ncu.py launches $out/${tag}_train_launches_raw.csv > $out/${tag}_train_launches.csv; head -40 $out/${tag}_train_launches.csv
ncu --set full --clock-control none --import-source on -k regex:"head_wgrad|bn_bwd_apply|col_reduce4|xv_tdnn_gemm" -s 60 -c 14 -o $out/${tag}_train_full -f \
  python bench.py --workload train --steps 3 --warmup 3 > $out/${tag}_ncu_full.log 2>&1
ls -la $out | tail -8
