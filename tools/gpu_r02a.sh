#!/bin/bash
# r02a: new training step (fused kernels) -- parity, A/B timings, launch list, sanitizer passes
out=gpurun_out; tag=r02a
nvidia-smi -L | head -2
timeout 900 python -m pytest tests -m gpu -q > $out/${tag}_pytest.log 2>&1; tail -15 $out/${tag}_pytest.log
python bench.py --workload train --steps 300 --warmup 20 > $out/${tag}_train.json 2> $out/${tag}_train.err; tail -c 600 $out/${tag}_train.json
XV_TRAIN_FUSED_STATS=0 python bench.py --workload train --steps 300 --warmup 20 > $out/${tag}_train_nofuse.json 2>/dev/null
XV_GEMM_2SM=0 python bench.py --workload train --steps 300 --warmup 20 > $out/${tag}_train_2sm0.json 2>/dev/null
XV_GEMM_2SM=0 XV_GEMM_CLUSTER=0 python bench.py --workload train --steps 300 --warmup 20 > $out/${tag}_train_nocluster.json 2>/dev/null
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02a_train*.json')):
    try:
        d=json.load(open(f)); print(f, 'ms_per_step', round(d['ms_per_step'],4), 'launches/step', d['gpu_launches']/d['steps'])
    except Exception as e: print(f, 'ERR', e)
PY
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file $out/${tag}_train_launches_raw.csv \
  python bench.py --workload train --steps 5 --warmup 3 > $out/${tag}_ncu_train.log 2>&1
python tools/summarize_ncu.py launches $out/${tag}_train_launches_raw.csv > $out/${tag}_train_launches.csv; head -40 $out/${tag}_train_launches.csv
for tool in memcheck racecheck synccheck; do
  timeout 420 compute-sanitizer --tool $tool --log-file $out/${tag}_${tool}.log python tools/sanitize_smoke.py > $out/${tag}_${tool}.out 2>&1
  echo "$tool rc=$?"; tail -3 $out/${tag}_${tool}.out; tail -4 $out/${tag}_${tool}.log
done
