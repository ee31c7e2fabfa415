#!/bin/bash
# r02k (1 GPU): ncu --set full with source for the headline GEMMs, ResNet-18 bench after the stem rewrite
out=gpurun_out; tag=r02k
python bench.py --workload resnet18 --steps 5 --warmup 3 --no-cpu-baseline > $out/${tag}_resnet18.json 2>/dev/null
python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/r02k_resnet18.json')); print('resnet18', d['value'], d['ms_per_step'], {k: round(v['ms_per_step'],2) for k,v in d['kernel_families_ms_per_step'].items()})
except Exception as e: print('resnet ERR', e)
PY
ncu --set full --clock-control none --import-source on -k regex:xv_tdnn_gemm_kernel -s 20 -c 5 -o $out/${tag}_gemm_full -f \
  python bench.py --steps 1 --warmup 3 --utts 1024 --no-legs --no-cpu-baseline --no-pool-bench > $out/${tag}_ncu.log 2>&1
ls -la $out/${tag}_gemm_full.ncu-rep; tail -3 $out/${tag}_ncu.log | cut -c1-300
