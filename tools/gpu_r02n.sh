#!/bin/bash
# r02n (1 GPU): weight gradients on a side stream -- training tests, A/B of the step time; ResNet-18 after the stem rewrite;
# ncu --set full with source for the headline GEMMs
out=gpurun_out; tag=r02n
timeout 600 python -m pytest tests/test_gpu_train.py tests/test_gpu_train_api.py tests/test_gpu_train_dp.py tests/test_gpu_wgrad.py -m gpu -q > $out/${tag}_pytest_train.log 2>&1; tail -8 $out/${tag}_pytest_train.log | cut -c1-400
for v in 0 1 0 1; do
  XV_TRAIN_WGRAD_STREAM=$v python bench.py --workload train --steps 300 --warmup 30 --no-cpu-baseline > $out/${tag}_train_wg$v.json 2>/dev/null
  python -c "import json; d=json.load(open('$out/${tag}_train_wg$v.json')); print('wgrad stream $v: ms_per_step', d['ms_per_step'], 'e2e', d['e2e']['value'])"
done
python bench.py --workload resnet18 --steps 5 --warmup 3 --no-cpu-baseline > $out/${tag}_resnet18.json 2>/dev/null
python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/r02n_resnet18.json')); print('resnet18', d['value'], d['ms_per_step'], {k: round(v['ms_per_step'],2) for k,v in d['kernel_families_ms_per_step'].items()})
except Exception as e: print('resnet ERR', e)
PY
ncu --set full --clock-control none --import-source on -k regex:xv_tdnn_gemm_kernel -s 20 -c 5 -o $out/${tag}_gemm_full -f \
  python bench.py --steps 1 --warmup 3 --utts 1024 --no-legs --no-cpu-baseline --no-pool-bench > $out/${tag}_ncu.log 2>&1
ls -la $out/${tag}_gemm_full.ncu-rep; tail -3 $out/${tag}_ncu.log | cut -c1-300
