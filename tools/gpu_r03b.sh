#!/bin/bash
# r03b (1 GPU): weighted (attention) pooled epilogue with register-held sums: parity tests, attention bench, headline check
out=gpurun_out; tag=r03b
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_kernels.py tests/test_gpu_dropin.py -m gpu -q > $out/${tag}_pytest.log 2>&1; tail -5 $out/${tag}_pytest.log | cut -c1-600
python bench.py --workload attention --steps 3 --warmup 3 --no-cpu-baseline > $out/${tag}_attention.json 2> $out/${tag}_attention.err; tail -2 $out/${tag}_attention.err
python -c "
import json; d=json.load(open('$out/${tag}_attention.json')); print('attention', d['value'], d['ms_per_step'], {k: (round(v['ms_per_step'],2), round(v.get('tflops',0))) for k,v in d['kernel_families_ms_per_step'].items()})"
python bench.py --steps 3 --warmup 3 --no-legs --no-cpu-baseline --no-pool-bench > $out/${tag}_headline.json 2>/dev/null
python -c "
import json; d=json.load(open('$out/${tag}_headline.json')); print('headline', d['value'], d['ms_per_step'], {k: (round(v['ms_per_step'],2), round(v.get('tflops',0))) for k,v in d['kernel_families_ms_per_step'].items()})"
