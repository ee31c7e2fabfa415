#!/bin/bash
# r03k (1 GPU): chunked extraction after the host-side rewrite: parity tests of the views path, default bench line with every leg
out=gpurun_out; tag=r03k
timeout 300 python -m pytest tests/test_gpu_dropin.py -x -q -m gpu > $out/${tag}_pytest.log 2>&1; tail -2 $out/${tag}_pytest.log | cut -c1-300
python bench.py > $out/${tag}_bench_n1.json 2> $out/${tag}_bench_n1.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r03k_bench_n1.json'))
print('headline', d['value'], 'e2e', d['e2e']['value'], 'frac', d['roofline']['frac'])
c=d['legs']['cfg1_chunked']; print('cfg1_chunked e2e', c['e2e']['ms_per_job'], c['e2e']['input_frames_per_s'], 'resident', c['resident']['ms_per_job'], c['resident']['input_frames_per_s'])
a=d['legs']['ark_e2e']; print('ark_e2e fm', a['fm']['frames_per_s'], a['fm']['reader_alone_frames_per_s'], 'cm', a['cm']['frames_per_s'])
print('train', d['legs']['train']['ms_per_step'], 'cfg0', d['legs']['cfg0']['resident']['ms_per_job'], d['legs']['cfg0']['e2e']['ms_per_job'])
PY
