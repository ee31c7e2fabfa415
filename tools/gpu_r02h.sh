#!/bin/bash
out=gpurun_out; tag=r02h
timeout 600 python -m pytest tests -m gpu -q -k "extract or kaldi or views or cli" > $out/${tag}_pytest.log 2>&1; tail -3 $out/${tag}_pytest.log
( time python bench.py > $out/${tag}_bench.json 2> $out/${tag}_bench.err ) 2>&1 | tail -3; tail -5 $out/${tag}_bench.err | cut -c1-300
python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/r02h_bench.json'))
    print('headline', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], 'frac', d['roofline']['frac'])
    print('LEG ark_e2e', json.dumps(d['legs'].get('ark_e2e'))[:1800])
    print({k: v.get('leg_wall_s') for k,v in d['legs'].items()})
except Exception as e: print('bench ERR', e)
PY
