#!/bin/bash
# r02r (1 GPU): whole GPU suite with the pipelined training API; default bench with every leg
out=gpurun_out; tag=r02r
timeout 1200 python -X faulthandler -m pytest tests -m gpu -q > $out/${tag}_pytest.log 2>&1; tail -8 $out/${tag}_pytest.log | cut -c1-400
( time python bench.py > $out/${tag}_bench.json 2> $out/${tag}_bench.err ) 2>&1 | tail -3; tail -5 $out/${tag}_bench.err | cut -c1-300
python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/r02r_bench.json'))
    print('headline', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], 'frac', d['roofline']['frac'], 'traffic', d['roofline'].get('traffic'), d['roofline'].get('traffic_source'))
    for k,v in d.get('legs',{}).items():
        s=json.dumps(v); print('LEG', k, s[:700])
except Exception as e: print('bench ERR', e)
PY
