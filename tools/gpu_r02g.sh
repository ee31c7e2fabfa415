#!/bin/bash
out=gpurun_out; tag=r02g
for t in 8 16; do XVK_BENCH_THREADS=$t python tools/bench_kaldi_io.py 4000 600 30 2>&1 | sed "s/^/thr$t: /" | head -3; done
XVK_BENCH_THREADS=16 XVK_MAP_POPULATE=1 python tools/bench_kaldi_io.py 4000 600 30 2>&1 | sed "s/^/thr16 populate: /" | head -3
( time python bench.py > $out/${tag}_bench.json 2> $out/${tag}_bench.err ) 2>&1 | tail -3; tail -5 $out/${tag}_bench.err | cut -c1-300
python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/r02g_bench.json'))
    print('headline', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], 'frac', d['roofline']['frac'])
    print('LEG ark_e2e', json.dumps(d['legs'].get('ark_e2e'))[:1800])
    print({k: v.get('leg_wall_s') for k,v in d['legs'].items()})
except Exception as e: print('bench ERR', e)
PY
