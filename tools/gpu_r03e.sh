#!/bin/bash
# r03e / r03h (2 GPUs): end-of-round validation -- whole GPU suite, smoke(), default bench at N=1 and N=2, reference arm
out=gpurun_out; tag=r03h
timeout 1500 python -X faulthandler -m pytest tests -x -q -m gpu > $out/${tag}_pytest.log 2>&1; tail -6 $out/${tag}_pytest.log | cut -c1-400
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $out/${tag}_smoke.log 2>&1; tail -2 $out/${tag}_smoke.log | cut -c1-300
( time python bench.py > $out/${tag}_bench_n1.json 2> $out/${tag}_bench_n1.err ) 2>&1 | tail -3
( time python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 > $out/${tag}_bench_n2.json 2> $out/${tag}_bench_n2.err ) 2>&1 | tail -3
( time python bench.py --impl reference --steps 2 --warmup 1 > $out/${tag}_bench_ref.json 2> $out/${tag}_bench_ref.err ) 2>&1 | tail -3
python - <<'PY'
import json
for n in (1, 2):
    try:
        d=json.load(open('gpurun_out/r03h_bench_n%d.json' % n))
        print('N=%d headline' % n, d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], 'frac', d['roofline']['frac'], 'launches', d['gpu_launches'], d['clocks'])
        for k,v in d.get('legs',{}).items():
            s=json.dumps(v); print('LEG', k, s[:420])
    except Exception as e: print('bench ERR', n, e)
try:
    d=json.load(open('gpurun_out/r03h_bench_ref.json')); print('reference arm', {k: d[k] for k in ('impl','value','unit','ms_per_step') if k in d}, d.get('cpu_baseline'))
except Exception as e: print('ref ERR', e)
PY
