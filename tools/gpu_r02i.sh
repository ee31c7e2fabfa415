#!/bin/bash
# r02i (8 GPUs): the driver's N=8 launch of the default bench (train_dp + ark_e2e legs)
out=gpurun_out; tag=r03j
nvidia-smi -L | wc -l
( time python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 8 > $out/${tag}_bench_n8.json 2> $out/${tag}_bench_n8.err ) 2>&1 | tail -3; tail -6 $out/${tag}_bench_n8.err | cut -c1-300
python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/r03j_bench_n8.json'))
    print('N=8 headline', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'])
    l=d['legs']
    t=l['train_dp']
    if 'error' in t: print('train_dp ERROR', t)
    else:
        print('weak', t['weak']['ms_per_step'], 'strong', t.get('strong_global64',{}).get('ms_per_step'), 'allreduce', t['weak'].get('allreduce'))
        print('parity', t['parity'])
    print('ark_e2e', json.dumps(l.get('ark_e2e'))[:1200])
except Exception as e: print('bench ERR', e)
PY
