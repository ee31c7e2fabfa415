#!/bin/bash
# r02l (2 GPUs): the no-alloc pooling entry point alone in a fresh process, then the whole GPU suite (it runs last there),
# then the N=2 default bench (bf16 gradient exchange) and the fp32-exchange train bench beside it
out=gpurun_out; tag=r02l
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $out/${tag}_noalloc_alone.log 2>&1; tail -5 $out/${tag}_noalloc_alone.log | cut -c1-400
timeout 1000 python -m pytest tests -m gpu -q > $out/${tag}_pytest.log 2>&1; tail -15 $out/${tag}_pytest.log | cut -c1-500
( time python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 > $out/${tag}_bench_n2.json 2> $out/${tag}_bench_n2.err ) 2>&1 | tail -3; tail -4 $out/${tag}_bench_n2.err | cut -c1-300
XV_TRAIN_COMM=fp32 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --workload train --steps 200 --warmup 20 > $out/${tag}_train_n2_fp32comm.json 2>/dev/null
python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/r02l_bench_n2.json'))
    t=d['legs']['train_dp']
    print('N=2 weak', t['weak']['ms_per_step'], 'strong', t['strong_global64']['ms_per_step'], t['weak']['allreduce'], t['parity'])
    print('ark', json.dumps(d['legs'].get('ark_e2e'))[:600])
    d=json.load(open('gpurun_out/r02l_train_n2_fp32comm.json')); print('fp32 comm N=2 weak', d['ms_per_step'], d.get('allreduce'))
except Exception as e: print('bench ERR', e)
PY
