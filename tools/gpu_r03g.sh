#!/bin/bash
# r03g (2 GPUs): reader without the per-batch key-split cost; waiting ranks off the GPU during the ark leg
out=gpurun_out; tag=r03g
timeout 600 python -m pytest tests/test_gpu_dropin.py -m gpu -q > $out/${tag}_pytest.log 2>&1; tail -3 $out/${tag}_pytest.log | cut -c1-300
( time python bench.py > $out/${tag}_bench_n1.json 2> $out/${tag}_bench_n1.err ) 2>&1 | tail -3
( time python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 > $out/${tag}_bench_n2.json 2> $out/${tag}_bench_n2.err ) 2>&1 | tail -3; tail -3 $out/${tag}_bench_n2.err | cut -c1-300
python - <<'PY'
import json
for n in (1, 2):
    try:
        d=json.load(open('gpurun_out/r03g_bench_n%d.json' % n))
        print('N=%d headline' % n, d['value'], 'e2e', d['e2e']['value'])
        a=d['legs']['ark_e2e']
        print('  ark_e2e', a['devices'], 'fm', a['fm']['frames_per_s'], 'reader alone', a['fm']['reader_alone_frames_per_s'], 'cm', a['cm']['frames_per_s'], a['cm']['reader_alone_frames_per_s'])
    except Exception as e: print('bench ERR', n, e)
PY
