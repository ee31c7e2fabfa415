#!/bin/bash
# r02t (1 GPU): ncu --set full of the bulk-staged row-streaming kernels inside a training step
out=gpurun_out; tag=r02t
ncu --set full --clock-control none --import-source on -k regex:bn_stream_kernel -s 28 -c 14 -o $out/${tag}_bn_stream_full -f \
  python bench.py --workload train --steps 2 --warmup 3 --no-cpu-baseline > $out/${tag}_ncu.log 2>&1
ls -la $out/${tag}_bn_stream_full.ncu-rep; tail -2 $out/${tag}_ncu.log | cut -c1-200
