#!/bin/bash
# Round-end measurement pass on one B200: GPU tests, every bench workload, the reference arm, then (separately, never
# for bench values) the ncu launch list and one full capture of the GEMM family.  Outputs under gpurun_out/$1_*.
tag=${1:-rXX}
out=gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > $out/${tag}_pytest_gpu.log 2>&1; tail -2 $out/${tag}_pytest_gpu.log
python bench.py > $out/${tag}_bench_tdnn.json 2> $out/${tag}_bench_tdnn.err
python bench.py --impl reference > $out/${tag}_bench_reference.json 2>/dev/null
python bench.py --workload attention > $out/${tag}_bench_attention.json 2>/dev/null
python bench.py --workload resnet18 --steps 5 --warmup 3 > $out/${tag}_bench_resnet18.json 2>/dev/null
python bench.py --workload train --steps 400 --warmup 20 > $out/${tag}_bench_train.json 2>/dev/null   # long enough for clock samples
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_launches_raw.csv \
  python bench.py --steps 1 --warmup 1 --utts 1024 > $out/${tag}_ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:xv_tdnn_gemm_kernel -s 12 -c 6 -o $out/${tag}_gemm_full -f \
  python bench.py --steps 1 --warmup 1 --utts 1024 > $out/${tag}_ncu_full.log 2>&1
ncu --set full --clock-control none -k regex:pool_accum_bulk_kernel -s 4 -c 3 -o $out/${tag}_pool_full -f \
  python bench.py --steps 1 --warmup 1 --utts 1024 --no-cpu-baseline > $out/${tag}_ncu_pool.log 2>&1
ls -la $out | tail -15
