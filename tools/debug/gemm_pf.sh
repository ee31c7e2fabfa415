for pf in 0 4 8 16 32; do
echo "== prefetch $pf"
XV_GEMM_PREFETCH=$pf XV_GEMM_TRACE=1 python bench.py --steps 2 --warmup 3 > gpurun_out/gemm_pf_$pf.log 2>&1
grep "gemm mma" gpurun_out/gemm_pf_$pf.log | tail -6 | head -5
tail -1 gpurun_out/gemm_pf_$pf.log | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(round(d['value']), round(d['ms_per_step'],2), {k:(round(v['ms_per_step'],2), round(v.get('tflops',0))) for k,v in d['kernel_families_ms_per_step'].items()})"
done
