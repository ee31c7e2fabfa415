for shape in "40 400 64 64 64" "40 400 32 128 128"; do
for res in 0 1; do
XV_CONV_HALO=0 python tools/debug/bench_conv.py $shape $res
for dbg in 0 3; do
XV_CONV_HALO_DEBUG=$dbg timeout 120 python tools/debug/bench_conv.py $shape $res
done; done; done
