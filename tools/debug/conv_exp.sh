# generic vs halo-tile convolution kernel on the two ResNet-18 stage shapes (F = 64), with and without a shortcut
for shape in "40 400 64 64 64" "40 400 32 128 128"; do
for res in 0 1; do
for halo in 0 2; do
XV_CONV_HALO=$halo timeout 120 python tools/debug/bench_conv.py $shape $res
done; done; done
