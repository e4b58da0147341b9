mkdir -p gpurun_out
B="python bench.py --windows 3000 --steps 2 --warmup 3 --no-cpu-baseline --no-hpos"
$B > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k1_tiled_kernel -s 2 -c 1 -o gpurun_out/r1c_k1_tiled46 $B > gpurun_out/ncu2.log 2>&1
