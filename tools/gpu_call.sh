mkdir -p gpurun_out
B="python bench.py --windows 3000 --steps 3 --warmup 3 --no-cpu-baseline --no-hpos"
$B > gpurun_out/b3k_m4.log 2>&1
DINDEL_B200_LIB=$PWD/dindel_tgi_b200/libdindel_b200_m3.so $B > gpurun_out/b3k_m3.log 2>&1
$B > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r1b_launches.csv $B > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k1_tiled_kernel -s 3 -c 1 -o gpurun_out/r1b_k1_tiled $B > gpurun_out/ncu2.log 2>&1
grep -h -o '"gcups": [0-9.]*' gpurun_out/b3k_m4.log gpurun_out/b3k_m3.log | head -4
