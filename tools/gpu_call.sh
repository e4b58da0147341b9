mkdir -p gpurun_out
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_100k.log 2>&1; tail -c 2500 gpurun_out/bench_100k.log
B="python bench.py --windows 3000 --steps 2 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r1e_launches.csv $B > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k1_tiled_kernel -s 2 -c 1 -o gpurun_out/r1e_k1_tiled46 $B > gpurun_out/ncu2.log 2>&1
