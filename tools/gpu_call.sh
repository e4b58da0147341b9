mkdir -p gpurun_out
python -m pytest tests/test_k1_gpu.py -x -q -k "breaks" > gpurun_out/r2p_breaks.log 2>&1; tail -3 gpurun_out/r2p_breaks.log
CMD2="python bench.py --windows 3000 --steps 2 --warmup 3 --no-cpu-baseline --no-extra --no-strong"
$CMD2 > gpurun_out/r2p_plain2.json 2> gpurun_out/r2p_plain2.err || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k1_tiled_kernel -s 6 -c 1 -o gpurun_out/r2_k1_j6_v5 $CMD2 > gpurun_out/r2p_ncu_j6.log 2>&1; echo "j6 rc $?"
