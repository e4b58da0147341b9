mkdir -p gpurun_out
python -m pytest tests/test_k1_gpu.py tests/test_cover.py -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -n 5 gpurun_out/pytest_gpu.log
B="python bench.py --cfg 4 --windows 600 --steps 2 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/b_cfg4_tiled.log 2>&1; echo tiled; grep -h -o '"gcups": [0-9.]*' gpurun_out/b_cfg4_tiled.log | head -1; grep -h -o '"ll_checksum": [-0-9.e]*' gpurun_out/b_cfg4_tiled.log; tail -2 gpurun_out/b_cfg4_tiled.log | cut -c1-300
DINDEL_B200_FORCE_GENERIC=1 $B > gpurun_out/b_cfg4_gen.log 2>&1; echo generic; grep -h -o '"gcups": [0-9.]*' gpurun_out/b_cfg4_gen.log | head -1; grep -h -o '"ll_checksum": [-0-9.e]*' gpurun_out/b_cfg4_gen.log
python bench.py --windows 20000 --steps 3 --warmup 3 --no-cpu-baseline --no-hpos > gpurun_out/b20k.log 2>&1; echo cfg2; grep -h -o '"gcups": [0-9.]*' gpurun_out/b20k.log | head -1
