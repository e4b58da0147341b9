mkdir -p gpurun_out
python -m pytest tests/test_k1_gpu.py -x -q > gpurun_out/r2b_k1tests.log 2>&1; tail -4 gpurun_out/r2b_k1tests.log
python bench.py --cfg 4 --windows 600 --steps 2 --warmup 3 --no-cpu-baseline --no-extra --no-strong > gpurun_out/r2b_cfg4.json 2> gpurun_out/r2b_cfg4.err; tail -c 1500 gpurun_out/r2b_cfg4.json
python bench.py --cfg 2 --windows 6000 --steps 3 --warmup 3 --no-cpu-baseline --no-extra --no-strong > gpurun_out/r2b_cfg2.json 2> gpurun_out/r2b_cfg2.err; tail -c 1500 gpurun_out/r2b_cfg2.json
