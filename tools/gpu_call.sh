mkdir -p gpurun_out
python -m pytest tests/test_read_stats.py -x -q > gpurun_out/r2d_k6tests.log 2>&1; tail -30 gpurun_out/r2d_k6tests.log
