mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -x -q > gpurun_out/r2c_noalloc_alone.log 2>&1; tail -15 gpurun_out/r2c_noalloc_alone.log
python -m pytest tests -m gpu -x -q > gpurun_out/r2c_gputests_full.log 2>&1; tail -6 gpurun_out/r2c_gputests_full.log
