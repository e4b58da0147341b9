mkdir -p gpurun_out
python -m pytest tests/test_cover.py -x -q > gpurun_out/r3g_cover.log 2>&1; tail -12 gpurun_out/r3g_cover.log
