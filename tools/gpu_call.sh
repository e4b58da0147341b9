mkdir -p gpurun_out
P=$(cat gpurun_out/r2_pipe_explog.json 2>/dev/null || tools/pipe_explog)
python tools/k23probe.py "$P" > gpurun_out/r2_k23probe_new512.json 2> gpurun_out/k23probe_new.err; cat gpurun_out/r2_k23probe_new512.json
python -m pytest tests/test_k23.py tests/test_infer_golden.py tests/test_process_gpu.py -m gpu -x -q 2>&1 | tail -3
