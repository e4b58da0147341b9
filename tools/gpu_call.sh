mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -n 15 gpurun_out/pytest_gpu.log
oracle/_ref/adapter_ref_test 40
