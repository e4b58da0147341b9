# scratch: the command list of the last gpurun call (`gpurun -- 'bash tools/gpu_call.sh'`); the round-end set is
#   pytest -m gpu, smoke(), default bench
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/gputests_full.log 2>&1; tail -4 gpurun_out/gputests_full.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; tail -c 400 gpurun_out/bench_n1.json
