mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2o_gputests_full.log 2>&1; tail -5 gpurun_out/r2o_gputests_full.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2o_smoke.log 2>&1; tail -2 gpurun_out/r2o_smoke.log
python bench.py > gpurun_out/r2o_bench_n1.json 2> gpurun_out/r2o_bench_n1.err; tail -c 300 gpurun_out/r2o_bench_n1.json; tail -2 gpurun_out/r2o_bench_n1.err
