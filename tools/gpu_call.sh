mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2s_gputests_full.log 2>&1; tail -4 gpurun_out/r2s_gputests_full.log
python bench.py > gpurun_out/r2s_bench_n1.json 2> gpurun_out/r2s_bench_n1.err; python -c "
import json; r=json.loads(open('gpurun_out/r2s_bench_n1.json').read().strip().splitlines()[-1]); print(r['value'], r['gcups'], r['roofline']['frac'], 'e2e', r['e2e']['value'], r['e2e']['ms_per_step']); print({k:(v['value'], v['e2e']['value']) for k,v in r['extra_configs'].items()}); print(r['strong']['n1']['ms'])"
