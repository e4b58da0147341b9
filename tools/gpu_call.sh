mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2_gputests_full.log 2>&1; tail -4 gpurun_out/r2_gputests_full.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke.log 2>&1; tail -2 gpurun_out/r2_smoke.log
python oracle/e2e/run_e2e.py --work /tmp/e2e1 --length 1000000 --seed 1 --arms cpu,gpu --windows-per-file 35 --procs 16 --gpu-one-file --heap-neutral > gpurun_out/r2_e2e_1mb.json 2> gpurun_out/r2_e2e_1mb.err
python -c "
import json; r=json.load(open('gpurun_out/r2_e2e_1mb.json')); print({k:(r[k]['cpu']['stage3_wall_s'], r[k]['gpu']['stage3_wall_s'], r[k]['glf_identical'], r[k]['vcf_identical'], r[k]['gpu']['batched']) for k in ('default','heap_neutral')})"
