mkdir -p gpurun_out
python -m pytest tests/test_k1_gpu.py -x -q > gpurun_out/r2w_k1tests.log 2>&1; tail -2 gpurun_out/r2w_k1tests.log
python bench.py --cfg 4 --windows 600 --steps 2 --warmup 3 --no-cpu-baseline --no-extra --no-strong > gpurun_out/r2w_cfg4.json 2> gpurun_out/r2w_cfg4.err
python -c "
import json; r=json.loads(open('gpurun_out/r2w_cfg4.json').read().strip().splitlines()[-1]); print('cfg4', r['ms_per_step'], r['gcups'], r['roofline']['frac'], r['e2e']['ll_checksum'])"
python bench.py --cfg 2 --windows 6000 --steps 3 --warmup 3 --no-cpu-baseline --no-extra --no-strong > gpurun_out/r2w_cfg2.json 2> gpurun_out/r2w_cfg2.err
python -c "
import json; r=json.loads(open('gpurun_out/r2w_cfg2.json').read().strip().splitlines()[-1]); print('cfg2', r['ms_per_step'], r['gcups'], r['roofline']['frac'], r['e2e']['ll_checksum'])"
