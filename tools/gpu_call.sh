mkdir -p gpurun_out
python bench.py --cfg 2 --windows 6000 --steps 2 --warmup 3 --no-cpu-baseline --no-extra --no-strong > gpurun_out/r2f_cfg2.json 2> gpurun_out/r2f_cfg2.err; python -c "
import json; r=json.loads(open('gpurun_out/r2f_cfg2.json').read().strip().splitlines()[-1]); print(r['inference_kernels'])"
python bench.py --cfg 3 --windows 300 --steps 2 --warmup 3 --no-cpu-baseline --no-extra --no-strong > gpurun_out/r2f_cfg3.json 2> gpurun_out/r2f_cfg3.err; python -c "
import json; r=json.loads(open('gpurun_out/r2f_cfg3.json').read().strip().splitlines()[-1]); print(r.get('inference_kernels'))"; tail -3 gpurun_out/r2f_cfg3.err
for st in 0 1; do for mode in "" "--pooled"; do
python oracle/e2e/run_e2e.py --work /tmp/e2e_$st$mode --length 1000000 --seed 1 --arms gpu --windows-per-file 35 --gpu-one-file --gpu-device-stats $st $mode > gpurun_out/r2f_e2e_stats${st}${mode}.json 2> gpurun_out/r2f_e2e_stats${st}${mode}.err
python -c "
import json,sys; r=json.load(open('gpurun_out/r2f_e2e_stats${st}${mode}.json'))['default']['gpu']; print('stats', '$st', '$mode', r['stage3_wall_s'], r['read_stats'], r['batched'])"
done; done
