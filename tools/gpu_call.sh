mkdir -p gpurun_out
nvidia-smi -L | head -4
python -m pytest tests/test_process_gpu.py tests/test_e2e.py -x -q -k "multi or several" > gpurun_out/r2t_multi_tests.log 2>&1; tail -4 gpurun_out/r2t_multi_tests.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 > gpurun_out/r2t_bench_n2.json 2> gpurun_out/r2t_bench_n2.err; python -c "
import json; r=json.loads(open('gpurun_out/r2t_bench_n2.json').read().strip().splitlines()[-1]); print(r['n_gpus'], r['value'], r['gcups'], 'e2e', r['e2e']['value']); s=r.get('strong'); print({k:(v['ms'], v['pairs_per_s']) for k,v in s.items() if isinstance(v, dict) and 'ms' in v}); print(s.get('efficiency'), s.get('identical'))"; tail -3 gpurun_out/r2t_bench_n2.err
