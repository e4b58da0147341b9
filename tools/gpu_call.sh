mkdir -p gpurun_out
for rep in 1 2; do for v in "" "_alt"; do
DINDEL_B200_LIB=$PWD/dindel_tgi_b200/libdindel_b200$v.so python bench.py --cfg 2 --windows 6000 --steps 3 --warmup 3 --no-cpu-baseline --no-extra --no-strong > gpurun_out/r2h_cfg2$v.$rep.json 2> gpurun_out/r2h_cfg2$v.$rep.err
python -c "
import json; r=json.loads(open('gpurun_out/r2h_cfg2$v.$rep.json').read().strip().splitlines()[-1]); print('variant [$v] rep $rep', r['gcups'], r['roofline']['frac'], r['e2e']['ll_checksum'])"
done; done
DINDEL_B200_LIB=$PWD/dindel_tgi_b200/libdindel_b200_alt.so python -m pytest tests/test_k1_gpu.py -x -q -k "every_shape_class or golden or random" 2>&1 | tail -2
