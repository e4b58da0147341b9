mkdir -p gpurun_out
B="python bench.py --cfg 2 --windows 6000 --steps 3 --warmup 3 --no-cpu-baseline --no-extra --no-strong"
for v in "" _alt; do
  [ -f dindel_tgi_b200/libdindel_b200$v.so ] || continue
  DINDEL_B200_LIB=$PWD/dindel_tgi_b200/libdindel_b200$v.so timeout 300 $B > gpurun_out/b6k$v.json 2> gpurun_out/b6k$v.err
  echo "variant '$v'"; python -c "
import json; r=json.load(open('gpurun_out/b6k$v.json')); print(r['gcups'], r['roofline']['frac'], r['e2e']['ll_checksum'])"
done
