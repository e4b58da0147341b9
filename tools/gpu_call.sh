mkdir -p gpurun_out
B="python bench.py --cfg 2 --windows 6000 --steps 3 --warmup 3 --no-cpu-baseline"
for v in "" _alt; do
  [ -f dindel_tgi_b200/libdindel_b200$v.so ] || continue
  DINDEL_B200_LIB=$PWD/dindel_tgi_b200/libdindel_b200$v.so $B > gpurun_out/b6k$v.log 2>&1
  echo "variant '$v'"; grep -h -o '"gcups": [0-9.]*' gpurun_out/b6k$v.log | head -1; grep -h -o '"ll_checksum": [-0-9.e]*' gpurun_out/b6k$v.log
done
