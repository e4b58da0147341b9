mkdir -p gpurun_out
python -m pytest tests/test_k1_gpu.py -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -n 3 gpurun_out/pytest_gpu.log
B="python bench.py --windows 20000 --steps 3 --warmup 3 --no-cpu-baseline --no-hpos"
for v in "" _prev _alt; do
  [ -f dindel_tgi_b200/libdindel_b200$v.so ] || continue
  DINDEL_B200_LIB=$PWD/dindel_tgi_b200/libdindel_b200$v.so $B > gpurun_out/b20k$v.log 2>&1
  echo "variant '$v'"; grep -h -o '"gcups": [0-9.]*' gpurun_out/b20k$v.log | head -1
done
