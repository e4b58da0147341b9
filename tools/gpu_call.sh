mkdir -p gpurun_out
python -m pytest tests/test_k1_gpu.py -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -n 3 gpurun_out/pytest_gpu.log
B="python bench.py --cfg 4 --windows 300 --steps 2 --warmup 3 --no-cpu-baseline --no-hpos"
for v in 0 1; do
  DINDEL_B200_BT_GLOBAL=$v $B > gpurun_out/b_cfg4_bt$v.log 2>&1
  echo "bt_global=$v"; grep -h -o '"gcups": [0-9.]*' gpurun_out/b_cfg4_bt$v.log | head -1; grep -h -o '"ll_checksum": [-0-9.e]*' gpurun_out/b_cfg4_bt$v.log
done
