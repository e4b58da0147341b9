mkdir -p gpurun_out/deep12
python -m pytest tests/test_k23.py tests/test_infer_golden.py -x -q > gpurun_out/r3j_k23.log 2>&1; tail -5 gpurun_out/r3j_k23.log
python oracle/e2e/run_e2e.py --work /tmp/deep --length 12000 --coverage 1000 --seed 5 --pooled --arms gpu --windows-per-file 2 --gpu-one-file > gpurun_out/r3j_deep.json 2> gpurun_out/r3j_deep.err
cp /tmp/deep/gpu.1.glf.txt gpurun_out/deep12/gpu.1.glf.txt; grep "^dindel_b200" /tmp/deep/gpu.1.log > gpurun_out/deep12/gpu.1.trace.log
cut -d' ' -f1 /tmp/deep/gpu.1.glf.txt | sort | uniq -c | sort -rn | head -4; cat gpurun_out/deep12/gpu.1.trace.log | cut -c1-250
