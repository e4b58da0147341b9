mkdir -p gpurun_out/deep12
python oracle/e2e/run_e2e.py --work /tmp/deep --length 12000 --coverage 1000 --seed 5 --pooled --arms gpu --windows-per-file 2 --gpu-one-file > gpurun_out/r3i_deep.json 2> gpurun_out/r3i_deep.err
cp /tmp/deep/gpu.1.glf.txt gpurun_out/deep12/; grep "^dindel_b200" /tmp/deep/gpu.1.log > gpurun_out/deep12/gpu.1.trace.log
cut -d' ' -f1 /tmp/deep/gpu.1.glf.txt | sort | uniq -c | sort -rn | head -6; cat gpurun_out/deep12/gpu.1.trace.log | cut -c1-250
