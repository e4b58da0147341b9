mkdir -p gpurun_out
python oracle/e2e/run_e2e.py --work /tmp/long --length 60000 --coverage 40 --seed 9 --read-len 250 --stage3-args "--maxLengthIndel 10 --maxHap 32" --arms cpu,gpu --windows-per-file 3 --gpu-one-file --heap-neutral > gpurun_out/r3k_long.json 2> gpurun_out/r3k_long.err
python -c "
import json; r=json.load(open('gpurun_out/r3k_long.json')); h=r['heap_neutral']; print('long reads', r['n_windows'], h['glf_identical'], h['vcf_identical'], h.get('glf_diff_first'), h['cpu']['stage3_wall_s'], h['gpu']['stage3_wall_s'], h['cpu']['windows_ok'], h['gpu']['windows_ok'], h['gpu']['inference'], h['cpu']['vcf_calls'])" || tail -5 gpurun_out/r3k_long.err
python oracle/e2e/run_e2e.py --work /tmp/longp --length 30000 --coverage 40 --seed 9 --read-len 250 --pooled --stage3-args "--maxLengthIndel 10 --maxHap 32" --arms cpu,gpu --windows-per-file 3 --gpu-one-file --heap-neutral > gpurun_out/r3k_longp.json 2> gpurun_out/r3k_longp.err
python -c "
import json; r=json.load(open('gpurun_out/r3k_longp.json')); h=r['heap_neutral']; print('long reads pooled', r['n_windows'], h['glf_identical'], h.get('glf_diff_first'), h['cpu']['stage3_wall_s'], h['gpu']['stage3_wall_s'], h['cpu']['windows_ok'], h['gpu']['windows_ok'], h['gpu']['inference'])" || tail -5 gpurun_out/r3k_longp.err
