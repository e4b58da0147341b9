mkdir -p gpurun_out
python oracle/e2e/run_e2e.py --work /tmp/dd --length 8000 --coverage 400 --seed 13 --filter-haplotypes --arms cpu,gpu --windows-per-file 2 --gpu-one-file --heap-neutral > gpurun_out/r3l_dd.json 2> gpurun_out/r3l_dd.err
python -c "
import json; r=json.load(open('gpurun_out/r3l_dd.json')); h=r['heap_neutral']; print('deep diploid filter', r['n_windows'], h['glf_identical'], h['vcf_identical'], h.get('glf_diff_first'), h['cpu']['stage3_wall_s'], h['gpu']['stage3_wall_s'], h['cpu']['windows_ok'], h['gpu']['windows_ok'], h['gpu']['inference'], h['gpu']['read_stats'])" || tail -5 gpurun_out/r3l_dd.err
python oracle/e2e/run_e2e.py --work /tmp/dl --length 40000 --coverage 40 --seed 14 --read-len 250 --lib-file --stage3-args "--maxLengthIndel 10" --arms cpu,gpu --windows-per-file 3 --gpu-one-file --heap-neutral > gpurun_out/r3l_dl.json 2> gpurun_out/r3l_dl.err
python -c "
import json; r=json.load(open('gpurun_out/r3l_dl.json')); h=r['heap_neutral']; print('long reads libFile', r['n_windows'], h['glf_identical'], h['vcf_identical'], h.get('glf_diff_first'), h['cpu']['stage3_wall_s'], h['gpu']['stage3_wall_s'], h['cpu']['windows_ok'], h['gpu']['windows_ok'])" || tail -5 gpurun_out/r3l_dl.err
