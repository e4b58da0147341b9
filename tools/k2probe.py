import sys, time
sys.path.insert(0, '/root/repo')
import numpy as np
from dindel_tgi_b200 import engine, synth
e = engine.Engine(engine.cli_params(), 0)
b = synth.make_batch(2, 20000, seed=5)
e.upload(b, want_hpos=False); e.run()
nh = np.diff(b.win_hap).astype(np.int64)
jw = np.arange(b.n_windows, dtype=np.int32)
st = [np.zeros(int(n * (n + 1) // 2)) for n in nh]
for i in range(3):
    e.pair_sums(jw, st)
    print("K2 20k windows: call %.2f ms kernel %.3f ms" % (e.last_call_ms, e.last_kernel_ms))
