"""K5 timing probe: dindel_b200_cigars over a configs[1]-shaped resident batch (device time of the kernel and the
wall time of the C call with host buffers)."""
import json
import sys

import numpy as np

sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))))

from dindel_tgi_b200 import engine, synth

nwin = int(sys.argv[1]) if len(sys.argv) > 1 else 6000
batch = synth.make_batch(2, nwin, seed=31)
e = engine.Engine(engine.cli_params(), 0)
e.compute_likelihoods(batch, want_hpos=False, keep_hpos=True)
hml = np.concatenate([np.arange(int(batch.hap_off[h + 1] - batch.hap_off[h]), dtype=np.int32) + 3
                      for h in range(batch.n_haps)])
nhw = np.diff(batch.win_hap)
h1, h2 = np.zeros(batch.n_windows, np.int32), (nhw - 1).astype(np.int32)
rs, ni = np.full(batch.n_windows, 100, np.int32), np.zeros(batch.n_haps, np.int32)
out = {}
for mode, name in ((engine.CIGAR_PAIR, "pair"), (engine.CIGAR_BEST, "best")):
    k, c = [], []
    for _ in range(6):
        e.cigars(mode, rs, hml, h1, h2, ni, cigar_cap=16)
        k.append(e.last_kernel_ms); c.append(e.last_call_ms)
    R = np.diff(batch.read_off)
    alg = int(2 * R.sum() + batch.n_reads * 4 * (3 + 16))    # hpos row of the chosen pair in, results out
    km = float(np.median(k[2:]))
    out[name] = {"reads": int(batch.n_reads), "kernel_ms": km, "call_ms": float(np.median(c[2:])),
                 "reads_per_s": batch.n_reads / km * 1e3, "algorithmic_bytes": alg, "GBps": alg / km / 1e6}
print(json.dumps(out))
