#!/usr/bin/env python
"""tools/k23probe.py — K2 / K3 at BASELINE configs[2] shape (1000 reads x 16 haplotypes per window): kernel times (CUDA
events inside the library), pair sums / EM sets per second, and the achieved exp / addLogs rates against the
register-resident rates of tools/pipe_explog (pass its JSON line as argv[1]).  DINDEL_B200_LIB selects another build of
the library (e.g. the round-1 kernels).  Prints one JSON line."""
import json
import sys
import os

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dindel_tgi_b200 import engine, synth  # noqa: E402

peaks = json.loads(sys.argv[1]) if len(sys.argv) > 1 else {}
nw = int(os.environ.get("K23_WINDOWS", "200"))
e = engine.Engine(engine.cli_params(), 0)
b = synth.make_batch(3, nw, seed=5)
e.upload(b, want_hpos=False)
e.run()
nh = np.diff(b.win_hap).astype(np.int64)
nr = np.diff(b.win_read).astype(np.int64)
jw = np.arange(b.n_windows, dtype=np.int32)
st = [np.zeros(int(n * (n + 1) // 2)) for n in nh]
for _ in range(3):
    sums = e.pair_sums(jw, st)
k2_ms = e.last_kernel_ms
n_addlogs = float(np.sum(nh * (nh + 1) // 2 * nr))
rng = np.random.default_rng(3)
sets, comps = [], []
for w in range(b.n_windows):
    for k in range(16):
        c = np.ones(nh[w], dtype=np.uint8) if k == 0 else (rng.random(nh[w]) < 0.7).astype(np.uint8)
        c[0] = 1
        sets.append(w); comps.append(c)
for _ in range(3):
    ll, it, pi, fr = e.bayes_em(sets, comps, 0.001, 1e-4)
k3_ms = e.last_kernel_ms
passes = float(np.sum(it))
n_exp = float(np.sum(it * np.array([nr[w] * nh[w] for w in sets])))
out = {"lib": os.path.basename(engine.LIB_PATH), "windows": nw, "k2_kernel_ms": k2_ms, "k2_addlogs_per_s": n_addlogs / (k2_ms * 1e-3),
       "k2_checksum": float(sum(float(x.sum()) for x in sums)),
       "k3_kernel_ms": k3_ms, "k3_sets": len(sets), "k3_sets_per_s": len(sets) / (k3_ms * 1e-3), "k3_mean_passes": passes / len(sets),
       "k3_read_hap_terms_per_s": n_exp / (k3_ms * 1e-3), "k3_loglik_checksum": float(ll.sum())}
if peaks:
    out["k2_frac_of_addlogs_peak"] = out["k2_addlogs_per_s"] / peaks["addlogs_per_s"]
    out["k3_frac_of_exp_peak"] = out["k3_read_hap_terms_per_s"] / peaks["exp_per_s"]
    out["peaks"] = peaks
print(json.dumps(out))
