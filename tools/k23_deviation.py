#!/usr/bin/env python
"""tools/k23_deviation.py — measured GPU-vs-reference deviations of K2 / K3 on the golden rl matrices
(tests/golden/infer_golden.json.gz): what the adapter's guard bands (include/dindel_b200.hpp: Inference) are sized from.
Prints one JSON line.  Needs a GPU."""
import gzip
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dindel_tgi_b200 import engine  # noqa: E402

gold = json.load(gzip.open(os.path.join(ROOT, "tests", "golden", "infer_golden.json.gz"), "rt"))
eng = engine.Engine(engine.cli_params(), 0)
tri = lambda nh: [(a, b) for a in range(nh) for b in range(a, nh)]
out = {}
dip = [w for w in gold if w["mode"] == "diploid"]
eng.load_likelihoods([np.array(w["rl"]) for w in dip], [w["nh"] for w in dip])
jobs, starts, want = [], [], []
for k, w in enumerate(dip):
    nh = w["nh"]; idx = {p: q for q, p in enumerate(tri(nh))}; n = nh * (nh + 1) // 2
    ref = np.full(n, np.nan)
    for h1, h2, ll, prior, c1, c2 in w["pairs"]:
        ref[idx[(h1, h2)]] = ll
    jobs.append(k); starts.append(np.zeros(n)); want.append(ref)
got = eng.pair_sums(jobs, starts)
rel, ratio = 0.0, 0.0
for j, k in enumerate(jobs):
    m = ~np.isnan(want[j])
    d = np.abs(got[j][m] - want[j][m])
    rel = max(rel, float((d / np.abs(want[j][m])).max()))
    b = np.array([engine.sum_error_bound(dip[k]["nr"], abs(x) + 1) for x in want[j][m]])
    ratio = max(ratio, float((d / b).max()))
out["k2_max_rel"] = rel
out["k2_max_fraction_of_bound"] = ratio
poo = [w for w in gold if w["mode"] == "pooled"]
eng.load_likelihoods([np.array(w["rl"]) for w in poo], [w["nh"] for w in poo])
sets, comps, ref = [], [], []
for k, w in enumerate(poo):
    for th, s in enumerate(w["sets"]):
        sets.append(k); comps.append(np.array(s["compatible"], dtype=np.uint8)); ref.append((w, th, s))
ll, it, pi, fr = eng.bayes_em(sets, comps, poo[0]["a0"], poo[0]["tol"])
d_ll = d_pi = d_fr = d_e = 0.0
margins = []
for j, (w, th, s) in enumerate(ref):
    nh = w["nh"]
    d_ll = max(d_ll, abs(ll[j] - s["loglik"]) / abs(s["loglik"]))
    c = np.array(s["compatible"]) == 1
    d_pi = max(d_pi, float(np.max(np.abs(pi[j][c] - np.array(s["pi"])[c]) / np.abs(np.array(s["pi"])[c]))))
    f = np.array(w["freqs"][th * nh:(th + 1) * nh])
    d_fr = max(d_fr, float(np.max(np.abs(fr[j][c] - f[c]) / f[c])))
    d_e = max(d_e, abs(eng.last_em_energy[j] - s["energy"]) / max(1.0, abs(s["energy"])))
    margins.append(float(eng.last_em_margin[j]))
out.update(k3_sets=len(ref), k3_iters_equal=bool(all(it[j] == ref[j][2]["iter"] for j in range(len(ref)))),
           k3_max_rel_loglik=d_ll, k3_max_rel_pi=d_pi, k3_max_rel_freq=d_fr, k3_max_rel_energy=d_e,
           k3_min_margin=min(margins), k3_margins_below_1e6=int(sum(m < 1e-6 for m in margins)))
print(json.dumps(out))
