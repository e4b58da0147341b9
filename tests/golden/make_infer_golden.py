#!/usr/bin/env python
"""tests/golden/make_infer_golden.py — regenerates tests/golden/infer_golden.json.gz.

Golden vectors of the reference's OWN inference arithmetic: the stock CPU driver with the third dump hook of
oracle/e2e/patch_dindel.py --dump (oracle/_ref/dindel_ref_dump, built by `make -C oracle e2e` from /root/reference)
is run on a small synthetic chr20 BAM (oracle/e2e/synth_genome.py), once with --doDiploid and once with --doPooled.
Per window it stores what DetInDel::diploidGLF (DInDel.cpp:2933-3662) resp.
DetInDel::estimateHaplotypeFrequenciesBayesEM (DInDel.cpp:2103-2930) computed:

  diploid  rl[nr*nh], filtered[nh]; per unfiltered pair h1<=h2 the read sum `ll` and prior (:3085-3091), the
           arg-max pairs with / without candidate indels, max_ll_indel / max_ll_noindel / qual (:3102-3120), per called
           site max_ll_altgeno / genoqual (:3240-3266); per variant position the per-pair `pr` and `ll` (:3366-3372),
           the grouped genotype likelihoods, maxll and its pair (:3377-3388)
  pooled   rl, poolID per read, a0, tol; per active-variant set compatible[], numah, EM passes, loglik, pi[], last
           energy (:2431-2541); logz and freqs (:2542-2543); per variant logp / freq / hapFreqs / hapHasVar
           (:2585-2602); per (variant, pool member, pair) prior and read sum (:2683-2692)

The K2 / K3 oracle restatements (oracle/dindel_oracle.c: ora_pair_sums, ora_bayes_em) and the CUDA kernels are pinned
on these (tests/test_k23.py).  Needs /root/reference at build time only.  The digamma both binaries call is the Boost
shim (oracle/shim/boost/math/special_functions/digamma.hpp -> ora_digamma): Boost itself is absent, so that one
function stays "parity unpinned" (DESIGN.md section 2).

    python tests/golden/make_infer_golden.py
"""
import gzip
import json
import os
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "oracle", "e2e"))
import run_e2e  # noqa: E402

DUMP_BIN = os.path.join(ROOT, "oracle", "_ref", "dindel_ref_dump")


def fl(xs):
    return [float(x) for x in xs]


def parse(path):
    wins, cur = [], None
    for line in open(path):
        t = line.split()
        k = t[0]
        if k == "DW":
            cur = {"mode": "diploid", "nh": int(t[1]), "nr": int(t[2]), "pairs": [], "gq": [], "site_pairs": [], "sites": []}
            wins.append(cur)
        elif k == "EW":
            cur = {"mode": "pooled", "nh": int(t[1]), "nr": int(t[2]), "a0": float(t[3]), "tol": float(t[4]), "sets": [],
                   "vars": [], "pool_pairs": []}
            wins.append(cur)
        elif k == "RL":
            cur["rl"] = fl(t[1:])
        elif k == "FILT":
            cur["filtered"] = [int(x) for x in t[1:]]
        elif k == "POOL":
            cur["pool"] = [int(x) for x in t[1:]]
        elif k == "PP":      # h1 h2 ll prior ncand1 ncand2
            cur["pairs"].append([int(t[1]), int(t[2]), float(t[3]), float(t[4]), int(t[5]), int(t[6])])
        elif k == "MAP":     # indel pair, noindel pair, max_ll_indel, max_ll_noindel, qual
            cur["map"] = [int(t[1]), int(t[2]), int(t[3]), int(t[4]), float(t[5]), float(t[6]), float(t[7])]
        elif k == "GQ":      # site, max_ll_altgeno, genoqual
            cur["gq"].append([int(t[1]), float(t[2]), float(t[3])])
        elif k == "SL":      # pos h1 h2 v1 v2 pr ll
            cur["site_pairs"].append([int(t[1]), int(t[2]), int(t[3]), int(t[4]), int(t[5]), float(t[6]), float(t[7])])
        elif k == "SM":      # pos hx1 hx2 maxll genotype:lik...
            cur["sites"].append([int(t[1]), int(t[2]), int(t[3]), float(t[4]), [[g.split(":")[0], float(g.split(":")[1])] for g in t[5:]]])
        elif k == "SET":     # th numah iter loglik logprior C ... PI ... E eOld
            c, p, e = t.index("C"), t.index("PI"), t.index("E")
            cur["sets"].append({"numah": int(t[2]), "iter": int(t[3]), "loglik": float(t[4]), "logprior": float(t[5]),
                                "compatible": [int(x) for x in t[c + 1:p]], "pi": fl(t[p + 1:e]), "energy": float(t[e + 1])})
        elif k == "LOGZ":
            cur["logz"] = float(t[1])
            cur["freqs"] = fl(t[3:])
        elif k == "VAR":     # idx logp freq HF ... HV ...
            hf, hv = t.index("HF"), t.index("HV")
            cur["vars"].append({"logp": float(t[2]), "freq": float(t[3]), "hapFreqs": fl(t[hf + 1:hv]), "hapHasVar": [int(x) for x in t[hv + 1:]]})
        elif k == "PG":      # idx b h1 h2 pr ll
            cur["pool_pairs"].append([int(t[1]), int(t[2]), int(t[3]), int(t[4]), float(t[5]), float(t[6])])
    return wins


def main():
    out = []
    with tempfile.TemporaryDirectory() as work:
        prep = run_e2e.prepare(work, 40000, seed=11)
        for mode, extra, keep in (("--doDiploid", [], 6), ("--doDiploid", ["--filterHaplotypes"], 2), ("--doPooled", [], 6)):
            dump = os.path.join(work, "dump_%s%d.txt" % (mode[2:], len(extra)))
            env = dict(os.environ, DINDEL_B200_DUMP_INFER=dump, **run_e2e.HEAP_NEUTRAL)
            for k, wf in enumerate(prep["window_files"]):
                subprocess.check_call([DUMP_BIN, "--analysis", "indels", mode, "--bamFile", prep["bam"], "--ref", prep["ref"],
                                       "--varFile", wf, "--outputFile", os.path.join(work, "d%s%d.%d" % (mode[2:], len(extra), k))] + extra,
                                      env=env, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            wins = [w for w in parse(dump) if (w["mode"] == "diploid" and "map" in w and w["sites"]) or (w["mode"] == "pooled" and w["sets"])]
            # the richest windows: most haplotypes, then most sites / sets; all with a filtered haplotype first
            wins.sort(key=lambda w: (-sum(w.get("filtered", [0])), -w["nh"], -len(w.get("sites", w.get("sets")))))
            out += wins[:keep]
    dst = os.path.join(ROOT, "tests", "golden", "infer_golden.json.gz")
    with gzip.open(dst, "wt") as f:
        json.dump(out, f, separators=(",", ":"))
    print("wrote %s: %d windows, %d bytes" % (dst, len(out), os.path.getsize(dst)))
    for w in out:
        if w["mode"] == "diploid":
            print("  diploid nh %d nr %d pairs %d sites %d filtered %s qual %.6g" % (w["nh"], w["nr"], len(w["pairs"]), len(w["sites"]), w["filtered"], w["map"][6]))
        else:
            print("  pooled nh %d nr %d sets %d iters %s logz %.6g pool_pairs %d" % (w["nh"], w["nr"], len(w["sets"]), [s["iter"] for s in w["sets"]], w["logz"], len(w["pool_pairs"])))


if __name__ == "__main__":
    main()
