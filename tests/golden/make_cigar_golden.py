#!/usr/bin/env python
"""tests/golden/make_cigar_golden.py — regenerates tests/golden/cigar_golden.json.gz.

Golden vectors of the reference's OWN realigned-BAM CIGAR derivation (--outputRealignedBAM): the per-read
haplotype choice at the two call sites (DInDel.cpp:502-511 pooled: first arg-max of ll over the haplotypes;
:597-610 diploid: the better of the MAP pair, ties within 1e-8 by countIndels) and DetInDel::getCIGAR
(DInDel.cpp:728-882).  The stock CPU driver with a dump hook after each getCIGAR call
(oracle/_ref/dindel_ref_dump, built by `make -C oracle e2e` from /root/reference +
oracle/e2e/patch_dindel.py --dump) is run with --outputRealignedBAM on a small synthetic chr20 BAM
(oracle/e2e/synth_genome.py), once with --doDiploid and once with --doPooled.  Per read the record holds the
inputs (ll of every haplotype, countIndels of every haplotype, the chosen pair's hpos, the haplotype's
hap.ml.hpos, refSeqPos) and the outputs (chosen haplotype, CIGAR::refPos, the CIGAR operations).
Needs /root/reference at build time only.

    python tests/golden/make_cigar_golden.py
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


def parse(path):
    recs = []
    for line in open(path):
        t = line.split()
        assert t[0] == "G"
        iL, iN, iP, iM, iC = (t.index(k, 9) for k in "LNPMC")
        recs.append({"mode": t[1], "h1": int(t[2]), "h2": int(t[3]), "nh": int(t[4]), "onHap": int(t[5]), "hmax": int(t[6]),
                     "refStart": int(t[7]), "refPos": int(t[8]),
                     "ll": [float(x) for x in t[iL + 1:iN]], "nInd": [int(x) for x in t[iN + 1:iP]],
                     "hpos": [int(x) for x in t[iP + 1:iM]], "hml": [int(x) for x in t[iM + 1:iC]],
                     "cigar": [int(x) for x in t[iC + 1:]]})
    return recs


def shape(rec):
    return "".join("MIDNSHP"[op] for op in rec["cigar"][0::2])


def main():
    out = []
    with tempfile.TemporaryDirectory() as work:
        prep = run_e2e.prepare(work, 40000, seed=11)
        for mode in ("--doDiploid", "--doPooled"):
            dump = os.path.join(work, "cig%s.txt" % mode)
            env = dict(os.environ, DINDEL_B200_DUMP_CIGAR=dump, **run_e2e.HEAP_NEUTRAL)
            for k, wf in enumerate(prep["window_files"]):
                subprocess.check_call([DUMP_BIN, "--analysis", "indels", mode, "--bamFile", prep["bam"], "--ref", prep["ref"],
                                       "--varFile", wf, "--outputFile", os.path.join(work, "c%s.%d" % (mode, k)),
                                       "--outputRealignedBAM"],
                                      env=env, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            recs = parse(dump)
            print("%s: %d reads dumped" % (mode, len(recs)))
            # keep every distinct CIGAR shape a few times, the near-ties of the diploid choice, and a plain sample
            seen, keep = {}, []
            for i, r in enumerate(recs):
                s = (shape(r), r["hmax"] == r["h1"] if r["mode"] == "D" else 0)
                tie = r["mode"] == "D" and abs(r["ll"][r["h1"]] - r["ll"][r["h2"]]) < 1e-8 and r["h1"] != r["h2"]
                if seen.get(s, 0) < (6 if len(s[0]) > 1 else 12) or (tie and seen.get("tie", 0) < 30):
                    seen[s] = seen.get(s, 0) + 1
                    if tie:
                        seen["tie"] = seen.get("tie", 0) + 1
                    keep.append(r)
            print("  kept %d, shapes: %s" % (len(keep), sorted(set(shape(r) for r in keep))))
            out += keep
    dst = os.path.join(ROOT, "tests", "golden", "cigar_golden.json.gz")
    with gzip.open(dst, "wt") as f:
        json.dump(out, f, separators=(",", ":"))
    print("wrote %s: %d reads, %d bytes" % (dst, len(out), os.path.getsize(dst)))


if __name__ == "__main__":
    main()
