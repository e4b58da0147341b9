#!/usr/bin/env python
"""tests/golden/make_cover_golden.py — regenerates tests/golden/cover_golden.json.gz.

Golden vectors of the reference's OWN DetInDel::filterHaplotypes (DInDel.cpp:1932-2100): the stock CPU
driver with a dump hook at the end of that function (oracle/_ref/dindel_ref_dump, built by
`make -C oracle e2e` from /root/reference + oracle/e2e/patch_dindel.py --dump) is run on a small synthetic
chr20 BAM (oracle/e2e/synth_genome.py), with and without --filterHaplotypes, under the heap-neutral
allocator settings of oracle/e2e/run_e2e.py (so that the function's stray hpos[size()] read is "no hit").
Every window's inputs (haplotypes, reads, per-pair offHapHMQ / numIndels / hpos, variant entries) and
outputs (filtered[], varCoverage) are stored.  Needs /root/reference at build time only.

    python tests/golden/make_cover_golden.py
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
    wins, cur = [], None
    for line in open(path):
        t = line.split()
        if t[0] == "W":
            cur = {"nh": int(t[1]), "nr": int(t[2]), "padCover": int(t[3]), "maxMismatch": int(t[4]), "doFilter": int(t[5]),
                   "haps": [], "reads": [], "strand": [], "pairs": [], "entries": [], "filtered": [], "cov": []}
            wins.append(cur)
        elif t[0] == "H":
            cur["haps"].append(t[1])
        elif t[0] == "R":
            cur["reads"].append(t[1]); cur["strand"].append(int(t[2]))
        elif t[0] == "P":
            cur["pairs"].append([int(x) for x in t[1:]])       # h r offHapHMQ numIndels hpos...
        elif t[0] == "E":
            cur["entries"].append([int(t[1]), t[2], int(t[3]), int(t[4]), int(t[5])])  # h key type left right
        elif t[0] == "F":
            cur["filtered"] = [int(x) for x in t[1:]]
        elif t[0] == "C":
            cur["cov"].append([t[1], int(t[2]), int(t[3])])
    return wins


def main():
    out = []
    with tempfile.TemporaryDirectory() as work:
        prep = run_e2e.prepare(work, 30000, seed=7)
        for extra in ([], ["--filterHaplotypes"]):
            dump = os.path.join(work, "dump%d.txt" % len(extra))
            env = dict(os.environ, DINDEL_B200_DUMP_COVER=dump, **run_e2e.HEAP_NEUTRAL)
            for k, wf in enumerate(prep["window_files"]):
                subprocess.check_call([DUMP_BIN, "--analysis", "indels", "--doDiploid", "--bamFile", prep["bam"], "--ref", prep["ref"],
                                       "--varFile", wf, "--outputFile", os.path.join(work, "d%d.%d" % (len(extra), k))] + extra,
                                      env=env, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            wins = parse(dump)
            # keep the windows with the most variant entries (and all with a filtered haplotype)
            wins.sort(key=lambda w: (-sum(w["filtered"]), -len(w["entries"])))
            out += wins[:5]
    dst = os.path.join(ROOT, "tests", "golden", "cover_golden.json.gz")
    with gzip.open(dst, "wt") as f:
        json.dump(out, f, separators=(",", ":"))
    print("wrote %s: %d windows, %d bytes" % (dst, len(out), os.path.getsize(dst)))
    for w in out:
        print("  nh %d nr %d entries %d doFilter %d filtered %s cov %s" % (w["nh"], w["nr"], len(w["entries"]), w["doFilter"], w["filtered"], w["cov"][:4]))


if __name__ == "__main__":
    main()
