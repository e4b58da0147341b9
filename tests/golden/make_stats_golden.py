#!/usr/bin/env python
"""tests/golden/make_stats_golden.py — regenerates tests/golden/stats_golden.json.gz.

Golden vectors of the reference's OWN read-statistics loops (the per-read aggregation behind a called site): the stock
CPU driver with the fourth dump hook of oracle/e2e/patch_dindel.py --dump (oracle/_ref/dindel_ref_dump, built by
`make -C oracle e2e` from /root/reference) is run on a small synthetic chr20 BAM (oracle/e2e/synth_genome.py), once
with --doDiploid and once with --doPooled.  Per window it stores

  reads    per read: Read::mapQual, onReverseStrand, isUnmapped(), poolID
  recs     per (read, haplotype), r-major: the MLAlignment fields the loops read (ll, offHap, indels.size(), nBQT,
           nmmBQT, mLogBQ, nMMLeft, nMMRight, firstBase, lastBase)
  jobs     one per executed loop: mode 1 = "dip.map" site (DInDel.cpp:3126-3170), mode 0 = variant position
           (:3396-3404, :3504-3554), mode 2 = pooled variant x pool member (:2697-2742); the candidate haplotypes with
           what the reference's map lookups found at the site (present, startRead, endRead, hapHasVar) and the values
           the loop left in its counters (before the sqrt / division that follow)

The K6 oracle restatement (oracle/dindel_oracle.c: ora_read_stats) and the CUDA kernel (csrc/k6_read_stats.cu) are
pinned on these (tests/test_read_stats.py).  Needs /root/reference at build time only.

    python tests/golden/make_stats_golden.py
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


def cands(t, n):
    """n x (hap present start end hasVar) -> [[hap, start, end, flags]]"""
    out = []
    for i in range(n):
        h, pres, st, en, hv = (int(x) for x in t[5 * i:5 * i + 5])
        out.append([h, st, en, (1 if pres else 0) | (2 if hv else 0)])
    return out


def parse(path):
    wins, cur = [], None
    for line in open(path):
        t = line.split()
        k = t[0]
        if k in ("DW", "EW"):
            cur = {"mode": "diploid" if k == "DW" else "pooled", "nh": int(t[1]), "nr": int(t[2]), "reads": [], "recs": {}, "jobs": []}
            wins.append(cur)
        elif k == "RD":       # r mapQual reverse unmapped pool
            cur["reads"].append([float(t[2]), int(t[3]), int(t[4]), int(t[5])])
        elif k == "LK":       # h r ll offHap nIndels nBQT nmmBQT mLogBQ nMMLeft nMMRight first last
            cur["recs"][(int(t[2]), int(t[1]))] = [float(t[3]), int(t[4]), int(t[5]), int(t[6]), int(t[7]), float(t[8]),
                                                   int(t[9]), int(t[10]), int(t[11]), int(t[12])]
        elif k == "RSM":      # pos m | 2 cands | O numf numr n msq numUnmappedRealigned
            o = t.index("O")
            cur["jobs"].append({"mode": 1, "pos": int(t[1]), "group": -1, "cands": cands(t[3:o], 2),
                                "out": {"nf": int(t[o + 1]), "nr": int(t[o + 2]), "n": int(t[o + 3]), "msq": float(t[o + 4]),
                                        "numUnmappedRealigned": int(t[o + 5])}})
        elif k == "RSS":      # pos | 2 cands | O nf nr n msq allmsq numOffBoth numIndels nBQT nmmBQT mLogBQ nMMLeft nMMRight nUR
            o = t.index("O")
            names = ["nf", "nr", "n", "msq", "allmsq", "numOffBoth", "numIndels", "nBQT", "nmmBQT", "mLogBQ", "nMMLeft", "nMMRight",
                     "numUnmappedRealigned"]
            out = {n: (float(v) if n in ("msq", "allmsq", "mLogBQ") else int(v)) for n, v in zip(names, t[o + 1:])}
            cur["jobs"].append({"mode": 0, "pos": int(t[1]), "group": -1, "cands": cands(t[2:o], 2), "out": out})
        elif k == "RSP":      # idx b pos | nh cands | O nf nr n msq
            o = t.index("O")
            cur["jobs"].append({"mode": 2, "pos": int(t[3]), "group": int(t[2]), "cands": cands(t[4:o], cur["nh"]),
                                "out": {"nf": int(t[o + 1]), "nr": int(t[o + 2]), "n": int(t[o + 3]), "msq": float(t[o + 4])}})
    for w in wins:
        w["recs"] = [w["recs"][(r, h)] for r in range(w["nr"]) for h in range(w["nh"])] if w["recs"] else []
    return wins


def main():
    out = []
    with tempfile.TemporaryDirectory() as work:
        prep = run_e2e.prepare(work, 40000, seed=11)
        for mode, keep in (("--doDiploid", 7), ("--doPooled", 5)):
            dump = os.path.join(work, "dump_%s.txt" % mode[2:])
            env = dict(os.environ, DINDEL_B200_DUMP_INFER=dump, **run_e2e.HEAP_NEUTRAL)
            for k, wf in enumerate(prep["window_files"]):
                subprocess.check_call([DUMP_BIN, "--analysis", "indels", mode, "--bamFile", prep["bam"], "--ref", prep["ref"],
                                       "--varFile", wf, "--outputFile", os.path.join(work, "s%s.%d" % (mode[2:], k))],
                                      env=env, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            wins = [w for w in parse(dump) if w["jobs"] and w["recs"]]
            # the richest windows: covered sites on both strands first, then most jobs, then most haplotypes
            wins.sort(key=lambda w: (-sum(1 for j in w["jobs"] if j["out"]["nf"] and j["out"]["nr"]), -len(w["jobs"]), -w["nh"]))
            out += wins[:keep]
    dst = os.path.join(ROOT, "tests", "golden", "stats_golden.json.gz")
    with gzip.open(dst, "wt") as f:
        json.dump(out, f, separators=(",", ":"))
    print("wrote %s: %d windows, %d bytes" % (dst, len(out), os.path.getsize(dst)))
    for w in out:
        print("  %s nh %d nr %d jobs %d (modes %s) covered %s unmapped %d" % (
            w["mode"], w["nh"], w["nr"], len(w["jobs"]), sorted(set(j["mode"] for j in w["jobs"])),
            [j["out"]["n"] for j in w["jobs"]][:8], sum(r[2] for r in w["reads"])))


if __name__ == "__main__":
    main()
