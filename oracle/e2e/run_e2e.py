#!/usr/bin/env python
"""oracle/e2e/run_e2e.py — TEST INFRASTRUCTURE ONLY.

End-to-end check of BASELINE.json configs[0] (`dindel --analysis indels --doDiploid` on a synthetic
chr20 BAM, 30x 100-bp reads, ~120-bp windows from makeWindows): the reference's own pipeline is run
twice over the same window files,
  arm "cpu": oracle/_ref/dindel_ref   (stock reference binary)
  arm "gpu": oracle/_ref/dindel_b200  (same sources + INTEGRATION.md's three-line patch: K1 on the B200)
and the `.glf.txt` files and the merged VCFs are compared byte for byte.

  stage 1  dindel_ref --analysis getCIGARindels          (README.md pipeline; GetCandidates.cpp:387)
  stage 2  make_windows.py (selectCandidates --minCount 2, makeWindows; window files = unit of parallelism)
  stage 3  <arm binary> --analysis indels --doDiploid    one process per window file, `procs` at a time
  stage 4  merge_diploid.py -> VCF

    python oracle/e2e/run_e2e.py --work /tmp/e2e --length 60000 --arms cpu,gpu
prints one JSON line with counts, wall times and the comparison verdict.
"""
import argparse
import json
import os
import subprocess
import sys
import time
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
REFDIR = os.path.join(os.path.dirname(HERE), "_ref")
sys.path.insert(0, HERE)
import make_windows  # noqa: E402
import merge_diploid  # noqa: E402

BIN = {"cpu": os.path.join(REFDIR, "dindel_ref"), "gpu": os.path.join(REFDIR, "dindel_b200"),
       "check": os.path.join(REFDIR, "dindel_b200_check")}
# DetInDel::filterHaplotypes reads hpos[b] for b == hpos.size() (DInDel.cpp:1995,2034: loop bound `<=`), one
# int of heap slack past the vector: the coverage counts (glf columns var_coverage_forward/reverse -> VCF NF/NR)
# then depend on what the allocator left there, in the stock binary as much as in the GPU-backed one.  With
# these two variables glibc poisons every freed chunk (no tcache), so the stray int is never a valid haplotype
# position in either arm ("no hit"); run_e2e reports the comparison both ways.
HEAP_NEUTRAL = {"GLIBC_TUNABLES": "glibc.malloc.tcache_count=0", "MALLOC_PERTURB_": "85"}
BAMTOOL = os.path.join(REFDIR, "bamtool")


def have_binaries(arms=("cpu",)):
    return os.path.exists(BAMTOOL) and all(os.path.exists(BIN[a]) for a in arms)


def sh(cmd, log, env=None):
    with open(log, "w") as f:
        rc = subprocess.call(cmd, stdout=f, stderr=subprocess.STDOUT, env=dict(os.environ, **env) if env else None)
    if rc != 0:
        raise RuntimeError("%s failed (rc %d), see %s:\n%s" % (cmd[0], rc, log, open(log).read()[-2000:]))


def prepare(work, length, seed, coverage=30.0, per_file=1000, min_count=2, read_len=100):
    """synthetic genome -> BAM -> candidates -> window files.  Returns dict of paths."""
    os.makedirs(work, exist_ok=True)
    sh([sys.executable, os.path.join(HERE, "synth_genome.py"), "--out", work, "--length", str(length),
        "--coverage", str(coverage), "--seed", str(seed), "--read-len", str(read_len)], os.path.join(work, "synth.log"))
    ref, sam, bam = (os.path.join(work, n) for n in ("ref.fa", "sample.sam", "sample.bam"))
    sh([BAMTOOL, "faidx", ref], os.path.join(work, "faidx.log"))
    sh([BAMTOOL, "sam2bam", sam, bam], os.path.join(work, "sam2bam.log"))
    s1 = os.path.join(work, "s1")
    sh([BIN["cpu"], "--analysis", "getCIGARindels", "--bamFile", bam, "--ref", ref, "--outputFile", s1],
       os.path.join(work, "stage1.log"))
    sel = os.path.join(work, "win.selected.txt")
    make_windows.select_candidates(s1 + ".variants.txt", sel, min_count)
    files = make_windows.split_and_merge(sel, os.path.join(work, "win"), 20, per_file)
    n_win = sum(1 for f in files for _ in open(f))
    return {"ref": ref, "bam": bam, "window_files": files, "n_windows": n_win, "libraries": s1 + ".libraries.txt",
            "n_candidates": sum(1 for _ in open(sel))}


def run_arm(arm, prep, work, procs, extra=(), env=None, tag="", mode="--doDiploid", one_file=False):
    """stage 3 + 4 for one arm.  Returns (glf paths, vcf path, wall seconds of stage 3, VCF calls).
    one_file: every window file concatenated into one and handled by ONE process (the GPU-backed binary batches the
    window loop itself: host threads for the CPU stages, one GPU pass per batch of windows)."""
    window_files = prep["window_files"]
    if one_file:
        allw = os.path.join(work, "win.all.txt")
        with open(allw, "w") as f:
            for wf in window_files:
                f.write(open(wf).read())
        window_files = [allw]

    def one(k_f):
        k, wf = k_f
        out = os.path.join(work, "%s%s.%d" % (arm, tag, k))
        sh([BIN[arm], "--analysis", "indels", mode, "--bamFile", prep["bam"], "--ref", prep["ref"],
            "--varFile", wf, "--outputFile", out] + list(extra), out + ".log", env)
        return out + ".glf.txt"
    t0 = time.perf_counter()
    with ThreadPoolExecutor(max(1, procs)) as ex:
        glfs = list(ex.map(one, enumerate(window_files, 1)))
    wall = time.perf_counter() - t0
    vcf = os.path.join(work, arm + tag + ".vcf")
    if mode == "--doDiploid":
        n_calls = merge_diploid.merge_output(glfs, prep["ref"], vcf)
    else:
        open(vcf, "w").close()
        n_calls = 0
    return glfs, vcf, wall, n_calls


def glf_stats(glfs):
    ok = err = 0
    for g in glfs:
        with open(g) as f:
            f.readline()
            seen = set()
            for line in f:
                d = line.split(" ", 2)
                if d[1] in seen:
                    continue
                seen.add(d[1])
                if d[0] == "ok":
                    ok += 1
                else:
                    err += 1
    return ok, err


def check_totals(glfs):
    """B200CHECK lines of the differential binary (patch_dindel.py --check): pairs compared, mismatches."""
    pairs = bad = 0
    for g in glfs:
        last = None
        for line in open(g[:-len(".glf.txt")] + ".log", errors="replace"):
            if line.startswith("B200CHECK window"):
                last = line.split()
        if last:
            pairs += int(last[last.index("total_pairs") + 1])
            bad += int(last[last.index("total_mismatches") + 1])
    return pairs, bad


def inference_totals(glfs):
    """`dindel_b200 inference:` lines of the GPU-backed binary (one per process, at exit): windows whose diploidGLF /
    Bayes EM numeric cores ran on K2 / K3, and how many of them the guard band sent back to the reference's CPU loops."""
    tot = {"windows": 0, "rechecked_on_host": 0, "em_sets": 0, "em_set_mismatch": 0}
    for g in glfs:
        for line in open(g[:-len(".glf.txt")] + ".log", errors="replace"):
            if line.startswith("dindel_b200 inference:"):
                t = line.split()
                for k in tot:
                    tot[k] += int(t[t.index(k) + 1])
    return tot


def read_stats_totals(glfs):
    """`dindel_b200 read stats:` lines: K6 calls of the GPU-backed binary and the jobs (site loops) they served."""
    tot = {"device_calls": 0, "jobs": 0}
    for g in glfs:
        for line in open(g[:-len(".glf.txt")] + ".log", errors="replace"):
            if line.startswith("dindel_b200 read stats:"):
                t = line.split()
                for k in tot:
                    tot[k] += int(t[t.index(k) + 1])
    return tot


def diff_lines(a_files, b_files, limit=5):
    out = []
    if len(a_files) != len(b_files):
        # one arm ran every window in one process: compare the concatenated bodies with the `index` column (the window's
        # line number inside ITS window file, DInDel.cpp:1345) masked
        def body(files):
            rows = []
            for f in files:
                for line in open(f).read().splitlines()[1:]:
                    d = line.split(" ")
                    d[1] = "*"
                    rows.append(" ".join(d))
            return rows
        la, lb = body(a_files), body(b_files)
        if len(la) != len(lb):
            out.append("%d vs %d lines" % (len(la), len(lb)))
        for i, (p, q) in enumerate(zip(la, lb)):
            if p != q:
                out.append("line %d cols %s" % (i + 1, [k for k, (u, v) in enumerate(zip(p.split(" "), q.split(" "))) if u != v]))
        return len(out), out[:limit]
    for x, y in zip(a_files, b_files):
        la, lb = open(x).read().splitlines(), open(y).read().splitlines()
        if len(la) != len(lb):
            out.append("%s: %d vs %d lines" % (os.path.basename(x), len(la), len(lb)))
        for i, (p, q) in enumerate(zip(la, lb)):
            if p != q:
                cols = [k for k, (u, v) in enumerate(zip(p.split(" "), q.split(" "))) if u != v]
                out.append("%s line %d cols %s" % (os.path.basename(x), i + 1, cols))
    return len(out), out[:limit]


def bam_records(path):
    """Alignment records of a BAM file (SAM/BAM spec 4.2): list of (core fields, cigar, rest-of-record bytes)."""
    import gzip
    import struct
    d = gzip.open(path, "rb").read()          # BGZF = concatenated gzip members
    assert d[:4] == b"BAM\1", path
    l_text, = struct.unpack_from("<i", d, 4)
    o = 8 + l_text
    n_ref, = struct.unpack_from("<i", d, o)
    o += 4
    for _ in range(n_ref):
        l_name, = struct.unpack_from("<i", d, o)
        o += 4 + l_name + 4
    recs = []
    while o < len(d):
        bs, = struct.unpack_from("<i", d, o)
        ref_id, pos, l_name, mapq, bin_, n_cig, flag, l_seq, nref, npos, tlen = struct.unpack_from("<iiBBHHHiiii", d, o + 4)
        q = o + 36
        name = d[q:q + l_name]
        cig = struct.unpack_from("<%dI" % n_cig, d, q + l_name)
        rest = d[q + l_name + 4 * n_cig:o + 4 + bs]
        recs.append({"name": name, "ref": ref_id, "pos": pos, "mapq": mapq, "flag": flag, "l_seq": l_seq, "nref": nref,
                     "npos": npos, "tlen": tlen, "cigar": cig, "rest": rest})
        o += 4 + bs
    return recs


def compare_realigned_bams(work, arm_a, arm_b, tag):
    """The per-window realigned BAMs (<out>.ra.<index>_<tid>_<left>_<right>.bam, DInDel.cpp:517-519 / :615-619) of two
    arms, record by record.  pos / isize of a wholly soft-clipped read are not compared: the reference never sets
    CIGAR::refPos for it (DInDel.cpp:782-786), so its own value is whatever the stack held."""
    import glob
    fa = sorted(glob.glob(os.path.join(work, "%s%s.*.ra.*.bam" % (arm_a, tag))))
    n_rec = n_bad = n_changed = 0
    first = None
    for f in fa:
        g = os.path.join(work, arm_b + os.path.basename(f)[len(arm_a):])
        if not os.path.exists(g):
            n_bad += 1
            first = first or "missing " + g
            continue
        ra, rb = bam_records(f), bam_records(g)
        if len(ra) != len(rb):
            n_bad += 1
            first = first or "record count " + f
            continue
        for x, y in zip(ra, rb):
            n_rec += 1
            clipped = len(x["cigar"]) == 1 and x["cigar"][0] == (x["l_seq"] << 4 | 4)
            keys = [k for k in x if not (clipped and k in ("pos", "tlen"))]
            n_changed += any(c & 15 in (1, 2, 4) for c in x["cigar"])
            if any(x[k] != y[k] for k in keys):
                n_bad += 1
                first = first or "%s %s: %s vs %s" % (os.path.basename(f), x["name"], {k: x[k] for k in keys if k != "rest"},
                                                      {k: y[k] for k in keys if k != "rest"})
    return {"files": len(fa), "records": n_rec, "records_with_indel_or_clip": n_changed, "mismatches": n_bad, "first": first}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--work", required=True)
    ap.add_argument("--length", type=int, default=60000)
    ap.add_argument("--coverage", type=float, default=30.0)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--arms", default="cpu,gpu")
    ap.add_argument("--procs", type=int, default=0, help="concurrent stage-3 processes (0 = host cores)")
    ap.add_argument("--gpu-procs", type=int, default=4, help="concurrent stage-3 processes sharing the GPU")
    ap.add_argument("--windows-per-file", type=int, default=1000)
    ap.add_argument("--filter-haplotypes", action="store_true", help="pass --filterHaplotypes to stage 3 (all arms)")
    ap.add_argument("--pooled", action="store_true", help="--doPooled instead of --doDiploid (Bayes EM path, configs[2]); "
                    "glf compared, no VCF (mergeOutputPooled.py is not restated)")
    ap.add_argument("--lib-file", action="store_true", help="pass --libFile <stage-1 libraries> to stage 3 (all arms): "
                    "insert-size prior + mate rescue of unmapped reads, as in the README pipeline")
    ap.add_argument("--realigned-bam", action="store_true", help="pass --outputRealignedBAM to stage 3 (all arms) and "
                    "compare the per-window realigned BAMs (the gpu arm's CIGARs come from K5, dindel_b200_cigars)")
    ap.add_argument("--heap-neutral", action="store_true", help="also compare with the allocator poisoning freed memory")
    ap.add_argument("--gpu-one-file", action="store_true", help="gpu arm: ONE process over all windows (its batched driver "
                    "uses host threads + one GPU pass per batch) instead of one process per window file")
    ap.add_argument("--gpu-threads", type=int, default=0, help="DINDEL_B200_THREADS of the gpu arm (0 = host cores)")
    ap.add_argument("--gpu-devices", default="", help="DINDEL_B200_DEVICES of the gpu arm: its batched driver shards every batch "
                    "over these CUDA devices through the in-process scheduler (e.g. 0,1)")
    ap.add_argument("--gpu-no-batch", action="store_true", help="gpu arm: the window-at-a-time loop (DINDEL_B200_NO_BATCH)")
    ap.add_argument("--gpu-device-stats", default="", help="DINDEL_B200_DEVICE_STATS of the gpu arm: 1 = the read statistics of every "
                    "called site from K6, 0 = always the reference's host loops (default: by size)")
    ap.add_argument("--read-len", type=int, default=100, help="read length of the synthetic BAM (configs[3]: 250)")
    ap.add_argument("--stage3-args", default="", help="extra arguments for stage 3 of every arm, e.g. '--maxLengthIndel 10 --maxHap 32'")
    ap.add_argument("--host-inference", action="store_true", help="gpu arm: diploidGLF / EM sums on the host (DINDEL_B200_HOST_INFERENCE)")
    a = ap.parse_args()
    arms = a.arms.split(",")
    procs = a.procs or len(os.sched_getaffinity(0))
    prep = prepare(a.work, a.length, a.seed, a.coverage, a.windows_per_file, read_len=a.read_len)
    res = {"length": a.length, "coverage": a.coverage, "n_candidates": prep["n_candidates"],
           "n_windows": prep["n_windows"], "window_files": len(prep["window_files"]), "procs": procs, "gpu_procs": a.gpu_procs,
           "filterHaplotypes": a.filter_haplotypes, "libFile": a.lib_file, "pooled": a.pooled}
    verdict = True
    for tag, env in (("", None), (".hn", HEAP_NEUTRAL)):
        if env and not a.heap_neutral:
            continue
        out, r = {}, {}
        for arm in arms:
            if arm == "check" and env:
                continue
            aenv = dict(env or {})
            if arm != "cpu":
                if a.gpu_threads:
                    aenv["DINDEL_B200_THREADS"] = str(a.gpu_threads)
                if a.gpu_devices:
                    aenv["DINDEL_B200_DEVICES"] = a.gpu_devices
                if a.gpu_no_batch:
                    aenv["DINDEL_B200_NO_BATCH"] = "1"
                if a.host_inference:
                    aenv["DINDEL_B200_HOST_INFERENCE"] = "1"
                if a.gpu_device_stats:
                    aenv["DINDEL_B200_DEVICE_STATS"] = a.gpu_device_stats
            glfs, vcf, wall, n_calls = run_arm(arm, prep, a.work, procs if arm == "cpu" else a.gpu_procs,
                                               extra=(["--filterHaplotypes"] if a.filter_haplotypes else []) +
                                               (["--libFile", prep["libraries"]] if a.lib_file else []) +
                                               (["--outputRealignedBAM"] if a.realigned_bam else []) + a.stage3_args.split(), env=aenv or None, tag=tag,
                                               mode="--doPooled" if a.pooled else "--doDiploid",
                                               one_file=a.gpu_one_file and arm == "gpu")
            ok, err = glf_stats(glfs)
            out[arm] = (glfs, vcf)
            r[arm] = {"stage3_wall_s": round(wall, 3), "vcf_calls": n_calls, "windows_ok": ok, "windows_skipped": err}
            if arm == "gpu":
                r[arm]["inference"] = inference_totals(glfs)
                r[arm]["read_stats"] = read_stats_totals(glfs)
                r[arm]["batched"] = [l.strip() for g in glfs for l in open(g[:-len(".glf.txt")] + ".log", errors="replace")
                                     if l.startswith("dindel_b200 batched:") or l.startswith("dindel_b200 host stages")][:8]
            if arm == "check":
                r[arm]["pairs_compared"], r[arm]["pair_mismatches"] = check_totals(glfs)
                verdict = verdict and r[arm]["pair_mismatches"] == 0 and r[arm]["pairs_compared"] > 0
        if "cpu" in out and "gpu" in out:
            n, first = diff_lines(out["cpu"][0], out["gpu"][0])
            r["glf_identical"], r["glf_diff_lines"], r["glf_diff_first"] = n == 0, n, first
            r["vcf_identical"] = open(out["cpu"][1], "rb").read() == open(out["gpu"][1], "rb").read()
            if a.realigned_bam:
                r["realigned_bam"] = compare_realigned_bams(a.work, "cpu", "gpu", tag)
                verdict = verdict and r["realigned_bam"]["mismatches"] == 0 and r["realigned_bam"]["records"] > 0
        res["heap_neutral" if env else "default"] = r
    # the verdict: identical output with the stray read neutralised when asked for, else as run
    last = res.get("heap_neutral", res["default"])
    verdict = verdict and last.get("glf_identical", True) and last.get("vcf_identical", True)
    res["ok"] = verdict
    print(json.dumps(res))
    return 0 if verdict else 1


if __name__ == "__main__":
    sys.exit(main())
