#!/usr/bin/env python
"""oracle/e2e/make_windows.py — TEST INFRASTRUCTURE ONLY.

Python-3 restatement of the two pipeline scripts between Dindel's stage 1 (`--analysis getCIGARindels`)
and stage 2 (`--analysis indels`); the reference's scripts are Python 2 only (`print` statements,
`has_key`) and python2 is absent from this image.  Both the CPU run and the GPU-backed run of the
end-to-end check consume the SAME window files written here, so only the likelihood engine differs.

  select_candidates : python/selectCandidates.py:14-64  (sum the counts of identical candidates at a
                      position, keep minCount <= count <= maxCount)
  split_and_merge   : python/makeWindows.py:129-207     (cluster candidate positions closer than minDist)
  write_windows     : python/makeWindows.py:34-125      (window = [min refStart - 60, max refEnd + 60],
                      <= 16 candidates per line, `variantsPerFile` windows per file: the reference's
                      unit of parallelism, python/makeWindows.py:48-55)
One deliberate difference: the reference iterates `set(...)` of a window's candidates in Python 2's
hash order (makeWindows.py:64); here they are written sorted, so the files are reproducible.
"""
import argparse
import sys


class Variant:
    """python/makeWindows.py:5-31"""

    def __init__(self, ref_pos, s):
        if len(s) < 2:
            raise ValueError("Unrecognized variant: " + s)
        if s[0] == "-":
            self.type, self.length = "del", len(s) - 1
            self.ref_start, self.ref_end = ref_pos, ref_pos + self.length - 1
        elif s[0] == "+":
            self.type, self.length = "ins", len(s) - 1
            self.ref_start, self.ref_end = ref_pos, ref_pos - 1
        elif len(s) == 4:
            self.type, self.length = "snp", 1
            self.ref_start, self.ref_end = ref_pos, ref_pos
        else:
            raise ValueError("Unrecognized variant: " + s)
        self.str = s


def select_candidates(in_path, out_path, min_count=2, max_count=10000000):
    cur_chr, cur_pos, counts, seen = None, -1, {}, set()
    with open(in_path) as fi, open(out_path, "w") as fo:
        def flush():
            for v in counts:  # insertion order (the reference: dict order of one position's variants)
                if min_count <= counts[v] <= max_count:
                    fo.write("%s %d %s # %d\n" % (cur_chr, cur_pos, v, counts[v]))
        for line in fi:
            dat = line.split()
            if not dat:
                continue
            chrom, pos = dat[0], int(dat[1])
            if chrom != cur_chr:
                if chrom in seen:
                    raise ValueError("Chromosome already processed; sort by chromosome, then position")
                if cur_chr is not None:
                    flush()
                cur_chr, cur_pos, counts = chrom, -1, {}
                seen.add(chrom)
            if pos < cur_pos:
                raise ValueError("Variants not sorted with respect to position")
            y = dat.index("#")
            if pos > cur_pos:
                if cur_pos != -1:
                    flush()
                cur_pos, counts = pos, {}
            for v, f in zip(dat[2:y], dat[y + 1:]):
                counts[v] = counts.get(v, 0) + int(f)
        if cur_chr is not None and cur_pos != -1:
            flush()


def write_windows(new_variants, chrom, prefix, per_file, start_idx, hap_width=60, max_var=16):
    idx, written, fo, files = start_idx, 10000000, None, []
    for pos in sorted(new_variants):
        if written > per_file:
            idx += 1
            if fo:
                fo.close()
            files.append("%s.%d.txt" % (prefix, idx))
            fo = open(files[-1], "w")
            written = 0
        else:
            written += 1
        vs = [Variant(p, s) for p, s in sorted(set(new_variants[pos]))]
        left = max(0, min(v.ref_start for v in vs) - hap_width)
        right = max(max(v.ref_end for v in vs), 0) + hap_width
        for k in range(0, len(vs), max_var):
            fo.write("%s %d %d" % (chrom, left, right))
            for v in vs[k:k + max_var]:
                fo.write(" %d,%s" % (v.ref_start, v.str))
            fo.write("\n")
    if fo:
        fo.close()
    return idx, files


def split_and_merge(var_path, prefix, min_dist=20, per_file=1000):
    variants, order = {}, []
    with open(var_path) as f:
        for line in f:
            dat = line.split()
            if not dat:
                continue
            chrom, pos = dat[0], int(dat[1])
            if chrom not in variants:
                variants[chrom] = {}
                order.append(chrom)
            lst = variants[chrom].setdefault(pos, [])
            i = 2
            while i < len(dat) and dat[i] != "#":
                lst.append(dat[i])
                i += 1
    start_idx, files = 0, []
    for chrom in order:
        positions = sorted(variants[chrom])
        new_pos = positions[:]
        done = False
        while not done:  # python/makeWindows.py:176-181
            done = True
            for p in range(1, len(positions)):
                if new_pos[p] != new_pos[p - 1] and new_pos[p] - positions[p - 1] <= min_dist:
                    new_pos[p] = new_pos[p - 1]
                    done = False
        clustered = {}
        for p, pos in enumerate(positions):
            for var in variants[chrom][pos]:
                clustered.setdefault(new_pos[p], []).append((pos, var))
        start_idx, fs = write_windows(clustered, chrom, prefix, per_file, start_idx)
        files += fs
    return files


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("-i", "--inputVarFile", required=True)
    ap.add_argument("-w", "--windowFilePrefix", required=True)
    ap.add_argument("-m", "--minDist", type=int, default=20)
    ap.add_argument("-n", "--numWindowsPerFile", type=int, default=1000)
    ap.add_argument("--minCount", type=int, default=0, help="run selectCandidates first (0 = skip)")
    a = ap.parse_args()
    src = a.inputVarFile
    if a.minCount > 0:
        src = a.windowFilePrefix + ".selected.txt"
        select_candidates(a.inputVarFile, src, a.minCount)
    files = split_and_merge(src, a.windowFilePrefix, a.minDist, a.numWindowsPerFile)
    sys.stderr.write("make_windows: %d window files\n" % len(files))


if __name__ == "__main__":
    main()
