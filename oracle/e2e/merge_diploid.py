#!/usr/bin/env python
"""oracle/e2e/merge_diploid.py — TEST INFRASTRUCTURE ONLY.

Python-3 restatement of the reference's stage 4, python/mergeOutputDiploid.py (Python 2 only), with the
helpers it imports: utils/Fasta.py:30-53 (indexed fetch), utils/AnalyzeSequence.py:5-19 (homopolymer
length), utils/Variant.py:2-30 (variant string classes), utils/FileUtils.py:4-80 (space-joined table
with a header line).  `.glf.txt` files -> one VCF 4.0 file.  Both arms of the end-to-end check (CPU
reference, GPU-backed reference) pass through this same code, so the VCFs differ iff the glf files do.

  get_vcf_string      : mergeOutputDiploid.py:35-155
  process_glf_file    : mergeOutputDiploid.py:158-238   (rows `ok` + `dip.map` + was_candidate_in_window,
                        int(float(qual)) >= 1, coverage from var_coverage_forward/reverse)
  merge_output        : mergeOutputDiploid.py:240-318   (header text, chromosome order 1..22,X,Y,others)
"""
import argparse
import os
import sys


class Fasta:
    def __init__(self, fname):
        self.fa = open(fname, "rb")
        self.ft = {}
        with open(fname + ".fai") as f:
            for line in f:
                d = line.split()
                if len(d) == 5:
                    self.ft[d[0]] = (int(d[1]), int(d[2]), int(d[3]), int(d[4]))

    def get(self, tid, pos1, n):
        _, offset, blen, llen = self.ft[tid]
        pos = int(pos1) - 1
        self.fa.seek(offset + (pos // blen) * llen + pos % blen, 0)
        seq = []
        while len(seq) < n:
            c = self.fa.read(1)
            if c == b"":
                break
            if c != b"\n":
                seq.append(c.decode())
        return seq


def homopolymer_length(seq, pos):
    n = 1
    for i in range(pos + 1, len(seq)):
        if seq[i] == seq[i - 1]:
            n += 1
        else:
            break
    for i in range(pos - 1, 0, -1):
        if seq[i] == seq[i + 1]:
            n += 1
        else:
            break
    return n


def variant_of(s):
    """-> (type, seq, length)   utils/Variant.py:3-24"""
    if s[0] == "-" and len(s) > 1:
        return "del", s[1:], len(s) - 1
    if s[0] == "+" and len(s) > 1:
        return "ins", s[1:], len(s) - 1
    if len(s) == 4 and s[1:3] == "=>":
        return "snp", s[3], 1
    if s[0] == "*" or "REF" in s or "ref" in s:
        return "ref", "", 0
    raise ValueError("Unrecognized variant: " + s)


def get_vcf_string(glf, fa, max_hp_len=10, filter_qual=0):
    filters = []
    pos, chrom = int(glf["pos"]), glf["chr"]
    hplen = homopolymer_length(fa.get(chrom, pos + 1 - 25, 50), 25)
    report_pos = pos
    max_del = 0
    for a in set(glf["nref_all"]):
        t, _, ln = variant_of(a)
        if t == "del" and ln > max_del:
            max_del = ln
    refseq = "".join(fa.get(chrom, report_pos, 1 + max_del))
    altseqs, alt_type = [], {}
    for a in glf["nref_all"]:
        t, sq, ln = variant_of(a)
        if t == "del":
            alt = refseq[0] + refseq[1 + ln:]
        elif t == "ins":
            alt = refseq[0] + sq + refseq[1:]
        elif t == "snp":
            alt = refseq[0] + sq[0] + refseq[2:]
        else:
            continue
        if alt not in altseqs:
            altseqs.append(alt)
            alt_type[alt] = t
    gtd = glf["genotype"].split(":")
    rec_gt = "%s:%d" % (gtd[0], int(float(gtd[1])))
    if all(alt_type[a] == "snp" for a in altseqs):
        report_pos += 1
        refseq = "".join(fa.get(chrom, report_pos, 1))
        altseqs = [a[1:] for a in altseqs]
    if hplen > max_hp_len:
        filters.append("hp%d" % max_hp_len)
    if glf["qual"] < filter_qual:
        filters.append("q%d" % filter_qual)
    altseqs = ["<DEL>" if "D" in a else a for a in altseqs]
    info = "DP=%d;NF=%d;NR=%d;NRS=%d;NFS=%d;HP=%d" % (
        int(glf["num_hap_reads"]), glf["num_cover_forward"], glf["num_cover_reverse"],
        glf["num_cover_forward_old"], glf["num_cover_reverse_old"], hplen)
    line = "%s\t%s\t.\t%s\t%s\t%s\t%s\t%s\t%s\t%s" % (
        chrom, report_pos, refseq, ",".join(altseqs), glf["qual"], ";".join(filters) if filters else "PASS",
        info, "GT:GQ", rec_gt)
    return line, report_pos


def process_glf_file(path, variants, fa, max_hp_len, filter_qual):
    with open(path) as f:
        labels = f.readline().rstrip("\n").rstrip().split(" ")
        for raw in f:
            cols = raw.rstrip("\n").split(" ")
            dat = dict(zip(labels, cols))
            if dat.get("msg") != "ok" or dat.get("analysis_type") != "dip.map":
                continue
            if dat["was_candidate_in_window"] != "1":
                continue
            glf = {"chr": dat["tid"], "pos": dat["realigned_position"], "qual": int(float(dat["qual"]))}
            if float(glf["qual"]) < 1.0:
                continue
            glf["nref_all"] = dat["nref_all"].split(",")
            if glf["nref_all"] == ["R=>D"]:
                continue
            glf["num_cover_forward"] = int(dat["var_coverage_forward"].split(",")[0])
            glf["num_cover_reverse"] = int(dat["var_coverage_reverse"].split(",")[0])
            glf["num_cover_forward_old"] = int(dat["num_cover_forward"])
            glf["num_cover_reverse_old"] = int(dat["num_cover_reverse"])
            glf["num_hap_reads"] = dat["num_reads"]
            glf["genotype"] = dat["glf"]
            line, rpos = get_vcf_string(glf, fa, max_hp_len, filter_qual)
            variants.setdefault(dat["tid"], {}).setdefault(rpos, []).append(line)


def merge_output(glf_files, ref_file, vcf_file, sample_id="SAMPLE", max_hp_len=10, filter_qual=20):
    fa = Fasta(ref_file)
    variants = {}
    for g in glf_files:
        if not os.path.exists(g):
            sys.stderr.write("File %s does not exist\n" % g)
            sys.exit(1)
        process_glf_file(g, variants, fa, max_hp_len, filter_qual)
    with open(vcf_file, "w") as fv:
        fv.write("##fileformat=VCFv4.0\n##source=Dindel\n##reference=%s\n" % os.path.basename(ref_file))
        fv.write('##INFO=<ID=DP,Number=1,Type=Integer,Description="Total number of reads in haplotype window">\n')
        fv.write('##INFO=<ID=HP,Number=1,Type=Integer,Description="Reference homopolymer tract length">\n')
        fv.write('##INFO=<ID=NF,Number=1,Type=Integer,Description="Number of reads covering non-ref variant on forward strand">\n')
        fv.write('##INFO=<ID=NR,Number=1,Type=Integer,Description="Number of reads covering non-ref variant on reverse strand">\n')
        fv.write('##INFO=<ID=NFS,Number=1,Type=Integer,Description="Number of reads covering non-ref variant site on forward strand">\n')
        fv.write('##INFO=<ID=NRS,Number=1,Type=Integer,Description="Number of reads covering non-ref variant site on reverse strand">\n')
        fv.write('##FORMAT=<ID=GT,Number=1,Type=String,Description="Genotype">\n')
        fv.write('##FORMAT=<ID=GQ,Number=1,Type=Integer,Description="Genotype quality">\n')
        fv.write('##ALT=<ID=DEL,Description="Deletion">\n')
        fv.write('##FILTER=<ID=q%d,Description="Quality below %d">\n' % (filter_qual, filter_qual))
        fv.write('##FILTER=<ID=hp%d,Description="Reference homopolymer length was longer than %d">\n' % (max_hp_len, max_hp_len))
        fv.write('##FILTER=<ID=fr0,Description="Non-ref allele is not covered by at least one read on both strands">\n')
        fv.write('##FILTER=<ID=wv,Description="Other indel in window had higher likelihood">\n')
        fv.write("#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\tFORMAT\t%s\n" % sample_id)
        order = [str(v) for v in range(1, 23)] + ["X", "Y"]
        order += [c for c in variants if c not in order]
        n = 0
        for c in order:
            for pos in sorted(variants.get(c, {})):
                for line in variants[c][pos]:
                    fv.write(line + "\n")
                    n += 1
    return n


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("-i", "--inputFiles", required=True, help="file listing the .glf.txt files")
    ap.add_argument("-o", "--outputFile", required=True)
    ap.add_argument("-r", "--refFile", required=True)
    ap.add_argument("-s", "--sampleID", default="SAMPLE")
    ap.add_argument("--maxHPLen", type=int, default=10)
    ap.add_argument("-f", "--filterQual", type=int, default=20)
    a = ap.parse_args()
    files = [l.split()[0] for l in open(a.inputFiles) if l.strip()]
    n = merge_output(files, a.refFile, a.outputFile, a.sampleID, a.maxHPLen, a.filterQual)
    sys.stdout.write("merge_diploid: %d calls -> %s\n" % (n, a.outputFile))


if __name__ == "__main__":
    main()
