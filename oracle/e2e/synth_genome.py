#!/usr/bin/env python
"""oracle/e2e/synth_genome.py — TEST INFRASTRUCTURE ONLY.

Generator G1 of SURVEY.md 8(d) / BASELINE.json configs[0]: a random "chr20" FASTA, one diploid sample
with ~1 indel / 2 kb and ~1 SNP / kb, 30x paired 100-bp Illumina-like reads (insert ~ N(300,30)),
written as a coordinate-sorted SAM with gapped-aligner-style CIGARs (indels in the read interior are
reported, indels within `clip` bases of a read end are soft-clipped the way a local aligner would).
oracle/_ref/bamtool turns the SAM into an indexed BAM.  Seeded and deterministic.

    python oracle/e2e/synth_genome.py --out DIR --length 1000000 --coverage 30 --seed 1
writes DIR/ref.fa, DIR/sample.sam, DIR/truth.txt
"""
import argparse
import os

import numpy as np

ACGT = np.frombuffer(b"ACGT", dtype=np.uint8)
COMP = np.zeros(256, dtype=np.uint8)
for a, b in zip(b"ACGTN", b"TGCAN"):
    COMP[a] = b


def make_reference(rng, L, hp_frac=0.08):
    ref = ACGT[rng.integers(0, 4, L)]
    # homopolymer runs (length 4..12) over ~hp_frac of the bases: the indel-error table matters there
    n_runs = int(L * hp_frac / 7)
    for s in rng.integers(0, L - 16, n_runs):
        ref[s:s + int(rng.integers(4, 13))] = ACGT[rng.integers(0, 4)]
    return ref


def make_variants(rng, ref, indel_rate, snp_rate, margin=400):
    """Non-overlapping variants, >= 25 bp apart; each with genotype (on hap 0, on hap 1)."""
    L = len(ref)
    n = int(L * (indel_rate + snp_rate))
    pos = np.sort(rng.choice(np.arange(margin, L - margin), size=n, replace=False))
    out, last = [], -100
    for p in pos:
        if p - last < 25:
            continue
        p = int(p)
        gt = [(1, 0), (0, 1), (1, 1)][int(rng.integers(0, 3))]
        if rng.random() < indel_rate / (indel_rate + snp_rate):
            ln = int(rng.integers(1, 6))
            if rng.random() < 0.5:
                out.append(("del", p, ln, b"", gt))
                last = p + ln
            else:
                if rng.random() < 0.5:  # tandem duplication of the following bases (typical insertion)
                    ins = bytes(ref[p:p + ln])
                else:
                    ins = bytes(ACGT[rng.integers(0, 4, ln)])
                out.append(("ins", p, ln, ins, gt))
                last = p
        else:
            alt = ACGT[(int(np.where(ACGT == ref[p])[0][0]) + int(rng.integers(1, 4))) % 4]
            out.append(("snp", p, 1, bytes([alt]), gt))
            last = p
    return out


def build_haplotype(ref, variants, which):
    """-> (hap bases, ref position of every hap base or -1 for inserted bases)."""
    seq, rpos, cur = [], [], 0
    for kind, p, ln, alt, gt in variants:
        if not gt[which]:
            continue
        seq.append(ref[cur:p]); rpos.append(np.arange(cur, p))
        if kind == "del":       # ref[p:p+ln] removed
            cur = p + ln
        elif kind == "ins":     # alt inserted before ref[p]
            seq.append(np.frombuffer(alt, dtype=np.uint8)); rpos.append(np.full(ln, -1))
            cur = p
        else:
            seq.append(np.frombuffer(alt, dtype=np.uint8)); rpos.append(np.array([p]))
            cur = p + 1
    seq.append(ref[cur:]); rpos.append(np.arange(cur, len(ref)))
    return np.concatenate(seq), np.concatenate(rpos)


def cigar_of(rp, clip):
    """rp: ref position per read base (-1 = inserted).  Returns (pos0, cigar string) or None."""
    R = len(rp)
    ops = []  # [op, len]
    prev = -1
    for k in range(R):
        p = int(rp[k])
        if p < 0:
            if ops and ops[-1][0] == "I":
                ops[-1][1] += 1
            else:
                ops.append(["I", 1])
        else:
            if prev >= 0 and p - prev > 1:
                ops.append(["D", p - prev - 1])
            if ops and ops[-1][0] == "M":
                ops[-1][1] += 1
            else:
                ops.append(["M", 1])
            prev = p
    # aligner-style ends: an indel closer than `clip` read bases to an end is soft-clipped away
    def clip_left(ops):
        changed = True
        lead = 0
        while changed and ops:
            changed = False
            if ops[0][0] == "I":
                lead += ops.pop(0)[1]; changed = True
            elif ops[0][0] == "D":
                ops.pop(0); changed = True
            elif ops[0][0] == "M" and len(ops) > 1 and ops[0][1] + lead < clip:
                lead += ops.pop(0)[1]; changed = True
        return lead
    lead = clip_left(ops)
    ops.reverse()
    trail = clip_left(ops)
    ops.reverse()
    if not ops or sum(l for o, l in ops if o == "M") < 30:
        return None
    first_m = lead  # index of the first aligned read base
    while rp[first_m] < 0:
        first_m += 1
    pos0 = int(rp[first_m])
    cig = ("%dS" % lead if lead else "") + "".join("%d%s" % (l, o) for o, l in ops) + ("%dS" % trail if trail else "")
    return pos0, cig


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", required=True)
    ap.add_argument("--length", type=int, default=1000000)
    ap.add_argument("--coverage", type=float, default=30.0)
    ap.add_argument("--read-len", type=int, default=100)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--chrom", default="20")
    ap.add_argument("--indel-rate", type=float, default=1 / 2000.0)
    ap.add_argument("--snp-rate", type=float, default=1 / 1000.0)
    ap.add_argument("--clip", type=int, default=6)
    a = ap.parse_args()
    os.makedirs(a.out, exist_ok=True)
    rng = np.random.default_rng(20100620 + a.seed)
    L, R = a.length, a.read_len
    ref = make_reference(rng, L)
    variants = make_variants(rng, ref, a.indel_rate, a.snp_rate)
    haps = [build_haplotype(ref, variants, w) for w in (0, 1)]

    with open(os.path.join(a.out, "ref.fa"), "w") as f:
        f.write(">%s\n" % a.chrom)
        s = ref.tobytes().decode()
        for i in range(0, L, 60):
            f.write(s[i:i + 60] + "\n")
    with open(os.path.join(a.out, "truth.txt"), "w") as f:
        for kind, p, ln, alt, gt in variants:
            f.write("%s %d %s %d %s %d|%d\n" % (a.chrom, p, kind, ln, alt.decode() or ".", gt[0], gt[1]))

    n_pairs = int(L * a.coverage / (2 * R))
    recs = []
    # base-quality profile: Q ~ clipped N(30, 8) decaying 0.1 per cycle toward the 3' end
    cyc = np.arange(R)
    for i in range(n_pairs):
        w = int(rng.integers(0, 2))
        hseq, hrp = haps[w]
        isz = max(R + 10, int(rng.normal(300, 30)))
        st = int(rng.integers(0, len(hseq) - isz))
        mq = (60, 40, 20)[int(np.searchsorted((0.85, 0.95), rng.random()))]
        mates = []
        for first in (True, False):
            s0 = st if first else st + isz - R
            reverse = not first
            bases = hseq[s0:s0 + R].copy()
            q = np.clip(rng.normal(30, 8, R) - 0.1 * (cyc if not reverse else cyc[::-1]), 2, 41).astype(np.int64)
            err = rng.random(R) < 10.0 ** (-q / 10.0)
            if err.any():
                idx = np.where(err)[0]
                cur = np.searchsorted(ACGT, bases[idx])  # ACGT is sorted
                bases[idx] = ACGT[(cur + rng.integers(1, 4, len(idx))) % 4]
            rp = hrp[s0:s0 + R].copy()
            # read-level indel error (p = 5e-4 per base): drop or duplicate one base
            if rng.random() < 5e-4 * R:
                k = int(rng.integers(10, R - 10))
                if rng.random() < 0.5 and s0 + R < len(hseq):
                    bases = np.concatenate([bases[:k], bases[k + 1:], hseq[s0 + R:s0 + R + 1]])
                    rp = np.concatenate([rp[:k], rp[k + 1:], hrp[s0 + R:s0 + R + 1]])
                else:
                    bases = np.concatenate([bases[:k], bases[k:k + 1], bases[k:R - 1]])
                    rp = np.concatenate([rp[:k], [-1], rp[k:R - 1]])
            cg = cigar_of(rp, a.clip)
            mates.append((cg, bases, q, reverse))
        if mates[0][0] is None or mates[1][0] is None:
            continue
        for m, (cg, bases, q, reverse) in enumerate(mates):
            other = mates[1 - m]
            flag = 0x1 | 0x2 | (0x10 if reverse else 0) | (0x20 if other[3] else 0) | (0x40 if m == 0 else 0x80)
            tl = mates[1][0][0] + R - mates[0][0][0]
            tlen = tl if m == 0 else -tl
            recs.append((cg[0], "r%07d" % i, flag, cg[1], mq, other[0][0], tlen,
                         bases.tobytes().decode(), (q + 33).astype(np.uint8).tobytes().decode()))
    recs.sort(key=lambda r: (r[0], r[1], r[2]))
    with open(os.path.join(a.out, "sample.sam"), "w") as f:
        f.write("@HD\tVN:1.0\tSO:coordinate\n@SQ\tSN:%s\tLN:%d\n" % (a.chrom, L))
        for pos0, name, flag, cig, mq, mpos0, tlen, seq, qual in recs:
            f.write("%s\t%d\t%s\t%d\t%d\t%s\t=\t%d\t%d\t%s\t%s\n" %
                    (name, flag, a.chrom, pos0 + 1, mq, cig, mpos0 + 1, tlen, seq, qual))
    print("synth_genome: %d bp, %d variants, %d reads -> %s" % (L, len(variants), len(recs), a.out))


if __name__ == "__main__":
    main()
