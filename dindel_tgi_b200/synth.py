"""Seeded synthetic window batches of the BASELINE.json shapes (generator G0).

Pair-level inputs for the likelihood engine: per window a reference segment, candidate haplotypes
(reference + alternates carrying 1-2 indels and 0-1 SNP), and reads sampled from a diploid pair
(or a pool) of those haplotypes with phred-distributed substitution errors, rare read indel
errors, mapping qualities and a mapped start.  Deterministic in (config, seed, n_windows).

  cfg 2: R=100, H=120+-5, nr~Poisson(55), nh~U{2..8}, indels 1-5 bp      (BASELINE configs[1])
  cfg 3: R=100, H~120,    nr=1000,        nh=16, pooled                   (configs[2])
  cfg 4: R=250, H=300+-20, nr~Poisson(40), nh=32, homopolymer-rich, indels 1-10 bp (configs[3])
"""
import numpy as np

from .engine import Batch

SEED0 = 20100620
ACGT = np.frombuffer(b"ACGT", dtype=np.uint8)

CONFIGS = {
    2: dict(R=100, H=120, dH=5, nr=("poisson", 55), nh=("uniform", 2, 8), max_indel=5, hp=False, pooled=False),
    3: dict(R=100, H=120, dH=2, nr=("fixed", 1000), nh=("fixed", 16), max_indel=5, hp=False, pooled=True),
    4: dict(R=250, H=300, dH=20, nr=("poisson", 40), nh=("fixed", 32), max_indel=10, hp=True, pooled=False),
}


def _draw(rng, spec, n):
    if spec[0] == "poisson":
        return np.maximum(rng.poisson(spec[1], n), 2).astype(np.int64)
    if spec[0] == "uniform":
        return rng.integers(spec[1], spec[2] + 1, n).astype(np.int64)
    return np.full(n, spec[1], dtype=np.int64)


def _random_seq(rng, n, hp):
    s = ACGT[rng.integers(0, 4, n)]
    if hp and n > 30:
        # homopolymer runs: geometric length (mean 6, cap 20) covering ~40 % of the bases
        pos = 0
        while pos < n:
            pos += int(rng.geometric(1.0 / 9.0))
            run = min(int(rng.geometric(1.0 / 6.0)), 20)
            if pos + run >= n:
                break
            s[pos:pos + run] = s[pos]
            pos += run
    return s


def _make_alt(rng, ref, max_indel, hp):
    """reference bytes -> alternate haplotype with 1-2 indels in the central third and 0-1 SNP."""
    h = bytearray(ref)
    n = len(h)
    for _ in range(int(rng.integers(1, 3))):
        p = int(rng.integers(n // 3, 2 * n // 3))
        if hp:  # slide into a homopolymer when one is near
            for q in range(p, min(p + 12, n - 2)):
                if h[q] == h[q + 1]:
                    p = q
                    break
        ln = int(rng.integers(1, max_indel + 1))
        if rng.random() < 0.5:
            if hp and rng.random() < 0.5:
                ins = bytes([h[p]]) * ln
            else:
                ins = bytes(ACGT[rng.integers(0, 4, ln)])
            h[p:p] = ins
        else:
            del h[p:p + ln]
    if rng.random() < 0.5:
        p = int(rng.integers(5, len(h) - 5))
        h[p] = int(ACGT[(np.searchsorted(ACGT, h[p]) + int(rng.integers(1, 4))) % 4])
    return bytes(h)


def make_batch(cfg=2, n_windows=100, seed=None, chunk=4096):
    """Generate `n_windows` windows of config `cfg` as an engine.Batch."""
    spec = CONFIGS[cfg]
    rng = np.random.Generator(np.random.MT19937(SEED0 + cfg if seed is None else seed))
    parts = []
    w_done = 0
    while w_done < n_windows:
        nw = min(chunk, n_windows - w_done)
        parts.append(_make_chunk(rng, spec, nw, w_done))
        w_done += nw
    return _concat(parts)


def _make_chunk(rng, spec, nw, w_base):
    R, max_indel, hp = spec["R"], spec["max_indel"], spec["hp"]
    Hs = spec["H"] + rng.integers(-spec["dH"], spec["dH"] + 1, nw)
    nhs = _draw(rng, spec["nh"], nw)
    nrs = _draw(rng, spec["nr"], nw)
    F = R  # flank length on each side of the window
    hap_list, win_hap = [], [0]
    # per window: the two (or pool of) source sequences reads are sampled from
    src_seqs, src_off, win_src = [], [0], []
    read_src = np.empty(int(nrs.sum()), dtype=np.int64)   # index into src_seqs
    read_hlen = np.empty(int(nrs.sum()), dtype=np.int64)  # length of the source haplotype
    rp = 0
    for w in range(nw):
        H = int(Hs[w])
        ctx = _random_seq(rng, H + 2 * F, hp)
        ref = bytes(ctx[F:F + H])
        haps, seen = [ref], {ref}
        tries = 0
        while len(haps) < nhs[w] and tries < 200:
            tries += 1
            alt = _make_alt(rng, ref, max_indel, hp)
            if alt not in seen and len(alt) > max_indel + 12:
                seen.add(alt)
                haps.append(alt)
        hap_list.extend(haps)
        win_hap.append(len(hap_list))
        nr = int(nrs[w])
        if spec["pooled"]:
            freqs = rng.uniform(0.01, 0.5, len(haps))
            freqs[0] = max(freqs[0], 0.5)
            pick = rng.choice(len(haps), size=nr, p=freqs / freqs.sum())
        else:
            dip = rng.integers(0, len(haps), 2)
            pick = dip[rng.integers(0, 2, nr)]
        first_src = len(src_seqs)
        used = {}
        for k in np.unique(pick):
            used[int(k)] = len(src_seqs)
            s = np.concatenate([ctx[:F], np.frombuffer(haps[int(k)], dtype=np.uint8), ctx[F + H:]])
            src_seqs.append(s)
            src_off.append(src_off[-1] + len(s))
        win_src.append(first_src)
        read_src[rp:rp + nr] = [used[int(k)] for k in pick]
        read_hlen[rp:rp + nr] = [len(haps[int(k)]) for k in pick]
        rp += nr
    nreads = rp
    src_all = np.concatenate(src_seqs)
    src_off = np.asarray(src_off, dtype=np.int64)
    # start of the read relative to the window's first base, overlap >= 20 bp with the haplotype
    lo = -(R - 20)
    hi = read_hlen - 20
    start_rel = lo + (rng.random(nreads) * (hi - lo + 1)).astype(np.int64)
    idx = (src_off[read_src] + F + start_rel)[:, None] + np.arange(R)[None, :]
    seq = src_all[idx]
    # base qualities: per-read level + per-base noise, decaying 0.1/base toward the 3' end
    level = rng.normal(31.0, 4.0, nreads).astype(np.float32)
    q = level[:, None] + rng.standard_normal((nreads, R), dtype=np.float32) * 6.0 \
        - 0.1 * np.arange(R, dtype=np.float32)[None, :]
    qual = np.clip(np.rint(q), 2, 41).astype(np.uint8)
    # substitution errors with p = 10^(-Q/10)
    perr = np.power(10.0, -qual.astype(np.float32) / 10.0)
    err = rng.random((nreads, R), dtype=np.float32) < perr
    ne = int(err.sum())
    if ne:
        cur = np.searchsorted(ACGT, seq[err])
        seq[err] = ACGT[(cur + rng.integers(1, 4, ne)) % 4]
    # a few 'N' calls at the lowest qualities
    nmask = (qual <= 2) & (rng.random((nreads, R), dtype=np.float32) < 0.3)
    seq[nmask] = ord("N")
    # read indel errors, 5e-4 per base: delete or insert one base, keep the length R
    nind = rng.binomial(nreads * R, 5e-4)
    for _ in range(int(nind)):
        r = int(rng.integers(0, nreads))
        p = int(rng.integers(1, R - 1))
        if rng.random() < 0.5:
            seq[r, p:R - 1] = seq[r, p + 1:R].copy()
        else:
            seq[r, p + 1:R] = seq[r, p:R - 1].copy()
            seq[r, p] = ACGT[rng.integers(0, 4)]
    mapq = rng.choice(np.array([60, 40, 20], dtype=np.uint8), size=nreads, p=[0.85, 0.10, 0.05])
    flags = (rng.random(nreads) < 0.5).astype(np.uint8) * 2
    jitter = np.where(rng.random(nreads) < 0.1, rng.integers(-2, 3, nreads), 0)
    left_pos = (100000 + 400 * (w_base + np.arange(nw))).astype(np.int64)
    win_of_read = np.repeat(np.arange(nw), nrs)
    read_start = (left_pos[win_of_read] + start_rel + jitter).astype(np.uint32)
    hap_off = np.concatenate([[0], np.cumsum([len(h) for h in hap_list])]).astype(np.int64)
    hap_seq = np.frombuffer(b"".join(hap_list), dtype=np.uint8)
    win_read = np.concatenate([[0], np.cumsum(nrs)]).astype(np.int32)
    read_off = (np.arange(nreads + 1, dtype=np.int64) * R)
    return Batch(left_pos.astype(np.uint32), np.asarray(win_hap, dtype=np.int32), win_read, hap_off, hap_seq,
                 read_off, seq.reshape(-1), qual.reshape(-1), mapq, read_start, flags)


def _concat(parts):
    if len(parts) == 1:
        return parts[0]
    hs = np.concatenate([p.hap_start for p in parts])
    wh, wr, ho, ro = [np.zeros(1, np.int64)], [np.zeros(1, np.int64)], [np.zeros(1, np.int64)], [np.zeros(1, np.int64)]
    nh = nr = hb = rb = 0
    for p in parts:
        wh.append(p.win_hap[1:].astype(np.int64) + nh)
        wr.append(p.win_read[1:].astype(np.int64) + nr)
        ho.append(p.hap_off[1:] + hb)
        ro.append(p.read_off[1:] + rb)
        nh += p.n_haps
        nr += p.n_reads
        hb += int(p.hap_off[-1])
        rb += int(p.read_off[-1])
    cat = lambda name: np.concatenate([getattr(p, name) for p in parts])
    return Batch(hs, np.concatenate(wh), np.concatenate(wr), np.concatenate(ho), cat("hap_seq"),
                 np.concatenate(ro), cat("read_seq"), cat("read_qual"), cat("read_mapq"), cat("read_start"),
                 cat("read_flags"))
