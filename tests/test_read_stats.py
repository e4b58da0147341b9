"""K6 (read statistics of a called site) pinned on the reference's OWN loops: tests/golden/stats_golden.json.gz holds,
for 12 windows of a synthetic chr20, what the stock CPU driver's aggregation loops read and computed (dip.map sites
DInDel.cpp:3126-3170, variant positions :3396-3404 + :3504-3554, pooled variants :2697-2742; dumped by
oracle/_ref/dindel_ref_dump, generator tests/golden/make_stats_golden.py).

  CPU   the oracle restatement (oracle/dindel_oracle.c: ora_read_stats) reproduces every dumped counter and every
        dumped double (msq, allmsq, mLogBQ: read-order FP64 sums) BIT FOR BIT
  GPU   the CUDA kernel (csrc/k6_read_stats.cu) through dindel_b200_read_stats does the same on the reference's
        records loaded with dindel_b200_load_records, and equals the oracle bit for bit on K1's own records of
        synthetic batches (unmapped reads, pool members, absent sites, 2 200-read windows: more than one tile)
Integer / byte work and ordered FP64 adds only: the tolerance is zero.
"""
import gzip
import json
import math
import os

import numpy as np
import pytest

from oracle.pyoracle import ORA_DTYPE

HERE = os.path.dirname(os.path.abspath(__file__))
INT_FIELDS = ("n", "nf", "nr", "numOffBoth", "numIndels", "nBQT", "nmmBQT", "nMMLeft", "nMMRight", "numUnmappedRealigned")
DBL_FIELDS = ("msq", "allmsq", "mLogBQ")
PAD = 2  # params.obsParams.padCover as the dindel CLI sets it (DInDel.cpp:4153; engine.cli_params())


@pytest.fixture(scope="module")
def golden():
    with gzip.open(os.path.join(HERE, "golden", "stats_golden.json.gz"), "rt") as f:
        return json.load(f)


def _ora_recs(w):
    recs = np.zeros(len(w["recs"]), dtype=ORA_DTYPE)
    for i, (ll, off, nind, nbqt, nmm, mlog, ml, mr, first, last) in enumerate(w["recs"]):
        recs[i]["ll"], recs[i]["offHap"], recs[i]["numIndels"], recs[i]["nBQT"], recs[i]["nmmBQT"] = ll, off, nind, nbqt, nmm
        recs[i]["mLogBQ"], recs[i]["nMMLeft"], recs[i]["nMMRight"], recs[i]["firstBase"], recs[i]["lastBase"] = mlog, ml, mr, first, last
    return recs


def _check(got, want, what):
    for k, v in want.items():
        if k in DBL_FIELDS:
            assert float(got[k]) == v, (what, k, float(got[k]), v)      # bit for bit
        else:
            assert int(got[k]) == v, (what, k, int(got[k]), v)


def test_golden_covers_the_three_loops(golden):
    modes = [j["mode"] for w in golden for j in w["jobs"]]
    assert modes.count(0) >= 10 and modes.count(1) >= 7 and modes.count(2) >= 20
    site = [j for w in golden for j in w["jobs"] if j["mode"] == 0]
    assert any(j["out"]["nf"] and j["out"]["nr"] for j in site)          # covered on both strands
    assert any(j["out"]["n"] == 0 for j in site)                           # a site no read covers
    assert any(j["out"]["numIndels"] for j in site) and any(j["out"]["nMMLeft"] + j["out"]["nMMRight"] for j in site)
    assert any(not (c[3] & 1) for j in site for c in j["cands"])          # a candidate without the site in its map
    pool = [j for w in golden for j in w["jobs"] if j["mode"] == 2]
    assert any(j["out"]["nf"] + j["out"]["nr"] for j in pool) and any(j["out"]["nf"] + j["out"]["nr"] == 0 for j in pool)


def test_oracle_read_stats_bit_exact_vs_reference(golden, port_oracle):
    n = 0
    for wi, w in enumerate(golden):
        recs = _ora_recs(w)
        mq, rev, unm, pool = (np.array([r[k] for r in w["reads"]]) for k in range(4))
        for ji, j in enumerate(w["jobs"]):
            got = port_oracle.read_stats(j["mode"], recs, w["nh"], mq, rev, unm, j["cands"], PAD,
                                         read_group=pool if j["group"] >= 0 else None, group=j["group"])
            _check(got, j["out"], "window %d job %d mode %d" % (wi, ji, j["mode"]))
            assert int(got["n_reads"]) == (w["nr"] if j["group"] < 0 else int((pool == j["group"]).sum()))
            n += 1
    assert n >= 50


# ---------------------------------------------------------------------------------------------------------------
# GPU
# ---------------------------------------------------------------------------------------------------------------
def _phred(map_qual):
    """Read.hpp:127-131 inverted: the phred byte whose conversion gives exactly this Read::mapQual."""
    ph = int(round(-10.0 * math.log10(1.0 - map_qual))) if map_qual < 1.0 - 1e-15 else 255
    for cand in (ph, ph - 1, ph + 1):
        q = 1.0 - math.pow(10.0, -float(cand) / 10.0)
        q = min(max(q, 1e-16), 1.0 - 1e-16)
        if q == map_qual:
            return cand
    raise AssertionError("no phred byte gives mapQual %r" % map_qual)


@pytest.mark.gpu
def test_gpu_read_stats_bit_exact_vs_reference_golden(golden):
    from dindel_tgi_b200 import engine
    e = engine.Engine(engine.cli_params(), 0)
    recs, nhs, mapq, flags, pools = [], [], [], [], []
    for w in golden:
        r = np.zeros(len(w["recs"]), dtype=engine.PAIR_DTYPE)
        for i, (ll, off, nind, nbqt, nmm, mlog, ml, mr, first, last) in enumerate(w["recs"]):
            r[i]["ll"], r[i]["offHap"], r[i]["numIndels"], r[i]["nBQT"], r[i]["nmmBQT"] = ll, off, nind, nbqt, nmm
            r[i]["mLogBQ"], r[i]["nMMLeft"], r[i]["nMMRight"], r[i]["firstBase"], r[i]["lastBase"] = mlog, ml, mr, first, last
        recs.append(r)
        nhs.append(w["nh"])
        mapq += [_phred(x[0]) for x in w["reads"]]
        flags += [(engine.READ_UNMAPPED if x[2] else 0) | (engine.READ_REVERSE if x[1] else 0) for x in w["reads"]]
        pools += [x[3] for x in w["reads"]]
    e.load_records(recs, nhs, mapq, flags)
    jobs = [(wi, j["mode"], j["group"], j["cands"]) for wi, w in enumerate(golden) for j in w["jobs"]]
    want = [j["out"] for w in golden for j in w["jobs"]]
    for rep in range(2):  # the second call reuses the staging buffers
        got = e.read_stats(jobs, read_group=pools)
        for k, (g, o) in enumerate(zip(got, want)):
            _check(g, o, "job %d" % k)
    e.close()


def _to_ora(pairs):
    o = np.zeros(len(pairs), dtype=ORA_DTYPE)
    for f in ("ll", "llOn", "llOff", "mLogBQ", "firstBase", "lastBase", "numIndels", "numMismatch", "nBQT", "nmmBQT", "nMMLeft",
              "nMMRight", "offHap", "offHapHMQ", "bMid", "status"):
        o[f] = pairs[f]
    return o


@pytest.mark.gpu
@pytest.mark.parametrize("cfg,nw,seed", [(2, 12, 31), (3, 3, 32)])
def test_gpu_read_stats_matches_oracle_on_k1_records(port_oracle, cfg, nw, seed):
    """K1's own records of synthetic batches (cfg 3: 1000 reads x 16 haplotypes per window, tiled x3 below to cross the
    1024-read tile), random jobs of the three modes with pool members, unmapped reads and absent sites."""
    from dindel_tgi_b200 import engine, synth
    rng = np.random.default_rng(seed)
    batch = synth.make_batch(cfg, nw, seed=seed)
    if cfg == 3:  # windows of 2 200 reads: more than one 1024-read tile per job
        wins = []
        for w in range(batch.n_windows):
            haps, reads, hs = batch.window(w)[:3]
            wins.append((hs, haps, reads + reads + reads[:200]))
        batch = engine.Batch.from_windows(wins)
    flags = batch.read_flags.copy()
    flags[rng.random(batch.n_reads) < 0.15] |= engine.READ_UNMAPPED
    flags[rng.random(batch.n_reads) < 0.5] |= engine.READ_REVERSE
    batch.read_flags = flags
    batch.read_mapq = rng.choice(np.array([0, 3, 20, 29, 37, 60, 255], dtype=np.uint8), batch.n_reads)
    e = engine.Engine(engine.cli_params(), 0)
    res = e.compute_likelihoods(batch, want_hpos=False)
    groups = rng.integers(0, 3, batch.n_reads).astype(np.int32)
    jobs, want = [], []
    for w in range(batch.n_windows):
        nh = int(batch.win_hap[w + 1] - batch.win_hap[w])
        r0, r1 = int(batch.win_read[w]), int(batch.win_read[w + 1])
        recs = _to_ora(res.pairs[res.win_pair[w]:res.win_pair[w + 1]])
        mq = np.array([port_oracle.phred_to_prob(int(p)) for p in batch.read_mapq[r0:r1]])
        rev = (flags[r0:r1] & engine.READ_REVERSE) != 0
        unm = (flags[r0:r1] & engine.READ_UNMAPPED) != 0
        H = int(np.min(np.diff(batch.hap_off[batch.win_hap[w]:batch.win_hap[w + 1] + 1])))
        for mode in (0, 1, 2, 0, 2):
            grp = -1 if mode != 2 or rng.random() < 0.3 else int(rng.integers(0, 3))
            hs = list(range(nh)) if mode == 2 else [int(x) for x in rng.integers(0, nh, 2)]
            cands = []
            for h in hs:
                st = int(rng.integers(5, max(6, H - 10)))
                cands.append((h, st, st + int(rng.integers(0, 6)), int(rng.integers(0, 4))))
            if mode != 2 and rng.random() < 0.3:
                cands[1] = (cands[0][0], cands[0][1], cands[0][2], 0)      # the homozygous pair: second candidate skipped
            jobs.append((w, mode, grp, cands))
            want.append(port_oracle.read_stats(mode, recs, nh, mq, rev, unm, cands, PAD, read_group=groups[r0:r1], group=grp))
    got = e.read_stats(jobs, read_group=groups)
    assert got.tobytes() == np.array(want, dtype=got.dtype).tobytes()
    assert got["n"].sum() > 0 and got["numUnmappedRealigned"].sum() > 0 and (got["nf"] > 0).any() and (got["nr"] > 0).any()
    e.close()
