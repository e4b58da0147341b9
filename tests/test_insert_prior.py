"""a8 with the insert-size prior (ObservationModelFB.cpp:279-293; active under --libFile, DInDel.cpp:4268-4272).

  * CPU: the oracle restatement equals the reference's own class (oracle/_ref) bit for bit on pairs of paired
    reads with a mapped mate, both mate orientations, a peaked library histogram, mates placed so that the prior
    moves the alignment decision (on / off haplotype) for some pairs.
  * GPU: dindel_b200_submit with read_ins_mode / read_ins_anchor / read_lib and dindel_b200_set_libraries equals
    the oracle bit for bit (tiled and generic kernels).
"""
import numpy as np
import pytest

from dindel_tgi_b200 import engine, synth
from oracle.pyoracle import OraParams
from tests import util

CLI = (5e-4, 1e-5, 100.0, 0.01, 0.95, 5, 2, -1, 0)


def _histogram():
    d = np.arange(1200, dtype=np.float64)
    return np.round(2000.0 * np.exp(-0.5 * ((d - 300.0) / 30.0) ** 2)) + (d % 7 == 0)


def _cases(batch, rng):
    """per read: (mode, anchor) with anchors around where a 300-bp insert would put the mate"""
    modes = rng.integers(0, 3, batch.n_reads).astype(np.uint8)
    anchors = np.zeros(batch.n_reads, dtype=np.int32)
    for w in range(batch.n_windows):
        for r in range(int(batch.win_read[w]), int(batch.win_read[w + 1])):
            start = int(batch.read_start[r])
            off = int(rng.choice([300, 300, 280, 340, 900, 5000, -250]))
            anchors[r] = start + off if modes[r] == 1 else start - off + 100
    return modes, anchors


def test_oracle_matches_reference_with_insert_prior(port_oracle, ref_oracle):
    rng = np.random.default_rng(5)
    counts = _histogram()
    probs, p95 = ref_oracle.library(counts)
    assert 0 < len(probs) <= len(counts) and abs(probs.sum() - 1.0) < 1e-6 and p95 > 0
    batch = synth.make_batch(2, 3, seed=321)
    modes, anchors = _cases(batch, rng)
    p = OraParams(*CLI)
    n = moved = 0
    for w in range(batch.n_windows):
        haps, reads, hs = batch.window(w)
        for r, (seq, qual, mapq, start, flags) in enumerate(reads[:14]):
            gr = int(batch.win_read[w]) + r
            q = np.array([port_oracle.phred_to_prob(x) for x in qual])
            mq = port_oracle.phred_to_prob(mapq)
            for hap in haps[:4]:
                a, ha, ma = port_oracle.pair_ins(hap, seq, q, mq, float(start), flags & 1, hs, p, modes[gr], anchors[gr], probs=probs, p95=p95)
                b, hb, mb = ref_oracle.pair_ins(hap, seq, q, mq, float(start), flags & 1, hs, p, modes[gr], anchors[gr], counts=counts)
                for f in util.CMP_FIELDS:
                    assert a[f] == b[f] or (np.isnan(a[f]) and np.isnan(b[f])), (f, a[f], b[f], modes[gr], anchors[gr])
                assert np.array_equal(ha, hb) and np.array_equal(ma, mb)
                base, _, _ = port_oracle.pair(hap, seq, q, mq, float(start), flags & 1, hs, p)
                moved += (base["ll"] != a["ll"])
                n += 1
    assert n >= 100 and moved >= n // 3


@pytest.mark.gpu
@pytest.mark.parametrize("force_generic", [False, True])
def test_gpu_insert_prior_matches_oracle(port_oracle, ref_oracle, monkeypatch, force_generic):
    if force_generic:
        monkeypatch.setenv("DINDEL_B200_FORCE_GENERIC", "1")
    rng = np.random.default_rng(6)
    counts = _histogram()
    probs, p95 = ref_oracle.library(counts)
    batch = synth.make_batch(2, 6, seed=654)
    modes, anchors = _cases(batch, rng)
    libs = rng.integers(0, 2, batch.n_reads).astype(np.int32)
    probs2 = np.full(50, 1.0 / 50)                       # a second, flat and short library (clamp to maxins-1)
    e = engine.Engine(engine.Params(*CLI), 0)
    e.set_libraries([probs, probs2], [p95, 1.0 / 50])
    res = e.compute_likelihoods(batch, want_hpos=True, ins=(modes, anchors, libs))
    p = OraParams(*CLI)
    moved = 0
    plain = e.compute_likelihoods(batch, want_hpos=False)
    for w in range(batch.n_windows):
        haps, reads, hs = batch.window(w)
        nh = len(haps)
        for r, (seq, qual, mapq, start, flags) in enumerate(reads):
            gr = int(batch.win_read[w]) + r
            q = np.array([port_oracle.phred_to_prob(x) for x in qual])
            pr, pp = (probs, p95) if libs[gr] == 0 else (probs2, 1.0 / 50)
            for h, hap in enumerate(haps):
                want, whp, _ = port_oracle.pair_ins(hap, seq, q, port_oracle.phred_to_prob(mapq), float(start), flags & 1, hs, p,
                                                    modes[gr], anchors[gr], probs=pr, p95=pp)
                i = int(res.win_pair[w]) + r * nh + h
                got = res.pairs[i]
                for f in ("ll", "llOn", "llOff", "mLogBQ", "firstBase", "lastBase", "numIndels", "offHap", "offHapHMQ"):
                    assert got[f] == want[f], (w, r, h, f, got[f], want[f])
                assert np.array_equal(res.hpos[i, :len(seq)], whp)
                moved += got["ll"] != plain.pairs[i]["ll"]
    assert moved > 100
    e.close()
