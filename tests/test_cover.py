"""a13 / K4: DetInDel::filterHaplotypes (DInDel.cpp:1932-2100).

  * CPU: the oracle restatement (oracle/dindel_oracle.c:ora_filter_haplotypes) reproduces the reference's OWN
    filterHaplotypes on golden windows dumped from the stock CPU driver (tests/golden/cover_golden.json.gz,
    generator tests/golden/make_cover_golden.py): filtered[] and every (nf, nr) of varCoverage, with and
    without --filterHaplotypes.  Plus hand-made cases for the `break` rule and the 'N' exemption.
  * GPU: dindel_b200_coverage (K4) over the resident K1 results equals the oracle run on the very hpos /
    offHapHMQ / numIndels the GPU produced, on synthetic windows with random variant entries (edge cases:
    empty and negative intervals, intervals past the haplotype end, OTHER entries, keys shared by haplotypes),
    with the hpos rows kept in HBM only (DINDEL_OUT_KEEP_HPOS) or also downloaded.  Bit-exact (integer work).
"""
import numpy as np
import pytest

from tests import util


def test_oracle_matches_reference_golden(port_oracle):
    wins = util.load_cover_golden()
    assert len(wins) >= 8 and any(w["doFilter"] for w in wins) and any(w["filtered"].any() for w in wins)
    n_cov = 0
    for w in wins:
        f, u, nf, nr = port_oracle.filter_haplotypes(w["haps"], w["reads"], w["hpos"], w["ohq"], w["ni"], w["strand"],
                                                     w["entries"], w["nk"], w["maxMismatch"], w["doFilter"])
        assert (f == w["filtered"]).all()
        assert set(w["cov"]) == set(w["keys"])
        for i, k in enumerate(w["keys"]):
            assert (int(nf[i]), int(nr[i])) == w["cov"][k], (k, nf[i], nr[i], w["cov"][k])
            n_cov += (nf[i] + nr[i]) > 0
    assert n_cov >= 10


def _toy(port_oracle, entries, nk, hap=b"ACGTACGTACGTACGTACGT", reads=None, do_filter=1, mm=2):
    reads = reads or [hap[2:18], hap[0:12]]
    nh, nr = 1, len(reads)
    stride = max(len(r) for r in reads)
    hpos = np.full((nr * nh, stride), -32768, dtype=np.int32)
    starts = [2, 0]
    for r, rd in enumerate(reads):
        hpos[r, :len(rd)] = np.arange(starts[r], starts[r] + len(rd))
    return port_oracle.filter_haplotypes([hap], reads, hpos, np.zeros(nr, np.uint8), np.zeros(nr, np.int32),
                                         np.array([0, 1], np.uint8), entries, nk, mm, do_filter)


def test_oracle_break_and_n_exemption(port_oracle):
    # entry 0 covered by both reads (one per strand); entry 1 (interval beyond read 1, inside read 0)
    f, u, nf, nr = _toy(port_oracle, [(0, 0, 1, 4, 9), (0, 1, 0, 10, 16)], 2)
    assert list(f) == [0] and (nf[0], nr[0]) == (1, 1) and (nf[1], nr[1]) == (1, 0)
    # first entry uncovered (interval reaches past every read) -> break: second entry never evaluated, hap filtered
    f, u, nf, nr = _toy(port_oracle, [(0, 0, 1, 1, 19), (0, 1, 0, 4, 9)], 2)
    assert list(f) == [1] and list(u) == [1] and (nf[1], nr[1]) == (0, 0)
    f, u, nf, nr = _toy(port_oracle, [(0, 0, 1, 1, 19), (0, 1, 0, 4, 9)], 2, do_filter=0)
    assert list(f) == [0] and list(u) == [1] and (nf[1], nr[1]) == (0, 0)
    # mismatches: read 0 with 3 mismatches inside the interval fails at maxMismatch 2; hap 'N' is exempt for DEL only
    hap = b"ACGTNNNTACGTACGTACGT"
    rd = bytearray(hap[2:18]); rd[2:5] = b"CCC"      # read bases over hap[4:7] = NNN
    f, u, nf, nr = _toy(port_oracle, [(0, 0, 0, 3, 9)], 1, hap=hap, reads=[bytes(rd), hap[0:12]])
    assert (nf[0], nr[0]) == (1, 1)                     # DEL: N exempt, read 1 matches N with N
    f, u, nf, nr = _toy(port_oracle, [(0, 0, 1, 3, 9)], 1, hap=hap, reads=[bytes(rd), hap[0:12]])
    assert (nf[0], nr[0]) == (0, 1)                     # INS: 3 mismatches > 2


def _random_entries(rng, batch, w, per_hap=None):
    h0, h1 = int(batch.win_hap[w]), int(batch.win_hap[w + 1])
    nk = int(rng.integers(1, 6)) if per_hap is None else int(rng.integers(20, 60))
    ents = []
    for h in range(h1 - h0):
        H = int(batch.hap_off[h0 + h + 1] - batch.hap_off[h0 + h])
        for _ in range(int(rng.integers(0, 4)) if per_hap is None else per_hap):
            ty = int(rng.choice([0, 1, 1, 0, 2]))
            kind = rng.random()
            if kind < 0.75:
                left = int(rng.integers(20, max(21, H - 30))); ln = int(rng.integers(2, 14))
            elif kind < 0.85:
                left = int(rng.integers(-4, 3)); ln = int(rng.integers(1, 8))        # negative / at the start
            elif kind < 0.95:
                left = H - int(rng.integers(1, 6)); ln = int(rng.integers(2, 10))    # runs past the haplotype
            else:
                left = int(rng.integers(20, 60)); ln = int(rng.integers(-2, 1))      # empty interval
            ents.append((h, int(rng.integers(0, nk)), ty, left, left + ln - 1))
    return ents, nk


@pytest.mark.gpu
@pytest.mark.parametrize("cfg,nwin,keep_only,do_filter", [(2, 40, True, 1), (2, 25, False, 0), (3, 2, True, 1), (4, 2, True, 1)])
def test_k4_matches_oracle(engine_cli, port_oracle, cfg, nwin, keep_only, do_filter):
    from dindel_tgi_b200 import synth
    rng = np.random.default_rng(1234 + cfg)
    batch = synth.make_batch(cfg, nwin, seed=4242 + cfg)
    res = engine_cli.compute_likelihoods(batch, want_hpos=True)     # hpos on the host for the oracle
    ents, nks = zip(*[_random_entries(rng, batch, w) for w in range(batch.n_windows)])
    strand = rng.integers(0, 2, batch.n_reads).astype(np.uint8)
    if keep_only:  # second submit: rows stay in HBM, nothing but the records crosses PCIe
        res2 = engine_cli.compute_likelihoods(batch, want_hpos=False, keep_hpos=True)
        assert res2.hpos is None and (res2.pairs["ll"] == res.pairs["ll"]).all()
    filt, unc, nf, nr = engine_cli.coverage(list(ents), list(nks), strand, max_mismatch=2, do_filter=do_filter)
    n_hits = n_filt = 0
    for w in range(batch.n_windows):
        haps, reads, _ = batch.window(w)
        h0, r0, p0, p1 = int(batch.win_hap[w]), int(batch.win_read[w]), int(res.win_pair[w]), int(res.win_pair[w + 1])
        of, ou, onf, onr = port_oracle.filter_haplotypes(
            haps, [r[0] for r in reads], res.hpos[p0:p1].astype(np.int32), res.pairs["offHapHMQ"][p0:p1],
            res.pairs["numIndels"][p0:p1].astype(np.int32), strand[r0:r0 + len(reads)], ents[w], nks[w], 2, do_filter)
        assert (filt[h0:h0 + len(haps)] == of).all(), w
        assert (unc[h0:h0 + len(haps)] == ou).all(), w
        assert (nf[w] == onf).all() and (nr[w] == onr).all(), (w, nf[w], onf, nr[w], onr)
        n_hits += int((onf + onr).sum()); n_filt += int(ou.sum())
    assert n_hits > 0 and n_filt > 0


@pytest.mark.gpu
def test_k4_deep_windows_keep_their_bit_sets_in_global_memory(engine_cli, port_oracle):
    """Deep pooled windows (configs[2]: 1000 reads x 16 haplotypes) with many variant entries: the covering-read bit sets
    (entries x 2 x ceil(reads/32) words) of a window no longer fit one CTA's shared memory (960 entries: 245 KB; the call
    used to refuse the whole batch).  Such windows run in a second launch with their sets in a global scratch, in the
    same call as ordinary windows; results against the oracle, window by window."""
    from dindel_tgi_b200 import synth
    rng = np.random.default_rng(99)
    batch = synth.make_batch(3, 4, seed=77)
    res = engine_cli.compute_likelihoods(batch, want_hpos=True)
    ents, nks = zip(*[_random_entries(rng, batch, w, per_hap) for w, per_hap in enumerate((None, 14, 60, None))])
    assert len(ents[2]) * 2 * 32 * 4 > 227 * 1024 and len(ents[1]) * 2 * 32 * 4 > 48 * 1024 and len(ents[0]) < 64
    strand = rng.integers(0, 2, batch.n_reads).astype(np.uint8)
    for do_filter in (1, 0):
        filt, unc, nf, nr = engine_cli.coverage(list(ents), list(nks), strand, max_mismatch=2, do_filter=do_filter)
        hits = 0
        for w in range(batch.n_windows):
            haps, reads, _ = batch.window(w)
            h0, r0, p0, p1 = int(batch.win_hap[w]), int(batch.win_read[w]), int(res.win_pair[w]), int(res.win_pair[w + 1])
            of, ou, onf, onr = port_oracle.filter_haplotypes(
                haps, [r[0] for r in reads], res.hpos[p0:p1].astype(np.int32), res.pairs["offHapHMQ"][p0:p1],
                res.pairs["numIndels"][p0:p1].astype(np.int32), strand[r0:r0 + len(reads)], ents[w], nks[w], 2, do_filter)
            assert (filt[h0:h0 + len(haps)] == of).all() and (unc[h0:h0 + len(haps)] == ou).all(), w
            assert (nf[w] == onf).all() and (nr[w] == onr).all(), w
            hits += int((onf + onr).sum())
        assert hits > 1000


@pytest.mark.gpu
def test_k4_needs_resident_hpos(engine_cli):
    from dindel_tgi_b200 import synth
    batch = synth.make_batch(2, 2, seed=5)
    engine_cli.compute_likelihoods(batch, want_hpos=False)
    with pytest.raises(RuntimeError, match="hpos"):
        engine_cli.coverage([[], []], [0, 0], np.zeros(batch.n_reads, np.uint8))
