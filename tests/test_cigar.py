"""Realigned-BAM CIGARs (--outputRealignedBAM): the per-read haplotype choice at the two call
sites (DInDel.cpp:503-509 pooled, :597-606 diploid) and DetInDel::getCIGAR (DInDel.cpp:728-882).

  * CPU: the oracle restatement (oracle/dindel_oracle.c: ora_choose_hap, ora_get_cigar) reproduces the reference's
    OWN results on golden reads dumped from the stock CPU driver (tests/golden/cigar_golden.json.gz, generator
    tests/golden/make_cigar_golden.py): chosen haplotype, CIGAR::refPos and every CIGAR operation, --doDiploid and
    --doPooled.  Plus hand-made cases for the branches the synthetic BAM does not reach.
  * GPU: dindel_b200_cigars (K5) over the resident K1 results equals the oracle run on the very ll / hpos the GPU
    produced, with synthetic haplotype-to-reference maps (insertions, deletions, LO / RO flanks).  Bit-exact.
"""
import numpy as np
import pytest

from tests import util

M, I, D, S = 0, 1, 2, 4
INS, LO, RO = -1, -3, -4


def test_oracle_matches_reference_golden(port_oracle):
    recs = util.load_cigar_golden()
    assert len(recs) >= 200
    shapes, n_h2 = set(), 0
    for g in recs:
        mode = 0 if g["mode"] == "D" else 1
        hx = port_oracle.choose_hap(mode, g["ll"], g["nInd"], g["h1"], g["h2"])
        assert hx == g["hmax"], g
        n_h2 += mode == 0 and hx == g["h2"] and g["h1"] != g["h2"]
        n, ref_pos, ops = port_oracle.get_cigar(g["hpos"], g["hml"], g["refStart"])
        want = list(zip(g["cigar"][0::2], g["cigar"][1::2]))
        assert n == len(want) and ops == want, (g, ops)
        if want != [(S, len(g["hpos"]))]:        # refPos of a wholly clipped read is never set by the reference
            assert ref_pos == g["refPos"], g
        assert sum(ln for op, ln in ops if op in (M, I, S)) == len(g["hpos"])
        shapes.add(tuple(op for op, _ in ops))
    assert n_h2 > 5 and {(M,), (S, M), (M, S), (M, I, M), (M, D, M), (S,)} <= shapes


def test_oracle_hand_cases(port_oracle):
    ident = list(range(40))
    # plain match at an offset, reference start added
    assert port_oracle.get_cigar(list(range(5, 15)), ident, 1000) == (1, 1005, [(M, 10)])
    # LO / RO flanks become soft clips
    assert port_oracle.get_cigar([LO, LO, 0, 1, 2, RO], ident, 7) == (3, 7, [(S, 2), (M, 3), (S, 1)])
    # read insertion, read deletion (haplotype positions skipped)
    assert port_oracle.get_cigar([3, 4, INS, INS, 5, 6], ident, 0)[2] == [(M, 2), (I, 2), (M, 2)]
    assert port_oracle.get_cigar([3, 4, 8, 9], ident, 0)[2] == [(M, 2), (D, 3), (M, 2)]
    # insertion followed by a jump on the reference: I then D (:851-859)
    assert port_oracle.get_cigar([3, 4, INS, 8, 9], ident, 0)[2] == [(M, 2), (I, 1), (D, 3), (M, 2)]
    # haplotype carrying a deletion (reference positions 10..14 missing) and an insertion (bases 20, 21 off reference)
    hml = list(range(10)) + list(range(15, 25)) + [INS, INS] + list(range(25, 35))
    n, ref_pos, ops = port_oracle.get_cigar(list(range(6, 30)), hml, 100)
    assert (ref_pos, ops) == (106, [(M, 4), (D, 5), (M, 10), (I, 2), (M, 8)])
    # read starting inside the haplotype's inserted bases: they are clipped (:794-797)
    assert port_oracle.get_cigar(list(range(20, 26)), hml, 0)[1:] == (25, [(S, 2), (M, 4)])
    # wholly off the haplotype / no base on the reference
    assert port_oracle.get_cigar([LO] * 5, ident, 0) == (1, -1, [(S, 5)])
    assert port_oracle.get_cigar([20, 21], hml, 0) == (1, -1, [(S, 2)])
    # capacity: the count is exact, the stored operations are truncated
    n, _, ops = port_oracle.get_cigar([3, 4, INS, 8, 9], ident, 0, cap=2)
    assert n == 4 and ops == [(M, 2), (I, 1)]
    # the haplotype choice: ties within 1e-8 go to the haplotype with fewer indels, else to h2 (:598-599)
    assert port_oracle.choose_hap(0, [-1.0, -1.0 - 1e-9, -3.0], [1, 0, 0], 0, 1) == 1
    assert port_oracle.choose_hap(0, [-1.0, -1.0 - 1e-9, -3.0], [0, 1, 0], 0, 1) == 0
    assert port_oracle.choose_hap(0, [-1.0, -1.0, -3.0], [1, 1, 0], 0, 1) == 1
    assert port_oracle.choose_hap(0, [-1.0, -2.0, -0.5], [1, 1, 0], 0, 1) == 0
    assert port_oracle.choose_hap(1, [-1.0, -0.5, -0.5], [0, 0, 0]) == 1          # first strict maximum


def _hap_ref_maps(rng, batch):
    """Synthetic haplotype-to-reference alignments (hap.ml.hpos): monotone reference positions with inserted runs
    (INS), deleted reference stretches (jumps) and, now and then, LO / RO flanks."""
    out = np.empty(int(batch.hap_off[-1]), dtype=np.int32)
    for h in range(batch.n_haps):
        o, H = int(batch.hap_off[h]), int(batch.hap_off[h + 1] - batch.hap_off[h])
        m, pos, b = np.empty(H, dtype=np.int32), int(rng.integers(0, 6)), 0
        while b < H:
            u = rng.random()
            if u < 0.02:
                k = min(int(rng.integers(1, 6)), H - b)
                m[b:b + k] = INS; b += k
                continue
            if u < 0.04:
                pos += int(rng.integers(1, 8))
            m[b] = pos; pos += 1; b += 1
        if rng.random() < 0.15:
            m[:int(rng.integers(1, 20))] = LO
        if rng.random() < 0.15:
            m[H - int(rng.integers(1, 20)):] = RO
        if rng.random() < 0.02:
            m[:] = INS                       # a haplotype with no base on the reference: every read wholly clipped
        out[o:o + H] = m
    return out


@pytest.mark.gpu
@pytest.mark.parametrize("cfg,nwin,cap", [(2, 40, 16), (2, 12, 2), (3, 2, 16), (4, 2, 16)])
def test_k5_matches_oracle(engine_cli, port_oracle, cfg, nwin, cap):
    from dindel_tgi_b200 import engine, synth
    rng = np.random.default_rng(77 + cfg + cap)
    batch = synth.make_batch(cfg, nwin, seed=999 + cfg)
    res = engine_cli.compute_likelihoods(batch, want_hpos=True)
    hml = _hap_ref_maps(rng, batch)
    n_ind = rng.integers(0, 3, batch.n_haps).astype(np.int32)
    ref_start = rng.integers(1000, 5_000_000, batch.n_windows).astype(np.int32)
    nhw = np.diff(batch.win_hap)
    h1 = np.array([rng.integers(0, n) for n in nhw], dtype=np.int32)
    h2 = np.array([rng.integers(0, n) for n in nhw], dtype=np.int32)
    h1[rng.random(batch.n_windows) < 0.1] = -1           # window without a MAP pair: skipped
    shapes, n_tie = set(), 0
    for mode in (engine.CIGAR_PAIR, engine.CIGAR_BEST):
        rh, rp, rn, ops = engine_cli.cigars(mode, ref_start, hml, h1, h2, n_ind, cigar_cap=cap)
        for w in range(batch.n_windows):
            h0, nh, r0, r1 = int(batch.win_hap[w]), int(nhw[w]), int(batch.win_read[w]), int(batch.win_read[w + 1])
            p0 = int(res.win_pair[w])
            for r in range(r0, r1):
                R = int(batch.read_off[r + 1] - batch.read_off[r])
                row = slice(p0 + (r - r0) * nh, p0 + (r - r0 + 1) * nh)
                ll = res.pairs["ll"][row]
                if (mode == engine.CIGAR_PAIR and h1[w] < 0) or (mode == engine.CIGAR_BEST and not res.read_on_hap[r]):
                    assert (rh[r], rn[r]) == (-1, 0), (mode, w, r)
                    continue
                hx = port_oracle.choose_hap(mode, ll, n_ind[h0:h0 + nh], int(h1[w]), int(h2[w]))
                assert rh[r] == hx, (mode, w, r)
                n_tie += mode == engine.CIGAR_PAIR and abs(ll[h1[w]] - ll[h2[w]]) < 1e-8 and h1[w] != h2[w]
                hp = res.hpos[p0 + (r - r0) * nh + hx][:R].astype(np.int32)
                ho = int(batch.hap_off[h0 + hx])
                n, ref_pos, want = port_oracle.get_cigar(hp, hml[ho:int(batch.hap_off[h0 + hx + 1])], int(ref_start[w]), cap=cap)
                assert rn[r] == n and rp[r] == ref_pos, (mode, w, r, rn[r], n, rp[r], ref_pos)
                got = [(int(x) & 15, int(x) >> 4) for x in ops[r][:max(0, min(n, cap))]]
                assert got == want, (mode, w, r, got, want)
                if 0 < n <= cap:
                    assert sum(ln for op, ln in got if op in (M, I, S)) == R
                    shapes.add(tuple(op for op, _ in got))
    if cfg == 2 and cap == 16:
        assert {(M,), (S, M), (M, S), (S,)} <= shapes and any(I in s for s in shapes) and any(D in s for s in shapes)


@pytest.mark.gpu
def test_k5_full_size_invariants(engine_cli):
    """configs[1]-sized batch: every realigned read's M + I + S lengths add up to the read, no operation is empty,
    neighbouring operations differ, and the chosen haplotype of PAIR mode is one of the pair."""
    from dindel_tgi_b200 import engine, synth
    batch = synth.make_batch(2, 2000, seed=31)
    engine_cli.compute_likelihoods(batch, want_hpos=False, keep_hpos=True)
    hml = np.concatenate([np.arange(int(batch.hap_off[h + 1] - batch.hap_off[h]), dtype=np.int32) + 3
                          for h in range(batch.n_haps)])
    nhw = np.diff(batch.win_hap)
    h1 = np.zeros(batch.n_windows, dtype=np.int32)
    h2 = (nhw - 1).astype(np.int32)
    rh, rp, rn, ops = engine_cli.cigars(engine.CIGAR_PAIR, np.full(batch.n_windows, 100, np.int32), hml, h1, h2,
                                        np.zeros(batch.n_haps, np.int32), cigar_cap=16)
    assert engine_cli.last_kernel_ms > 0
    R = np.diff(batch.read_off)
    assert (rn != 0).all()
    ok = (rn >= 1) & (rn <= 16)              # the rest: the reference's throw cases / more than cap operations
    assert ok.mean() > 0.999
    k = (np.arange(16)[None, :] < rn[:, None]) & ok[:, None]
    op, ln = ops & 15, (ops >> 4).astype(np.int64)
    assert (ln[k] > 0).all() and (ops[ok][~k[ok]] == 0).all()
    assert ((ln * (k & np.isin(op, (M, I, S)))).sum(axis=1)[ok] == R[ok]).all()
    assert not (k[:, 1:] & (op[:, 1:] == op[:, :-1])).any()
    win_of_read = np.repeat(np.arange(batch.n_windows), np.diff(batch.win_read))
    assert ((rh == h1[win_of_read]) | (rh == h2[win_of_read])).all()
    aligned = rp >= 0
    assert aligned.mean() > 0.5 and (rp[aligned] >= 103).all()


@pytest.mark.gpu
def test_k5_needs_resident_hpos(engine_cli):
    from dindel_tgi_b200 import engine, synth
    batch = synth.make_batch(2, 2, seed=5)
    engine_cli.compute_likelihoods(batch, want_hpos=False)
    with pytest.raises(RuntimeError, match="hpos"):
        engine_cli.cigars(engine.CIGAR_BEST, np.zeros(2, np.int32), np.zeros(int(batch.hap_off[-1]), np.int32))
