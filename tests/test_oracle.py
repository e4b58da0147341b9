"""CPU tests: the plain-C restatement (oracle/dindel_oracle.c) is pinned against
 (1) the reference's smain() known-answer fixtures (DInDel.cpp:3993-4066),
 (2) golden vectors produced by the reference's own compiled code (tests/golden/k1_golden.npz),
 (3) the reference's own code itself (oracle/_ref) on fresh random pairs, when it is built."""
import numpy as np
import pytest

from oracle.pyoracle import OraParams
from tests import util

DEFAULT = OraParams(*util.PSETS["default"])
CLI = OraParams(*util.PSETS["cli"])

HAP1 = b"ATCGATTCGTGATATATATATTCAATGTAGTCGCTAG"
HAP3 = b"ATCGATTCGTGTTTTTTCAATGTAGTCGCTAG"
HAP4 = (b"AAAATCACCAACACTTCATAATCTATTTTTTCCCCTGAGGAACTTCCTAAAATGAATAAAAAAAAACCCCAGCCACATCTGCATTTGCAAACAGGAAACTC"
        b"TGCAAGCCATACTAAGACCAAAGCTTAGTT")
KATS = [
    # hap, read, posStat.first, ll, llOff, firstBase, lastBase, numIndels, hpos
    (HAP1, b"ATCGATTCGTGATAATATTCAATGTAGTCGCTAG", 0.0, -12.208726495298166, -23.458611539743039, 0, 36, 1,
     list(range(0, 14)) + list(range(17, 37))),
    (HAP1, b"ATCGATTCGTGATATATATATAATTCAATGTAGTCGCTAG", 0.0, -11.994036925699257, -23.545031394973467, 0, 36, 1,
     list(range(0, 21)) + [-1, -1, -1] + list(range(21, 37))),
    (HAP3, b"ATCGATTCGTGTTTTTCAATGTAGTCGCTAG", 0.0, -8.0612584595631169, -23.415401612127827, 0, 31, 1,
     list(range(0, 15)) + list(range(16, 32))),
    (HAP4, b"CAAACAGGAAACTCTGCAAGCCATACTAAGACCAAAGCTTAGTTA", 87.0, -0.34501268318605621, -23.593597157340653,
     87, 130, 0, list(range(87, 131)) + [-4]),
]


@pytest.mark.parametrize("k", range(len(KATS)))
def test_kat_smain(port_oracle, k):
    hap, read, ps, ll, lloff, fb, lb, nind, hpos = KATS[k]
    rec, hp, ms = port_oracle.pair(hap, read, np.full(len(read), 0.99), 1 - 1e-16, ps, 0, 0, DEFAULT)
    assert rec["ll"] == ll and rec["llOn"] == ll and rec["llOff"] == lloff
    assert (rec["firstBase"], rec["lastBase"], rec["numIndels"], rec["offHap"]) == (fb, lb, nind, 0)
    assert list(hp) == hpos


@pytest.mark.parametrize("case", util.golden_cases())
def test_port_matches_reference_golden(port_oracle, case):
    batch, pset, want, want_hpos = util.load_golden(case)
    got, got_hpos = util.oracle_batch(port_oracle, batch, OraParams(*util.PSETS[pset]))
    util.assert_records_equal(got, want, case)
    util.assert_hpos_equal(got_hpos, want_hpos, batch, case)


def test_port_matches_compiled_reference_random(port_oracle, ref_oracle):
    from dindel_tgi_b200 import synth
    batch = synth.make_batch(2, 6, seed=4242)
    for params in (CLI, DEFAULT):
        a, ah = util.oracle_batch(port_oracle, batch, params)
        b, bh = util.oracle_batch(ref_oracle, batch, params)
        util.assert_records_equal(a, b, "random cfg2")
        util.assert_hpos_equal(ah, bh, batch, "random cfg2")


def test_port_edge_cases_match_reference(port_oracle, ref_oracle):
    """unmapped read, read left / right of the window, forced bMid, forceReadOnHaplotype,
    N in hap and read, hapSize error."""
    rng = np.random.default_rng(5)
    hap = bytes(rng.choice(np.frombuffer(b"ACGT", np.uint8), 60))
    hapn = hap[:20] + b"N" + hap[21:]
    read = hap[10:45]
    readn = read[:5] + b"N" + read[6:]
    q = np.array([port_oracle.phred_to_prob(p) for p in rng.integers(2, 42, len(read))])
    mq = port_oracle.phred_to_prob(37)
    cases = [
        (hap, read, 1010.0, 0, 1000, DEFAULT), (hap, read, 1010.0, 1, 1000, DEFAULT),
        (hap, read, 5000.0, 0, 1000, DEFAULT), (hap, read, 10.0, 0, 1000, DEFAULT),
        (hapn, readn, 1010.0, 0, 1000, CLI), (hap, read, 990.0, 0, 1000, CLI),
        (hap, read, 1010.0, 0, 1000, OraParams(1e-4, 1e-4, 100.0, 0.01, 0.95, 10, 5, 3, 0)),
        (hap, read, 1010.0, 0, 1000, OraParams(1e-4, 1e-4, 100.0, 0.01, 0.95, 10, 5, -1, 1)),
        (hap, read, 1010.0, 0, 1000, OraParams(1e-4, 1e-4, 30.0, 0.01, 0.95, 10, 5, -1, 0)),
        (hap[:8], read, 1010.0, 0, 1000, DEFAULT),  # maxLengthDel 10 > H 8 -> "hapSize error."
        (hap, read[:1], 1010.0, 0, 1000, CLI),
    ]
    for hp_, rd, ps, um, hs, par in cases:
        a = port_oracle.pair(hp_, rd, q[:len(rd)], mq, ps, um, hs, par)
        b = ref_oracle.pair(hp_, rd, q[:len(rd)], mq, ps, um, hs, par)
        assert a[0].tobytes() == b[0].tobytes(), (a[0], b[0])
        if a[0]["status"] == 0:
            assert (a[1] == b[1]).all() and (a[2] == b[2]).all()
    assert port_oracle.pair(hap[:8], read, q, mq, 1010.0, 0, 1000, DEFAULT)[0]["status"] == 1
