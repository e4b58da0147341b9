"""K2 / K3 pinned on the reference's OWN arithmetic: tests/golden/infer_golden.json.gz holds what the stock CPU driver's
DetInDel::diploidGLF (DInDel.cpp:3085-3120, :3240-3266, :3349-3390) and estimateHaplotypeFrequenciesBayesEM
(:2431-2545, :2585-2602, :2683-2694) computed on 14 windows of a synthetic chr20 (dumped by oracle/_ref/dindel_ref_dump,
generator tests/golden/make_infer_golden.py).

  CPU   the oracle restatements (ora_pair_sums, ora_bayes_em; same glibc, same summation order) reproduce every dumped
        double BIT FOR BIT, iteration counts included
  GPU   the CUDA kernels agree to 1e-9 relative (CUDA libm vs glibc at the ulp level, reduction order over reads) with
        identical EM pass counts and identical arg-max pairs; the library's own guard band (dindel_b200_pair_sums /
        dindel_b200_bayes_em report an error bound resp. a stopping-rule margin) covers the observed differences, which
        is what lets the host adapter keep glf / VCF text byte-identical (tests/test_e2e.py)
"""
import gzip
import json
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
RTOL = 1e-9


@pytest.fixture(scope="module")
def golden():
    with gzip.open(os.path.join(HERE, "golden", "infer_golden.json.gz"), "rt") as f:
        return json.load(f)


def _tri(nh):
    return [(h1, h2) for h1 in range(nh) for h2 in range(h1, nh)]


def test_golden_has_both_modes(golden):
    dip = [w for w in golden if w["mode"] == "diploid"]
    poo = [w for w in golden if w["mode"] == "pooled"]
    assert len(dip) >= 6 and len(poo) >= 5
    assert any(sum(w["filtered"]) for w in dip)                       # --filterHaplotypes windows
    its = [s["iter"] for w in poo for s in w["sets"]]
    assert min(its) <= 3 and max(its) == 27 and len(set(its)) >= 10   # early stop, the 27-pass cap and many in between


def test_oracle_pair_sums_bit_exact_vs_reference(golden, port_oracle):
    n = 0
    for w in golden:
        if w["mode"] != "diploid":
            continue
        nh, rl = w["nh"], np.array(w["rl"])
        idx = {p: q for q, p in enumerate(_tri(nh))}
        zero = np.zeros(nh * (nh + 1) // 2)
        s0 = port_oracle.pair_sums(rl, nh, zero)                       # DInDel.cpp:3086-3089
        for h1, h2, ll, prior, c1, c2 in w["pairs"]:
            assert s0[idx[(h1, h2)]] == ll, (h1, h2)
            n += 1
        # per variant position: start = prior - logPriorPos (DInDel.cpp:3366-3372)
        by_pos = {}
        for pos, h1, h2, v1, v2, pr, ll in w["site_pairs"]:
            by_pos.setdefault(pos, []).append((h1, h2, pr, ll))
        for pos, rows in by_pos.items():
            start = zero.copy()
            for h1, h2, pr, ll in rows:
                start[idx[(h1, h2)]] = pr
            s = port_oracle.pair_sums(rl, nh, start)
            for h1, h2, pr, ll in rows:
                assert s[idx[(h1, h2)]] == ll, (pos, h1, h2)
                n += 1
    assert n > 1500


def test_oracle_diploid_decisions_vs_reference(golden, port_oracle):
    """arg-max pairs (strict >, first wins), qual and genotype grouping recomputed from the dumped sums."""
    for w in golden:
        if w["mode"] != "diploid":
            continue
        best_i = best_n = -np.inf
        pi_ = pn_ = (-1, -1)
        for h1, h2, ll, prior, c1, c2 in w["pairs"]:
            post = ll + prior
            if post > best_i and (c1 > 0 or c2 > 0):
                best_i, pi_ = post, (h1, h2)
            if post > best_n and c1 == 0 and c2 == 0:
                best_n, pn_ = post, (h1, h2)
        m = w["map"]
        assert (pi_[0], pi_[1], pn_[0], pn_[1]) == tuple(m[:4])
        assert best_i == m[4] and best_n == m[5]
        qual = -10.0 * (best_n - port_oracle.addlogs(best_i, best_n)) / np.log(10.0)
        assert qual == m[6] or (np.isnan(qual) and np.isnan(m[6]))
        by_pos = {}
        for pos, h1, h2, v1, v2, pr, ll in w["site_pairs"]:
            by_pos.setdefault(pos, []).append((h1, h2, v1, v2, ll))
        for pos, hx1, hx2, maxll, gl in w["sites"]:
            gen, best, arg = {}, -np.inf, None
            for h1, h2, v1, v2, ll in by_pos[pos]:
                g = "%d/%d" % (min(v1, v2), max(v1, v2))
                gen[g] = ll if g not in gen else port_oracle.addlogs(gen[g], ll)
                if ll > best:
                    best, arg = ll, (h1, h2)
            assert arg == (hx1, hx2) and best == maxll
            assert {g: v for g, v in gl} == gen


def test_oracle_bayes_em_bit_exact_vs_reference(golden, port_oracle):
    n = 0
    for w in golden:
        if w["mode"] != "pooled":
            continue
        nh, rl = w["nh"], np.array(w["rl"])
        logz = -np.inf
        for th, s in enumerate(w["sets"]):
            ll, it, pi, fr = port_oracle.bayes_em(rl, nh, np.array(s["compatible"], dtype=np.uint8), w["a0"], w["tol"])
            assert it == s["iter"], (th, it, s["iter"])
            assert ll == s["loglik"]
            assert list(pi) == s["pi"]
            assert list(fr) == w["freqs"][th * nh:(th + 1) * nh]
            logz = port_oracle.addlogs(logz, ll + s["logprior"])
            n += 1
        assert logz == w["logz"]
        # pooled genotype likelihoods (DInDel.cpp:2683-2692): start = prior, reads of pool member b
        idx = {p: q for q, p in enumerate(_tri(nh))}
        by_vb = {}
        for vi, b, h1, h2, pr, ll in w["pool_pairs"]:
            by_vb.setdefault((vi, b), []).append((h1, h2, pr, ll))
        for (vi, b), rows in by_vb.items():
            start = np.zeros(nh * (nh + 1) // 2)
            for h1, h2, pr, ll in rows:
                start[idx[(h1, h2)]] = pr
            got = port_oracle.pair_sums(rl, nh, start, np.array(w["pool"], dtype=np.int32), b)
            for h1, h2, pr, ll in rows:
                assert got[idx[(h1, h2)]] == ll
    assert n >= 60


# ---------------------------------------------------------------------------------------------------------------
# GPU: K2 / K3 on the reference's own rl matrices (dindel_b200_load_likelihoods makes them the resident batch)
# ---------------------------------------------------------------------------------------------------------------
@pytest.mark.gpu
def test_gpu_k2_vs_reference_golden(golden, engine_cli):
    from dindel_tgi_b200 import engine
    dip = [w for w in golden if w["mode"] == "diploid"]
    engine_cli.load_likelihoods([np.array(w["rl"]) for w in dip], [w["nh"] for w in dip])
    jobs, starts, want = [], [], []
    for k, w in enumerate(dip):
        nh = w["nh"]
        idx = {p: q for q, p in enumerate(_tri(nh))}
        npairs = nh * (nh + 1) // 2
        ref = np.full(npairs, np.nan)
        for h1, h2, ll, prior, c1, c2 in w["pairs"]:
            ref[idx[(h1, h2)]] = ll
        jobs.append(k); starts.append(np.zeros(npairs)); want.append(ref)
        by_pos = {}
        for pos, h1, h2, v1, v2, pr, ll in w["site_pairs"]:
            by_pos.setdefault(pos, []).append((h1, h2, pr, ll))
        for pos, rows in by_pos.items():
            st, ref = np.zeros(npairs), np.full(npairs, np.nan)
            for h1, h2, pr, ll in rows:
                st[idx[(h1, h2)]], ref[idx[(h1, h2)]] = pr, ll
            jobs.append(k); starts.append(st); want.append(ref)
    got = engine_cli.pair_sums(jobs, starts)
    n = 0
    worst = 0.0
    for j, k in enumerate(jobs):
        m = ~np.isnan(want[j])       # filtered pairs are not evaluated by the reference
        g, r = got[j][m], want[j][m]
        assert np.allclose(g, r, rtol=RTOL, atol=0), (j, np.abs(g - r).max())
        # the library's own guard band covers the observed difference with room to spare
        bound = np.array([engine.sum_error_bound(dip[k]["nr"], abs(x) + abs(s) + 1) for x, s in zip(r, starts[j][m])])
        assert (np.abs(g - r) <= bound).all()
        worst = max(worst, float((np.abs(g - r) / bound).max()))
        assert int(np.argmax(g)) == int(np.argmax(r))          # strict-> arg-max pair (DInDel.cpp:3102-3110, :3384)
        n += int(m.sum())
    assert n > 1500
    assert worst < 0.5, worst     # observed difference stays below half the bound


@pytest.mark.gpu
def test_gpu_k3_vs_reference_golden(golden, engine_cli):
    poo = [w for w in golden if w["mode"] == "pooled"]
    engine_cli.load_likelihoods([np.array(w["rl"]) for w in poo], [w["nh"] for w in poo])
    sets, comps, ref = [], [], []
    for k, w in enumerate(poo):
        for th, s in enumerate(w["sets"]):
            sets.append(k); comps.append(np.array(s["compatible"], dtype=np.uint8)); ref.append((w, th, s))
    ll, it, pi, fr = engine_cli.bayes_em(sets, comps, poo[0]["a0"], poo[0]["tol"])
    margin, energy = engine_cli.last_em_margin, engine_cli.last_em_energy
    worst_e = 0.0
    for j, (w, th, s) in enumerate(ref):
        nh = w["nh"]
        assert it[j] == s["iter"], (j, it[j], s["iter"], margin[j])
        assert abs(ll[j] - s["loglik"]) <= RTOL * abs(s["loglik"])
        assert np.allclose(pi[j], s["pi"], rtol=RTOL, atol=0)
        assert np.allclose(fr[j], w["freqs"][th * nh:(th + 1) * nh], rtol=1e-8, atol=1e-300)
        # energy of the last pass: the quantity the stopping rule compares (DInDel.cpp:2515); the adapter re-runs a set
        # on the host when its margin is below 1e-7 * max(1, |E|): the observed GPU-vs-reference difference must stay
        # far inside that guard
        de = abs(energy[j] - s["energy"])
        worst_e = max(worst_e, de / max(1.0, abs(s["energy"])))
        assert margin[j] >= 0.0
    assert len(ref) >= 60
    assert worst_e < 1e-10, worst_e
