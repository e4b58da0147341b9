"""K2 / K3: CPU tests of the restatements (against independent numpy/scipy formulations) and GPU
parity tests through the C ABI.  Tolerance: 1e-9 relative (CUDA libm vs glibc at the ulp level,
and K3's reduction tree over reads); arg-max pairs identical."""
import numpy as np
import pytest

from tests import util

RTOL = 1e-9


def _rl(rng, nr, nh):
    """log-likelihood matrix shaped like K1 output: most reads fit 1-2 haplotypes."""
    base = -rng.random((nr, 1)) * 5 - 0.2
    pen = rng.choice([0.0, 3.4, 12.0, 25.0], size=(nr, nh), p=[0.45, 0.25, 0.2, 0.1])
    return (base - pen - rng.random((nr, nh)) * 1e-3).reshape(-1)


def test_oracle_pair_sums_vs_numpy(port_oracle):
    rng = np.random.default_rng(1)
    nr, nh = 57, 6
    rl = _rl(rng, nr, nh)
    start = rng.random(nh * (nh + 1) // 2) * -3
    got = port_oracle.pair_sums(rl, nh, start)
    m = rl.reshape(nr, nh)
    q = 0
    for h1 in range(nh):
        for h2 in range(h1, nh):
            want = start[q] + np.sum(np.log(0.5) + np.logaddexp(m[:, h1], m[:, h2]))
            assert abs(got[q] - want) <= 1e-12 * abs(want)
            q += 1
    grp = rng.integers(0, 3, nr)
    got1 = port_oracle.pair_sums(rl, nh, start, grp, 1)
    want1 = start[0] + np.sum((np.log(0.5) + np.logaddexp(m[:, 0], m[:, 0]))[grp == 1])
    assert abs(got1[0] - want1) <= 1e-12 * abs(want1)


def test_oracle_bayes_em_vs_python_restatement(port_oracle):
    from scipy.special import digamma
    rng = np.random.default_rng(2)
    nr, nh, a0, tol = 200, 5, 0.001, 1e-4
    rl = _rl(rng, nr, nh)
    comp = np.array([1, 1, 0, 1, 1], dtype=np.uint8)
    ll, it, pi, fr = port_oracle.bayes_em(rl, nh, comp, a0, tol)
    m = rl.reshape(nr, nh)
    numah = int(comp.sum())
    lpi = np.where(comp == 1, np.log(1.0 / numah), -100.0)
    e_old, k = -np.inf, 0
    while True:   # DInDel.cpp:2431-2520 in vector form
        zl = lpi[None, :] + m
        ln = np.logaddexp.reduce(zl, axis=1)
        z = np.exp(zl - ln[:, None])
        nk = z.sum(axis=0)
        ahat = np.sum((nk + a0)[comp == 1])
        lpi = np.where(comp == 1, digamma(nk + a0) - digamma(ahat), -100.0)
        p = np.where(comp == 1, np.log((a0 + nk) / (numah * a0 + nr)), -100.0)
        e_new = np.sum(z * (p[None, :] + m))
        done = abs(e_old - e_new) < tol or k > 25
        e_old = e_new
        k += 1
        if done:
            break
    assert it == k
    assert abs(ll - ln.sum()) <= 1e-10 * abs(ll)
    assert np.allclose(pi, p, rtol=1e-10, atol=0)
    assert np.allclose(fr, np.exp(p) / np.exp(p).sum(), rtol=1e-10, atol=0)
    assert 1 <= it <= 27


@pytest.mark.gpu
def test_gpu_pair_sums_and_argmax(port_oracle):
    from dindel_tgi_b200 import engine, synth
    batch = synth.make_batch(2, 12, seed=55)
    e = engine.Engine(engine.cli_params(), 0)
    res = e.compute_likelihoods(batch, want_hpos=False)
    rng = np.random.default_rng(3)
    nh = np.diff(batch.win_hap)
    jobs, starts, groups = [], [], []
    rg = rng.integers(0, 2, batch.n_reads).astype(np.int32)
    for w in range(batch.n_windows):
        npairs = nh[w] * (nh[w] + 1) // 2
        jobs += [w, w, w]
        starts += [np.zeros(npairs), -rng.random(npairs) * 20, np.zeros(npairs)]   # :3087, :3375, :2680
        groups += [-1, -1, 1]
    got = e.pair_sums(jobs, starts, groups, rg)
    for j, w in enumerate(jobs):
        rl = res.pairs["ll"][res.win_pair[w]:res.win_pair[w + 1]]
        want = port_oracle.pair_sums(rl, int(nh[w]), starts[j], rg[batch.win_read[w]:batch.win_read[w + 1]], groups[j])
        assert np.allclose(got[j], want, rtol=RTOL, atol=0), (j, got[j], want)
        assert int(np.argmax(got[j])) == int(np.argmax(want))      # MAP pair (DInDel.cpp:3102-3110)
    e.close()


@pytest.mark.gpu
def test_gpu_bayes_em(port_oracle):
    from dindel_tgi_b200 import engine, synth
    b3 = synth.make_batch(3, 1, seed=66)
    hs, haps, reads = b3.window(0)[2], b3.window(0)[0], b3.window(0)[1]
    batch = engine.Batch.from_windows([(hs, haps, reads[:400]), (hs + 1000, haps[:5], reads[400:460])])
    e = engine.Engine(engine.cli_params(), 0)
    res = e.compute_likelihoods(batch, want_hpos=False)
    rng = np.random.default_rng(4)
    nh = np.diff(batch.win_hap)
    sets, comps = [], []
    for w in range(batch.n_windows):
        for _ in range(4):
            c = (rng.random(nh[w]) < 0.7).astype(np.uint8)
            c[0] = 1
            sets.append(w)
            comps.append(c)
        sets.append(w)
        comps.append(np.ones(nh[w], dtype=np.uint8))
    ll, it, pi, fr = e.bayes_em(sets, comps, 0.001, 1e-4)
    for j, w in enumerate(sets):
        rl = res.pairs["ll"][res.win_pair[w]:res.win_pair[w + 1]]
        wll, wit, wpi, wfr = port_oracle.bayes_em(rl, int(nh[w]), comps[j], 0.001, 1e-4)
        assert it[j] == wit
        assert abs(ll[j] - wll) <= RTOL * abs(wll)
        assert np.allclose(pi[j], wpi, rtol=RTOL, atol=0)
        assert np.allclose(fr[j], wfr, rtol=1e-8, atol=1e-300)
    e.close()


@pytest.mark.gpu
@pytest.mark.parametrize("nh,nr", [(80, 700), (65, 40), (200, 1300)])
def test_gpu_bayes_em_wide_windows(port_oracle, nh, nr):
    """More than 64 haplotypes per window (1000x pooled data has windows with 72-92): the warp-per-read kernel
    (k3_bayes_em_wide) against the oracle on a loaded likelihood matrix: identical pass counts, 1e-9 relative."""
    from dindel_tgi_b200 import engine
    rng = np.random.default_rng(nh * 1000 + nr)
    # reads drawn from a few true haplotypes: ll close to 0 on them, penalties elsewhere (the shape of a real matrix)
    truth = rng.integers(0, nh, nr)
    rl = -rng.gamma(2.0, 6.0, (nr, nh)) - 3.0
    rl[np.arange(nr), truth] = -rng.gamma(1.0, 0.3, nr)
    rl[:, : nh // 3] += 2.5                                   # a block of near-equivalent haplotypes: slow EM convergence
    e = engine.Engine(engine.cli_params(), 0)
    e.load_likelihoods([rl.ravel(), rl[:50, :8].ravel().copy()], [nh, 8])
    sets, comps = [], []
    for k in range(6):
        c = (rng.random(nh) < (0.3 + 0.1 * k)).astype(np.uint8)
        c[int(truth[0])] = 1
        sets.append(0); comps.append(c)
    sets.append(0); comps.append(np.ones(nh, dtype=np.uint8))
    sets.append(1); comps.append(np.ones(8, dtype=np.uint8))    # a narrow window in the same call
    ll, it, pi, fr = e.bayes_em(sets, comps, 0.001, 1e-4)
    mats = {0: (rl.ravel(), nh), 1: (rl[:50, :8].ravel(), 8)}
    its = []
    for j, w in enumerate(sets):
        m, n = mats[w]
        wll, wit, wpi, wfr = port_oracle.bayes_em(m, n, comps[j], 0.001, 1e-4)
        assert it[j] == wit, (j, it[j], wit)
        assert abs(ll[j] - wll) <= RTOL * abs(wll)
        assert np.allclose(pi[j], wpi, rtol=RTOL, atol=0)
        assert np.allclose(fr[j], wfr, rtol=1e-8, atol=1e-300)
        its.append(wit)
    assert len(set(its)) >= 2
    e.close()


def test_oracle_digamma_against_cephes(port_oracle):
    """boost::math::digamma (DInDel.cpp:2466, 2472) is third party and absent here, so `ora_digamma` cannot be pinned
    against it; it is cross-checked against an independent published implementation instead (Cephes psi through
    scipy.special.digamma) over the arguments the EM can produce (a0 = 0.001 up to nk + a0 for 10^5 reads): absolute
    error below 1e-14 * max(1, |psi|), i.e. the two agree to the last digits everywhere, the root at 1.4616 included."""
    from scipy.special import digamma
    xs = np.concatenate([np.logspace(-3, 5, 4000), np.arange(1, 2000) + 0.001, [1.4616321449683623, 0.001, 1.0, 2.0]])
    got = np.array([port_oracle.digamma(float(x)) for x in xs])
    want = digamma(xs)
    assert (np.abs(got - want) <= 1e-14 * np.maximum(1.0, np.abs(want))).all()
    assert abs(port_oracle.digamma(1.0) + 0.5772156649015329) < 1e-15      # -gamma
