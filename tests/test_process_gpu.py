"""dindel_b200_process (one pass: K1 + K4 + K2 + K3, streamed through bounded device memory in ranges of windows) and
dindel_b200_multi_process (the window scheduler over the GPUs of one box): GPU parity tests through the C ABI.

  * a batch streamed in several ranges / chunks gives byte-identical buffers to the whole batch resident at once,
    for every output (pair records, hpos rows, window status, onHap, filtered[], coverage counts, pair sums, EM)
  * the stage outputs equal the stand-alone entries (dindel_b200_coverage / pair_sums / bayes_em) and the oracle
  * the same batch on 1 vs N devices -> byte-identical buffers (SURVEY 8e Test); N contexts on ONE device exercise the
    scheduler's threads and sharding where a second GPU is absent
  * the generic kernel inside a pipelined multi-chunk submit on a fresh context (stale CSR reads)
"""
import os

import numpy as np
import pytest

from tests import util
from tests.test_cover import _random_entries


def _fresh_engine(env):
    """Engine whose context was created under the given environment (the library reads its knobs at create)."""
    from dindel_tgi_b200 import engine
    old = {k: os.environ.get(k) for k in env}
    os.environ.update({k: str(v) for k, v in env.items()})
    try:
        return engine.Engine(engine.cli_params(), 0)
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


def _job_inputs(batch, seed):
    rng = np.random.default_rng(seed)
    ents, nks = zip(*[_random_entries(rng, batch, w) for w in range(batch.n_windows)])
    strand = rng.integers(0, 2, batch.n_reads).astype(np.uint8)
    return (list(ents), list(nks), strand, 2, 1)


def _assert_same(a, b, what):
    assert a.pairs.tobytes() == b.pairs.tobytes(), what + ": pair records"
    for f in ("win_pair", "win_status", "read_on_hap", "hap_filtered", "hap_uncovered", "key_nf", "key_nr", "pair_sums",
              "win_sum_off"):
        if hasattr(a, f):
            assert np.array_equal(getattr(a, f), getattr(b, f)), what + ": " + f
    if a.hpos is not None:
        assert a.hpos.shape == b.hpos.shape


@pytest.mark.gpu
def test_process_streamed_ranges_identical_to_resident(engine_cli, port_oracle):
    from dindel_tgi_b200 import synth
    batch = synth.make_batch(2, 48, seed=321)
    cover = _job_inputs(batch, 7)
    whole = engine_cli.process(batch, want_pairs=True, want_hpos=True, cover=cover, pair_sums=True)
    assert whole.timing.n_ranges == 1
    # stand-alone entries on the same resident batch
    filt, unc, nf, nr = engine_cli.coverage(cover[0], cover[1], cover[2], max_mismatch=2, do_filter=1)
    assert np.array_equal(filt, whole.hap_filtered[:batch.n_haps]) and np.array_equal(unc, whole.hap_uncovered[:batch.n_haps])
    assert np.array_equal(np.concatenate(nf), whole.key_nf[:int(whole.win_key[-1])])
    nh = np.diff(batch.win_hap)
    for w in range(batch.n_windows):
        rl = whole.pairs["ll"][whole.win_pair[w]:whole.win_pair[w + 1]]
        npairs = nh[w] * (nh[w] + 1) // 2
        want = port_oracle.pair_sums(rl, int(nh[w]), np.zeros(npairs))
        got = whole.pair_sums[whole.win_sum_off[w]:whole.win_sum_off[w + 1]]
        assert np.allclose(got, want, rtol=1e-9, atol=0), w
    # the oracle on the same windows: K1 parity through the new path
    from oracle.pyoracle import params_from
    want, want_hpos = util.oracle_batch(port_oracle, batch, params_from(engine_cli.params))
    util.assert_records_equal(whole.pairs, want, "process")
    util.assert_hpos_equal(whole.hpos, want_hpos, batch, "process")
    # ~1/5 of the batch per range, two chunks per range
    per_win = (batch.input_bytes() + whole.pairs.nbytes + whole.hpos.nbytes) // batch.n_windows
    eng = _fresh_engine({"DINDEL_B200_MEM_BUDGET": 10 * per_win, "DINDEL_B200_CHUNKS": 2})
    try:
        part = eng.process(batch, want_pairs=True, want_hpos=True, cover=cover, pair_sums=True)
        assert part.timing.n_ranges >= 4 and part.timing.n_chunks >= 2 * part.timing.n_ranges - 1
        _assert_same(whole, part, "ranges")
        util.assert_hpos_equal(part.hpos, whole.hpos.astype(np.int32), batch, "ranges")
        # streamed -> not resident: the stand-alone stages refuse instead of reading a stale range
        with pytest.raises(RuntimeError, match="resident"):
            eng.coverage(cover[0], cover[1], cover[2])
        # aggregates only: no pair record crosses PCIe, stage outputs unchanged
        agg = eng.process(batch, want_pairs=False, want_hpos=False, cover=cover, pair_sums=True)
        assert agg.pairs is None
        for f in ("hap_filtered", "key_nf", "key_nr", "pair_sums", "win_status", "read_on_hap"):
            assert np.array_equal(getattr(agg, f), getattr(whole, f)), f
    finally:
        eng.close()


@pytest.mark.gpu
def test_process_pooled_stages(engine_cli, port_oracle):
    """per pool-member pair sums (DInDel.cpp:2683-2694) and the EM of sets known up front, inside one process call."""
    from dindel_tgi_b200 import synth
    batch = synth.make_batch(2, 10, seed=77)
    rng = np.random.default_rng(5)
    G = 3
    rg = rng.integers(0, G, batch.n_reads).astype(np.int32)
    nh = np.diff(batch.win_hap)
    sets, comps = [], []
    for w in range(batch.n_windows):
        for _ in range(3):
            c = (rng.random(nh[w]) < 0.7).astype(np.uint8)
            c[0] = 1
            sets.append(w); comps.append(c)
    res = engine_cli.process(batch, pair_sums=True, n_groups=G, read_group=rg, em=(sets, comps, 0.001, 1e-4))
    ll, it, pi, fr = engine_cli.bayes_em(sets, comps, 0.001, 1e-4)      # stand-alone on the resident batch
    assert np.array_equal(ll, res.em_loglik) and np.array_equal(it, res.em_iters)
    assert np.array_equal(np.concatenate(pi), res.em_pi) and np.array_equal(engine_cli.last_em_margin, res.em_margin)
    for w in range(batch.n_windows):
        rl = res.pairs["ll"][res.win_pair[w]:res.win_pair[w + 1]]
        npairs = int(nh[w] * (nh[w] + 1) // 2)
        for g in range(G):
            want = port_oracle.pair_sums(rl, int(nh[w]), np.zeros(npairs), rg[batch.win_read[w]:batch.win_read[w + 1]], g)
            o = int(res.win_sum_off[w]) * G + g * npairs
            assert np.allclose(res.pair_sums[o:o + npairs], want, rtol=1e-9, atol=1e-300), (w, g)
    for j, w in enumerate(sets):
        rl = res.pairs["ll"][res.win_pair[w]:res.win_pair[w + 1]]
        wll, wit, wpi, wfr = port_oracle.bayes_em(rl, int(nh[w]), comps[j], 0.001, 1e-4)
        assert it[j] == wit and abs(ll[j] - wll) <= 1e-9 * abs(wll)
    # ranges: the set list is sliced per range
    eng = _fresh_engine({"DINDEL_B200_MEM_BUDGET": 3 * (batch.input_bytes() + res.pairs.nbytes) // batch.n_windows})
    try:
        part = eng.process(batch, pair_sums=True, n_groups=G, read_group=rg, em=(sets, comps, 0.001, 1e-4))
        assert part.timing.n_ranges >= 3
        for f in ("pair_sums", "em_loglik", "em_iters", "em_pi", "em_freqs", "em_margin"):
            assert np.array_equal(getattr(part, f), getattr(res, f)), f
        assert part.pairs.tobytes() == res.pairs.tobytes()
    finally:
        eng.close()


@pytest.mark.gpu
def test_generic_kernel_multichunk_on_fresh_context(port_oracle):
    """with the generic kernel in a pipelined submit the window lookup must only read the CSR slices
    its own chunk uploaded.  Fresh context (no stale values from an earlier batch), 4 chunks, 2 lanes."""
    from dindel_tgi_b200 import synth
    from oracle.pyoracle import params_from
    batch = synth.make_batch(2, 24, seed=99)
    one = _fresh_engine({"DINDEL_B200_FORCE_GENERIC": 1, "DINDEL_B200_CHUNKS": 1})
    try:
        a = one.compute_likelihoods(batch, want_hpos=True)
    finally:
        one.close()
    four = _fresh_engine({"DINDEL_B200_FORCE_GENERIC": 1, "DINDEL_B200_CHUNKS": 4})
    try:
        b = four.compute_likelihoods(batch, want_hpos=True)
        assert b.timing.n_chunks == 4
        assert a.pairs.tobytes() == b.pairs.tobytes()
        util.assert_hpos_equal(b.hpos, a.hpos.astype(np.int32), batch, "generic x4")
        want, want_hpos = util.oracle_batch(port_oracle, batch, params_from(four.params))
        util.assert_records_equal(b.pairs, want, "generic x4")
    finally:
        four.close()


@pytest.mark.gpu
def test_empty_batch_is_ok(engine_cli):
    from dindel_tgi_b200 import engine
    empty = engine.Batch([], [0], [0], [0], [], [0], [], [], [], [], [])
    res = engine_cli.compute_likelihoods(empty, want_hpos=False)
    assert len(res.pairs) == 0 and res.win_pair[0] == 0


def _multi_vs_single(devices):
    from dindel_tgi_b200 import engine, synth
    batch = synth.make_batch(2, 64, seed=11)
    cover = _job_inputs(batch, 3)
    e = engine.Engine(engine.cli_params(), devices[0])
    try:
        single = e.process(batch, want_pairs=True, want_hpos=True, cover=cover, pair_sums=True)
    finally:
        e.close()
    m = engine.MultiEngine(engine.cli_params(), devices)
    try:
        multi = m.process(batch, want_pairs=True, want_hpos=True, cover=cover, pair_sums=True)
        _assert_same(single, multi, "multi %s" % (devices,))
        util.assert_hpos_equal(multi.hpos, single.hpos.astype(np.int32), batch, "multi")
        # shards: contiguous, cover every window, near-equal cost
        assert multi.shards[0][0] == 0 and multi.shards[-1][1] == batch.n_windows
        assert all(multi.shards[k][1] == multi.shards[k + 1][0] for k in range(len(devices) - 1))
        cells = [t.n_cells for _, _, t in multi.shards]
        assert sum(cells) == batch.n_cells() and max(cells) <= 1.5 * (sum(cells) / len(devices))
        # a second job on the same scheduler (worker threads are persistent)
        again = m.process(batch, want_pairs=True, want_hpos=False)
        assert again.pairs.tobytes() == single.pairs.tobytes()
    finally:
        m.close()


@pytest.mark.gpu
def test_multi_process_two_contexts_one_device():
    _multi_vs_single((0, 0))
    _multi_vs_single((0, 0, 0))


@pytest.mark.gpu
def test_multi_process_validates_large_batches_in_slices():
    """Above 200 k reads the multi-device call checks the CSR arrays on several host threads (validate_parallel): a valid
    batch gives the single-context records, a batch whose read_off / hap_off / win_read is not monotone somewhere in the
    middle of a slice is refused as a whole, whichever slice holds the defect."""
    from dindel_tgi_b200 import engine, synth
    batch = synth.make_batch(2, 4000, seed=12)
    assert batch.n_reads >= 200000
    e = engine.Engine(engine.cli_params(), 0)
    single = e.compute_likelihoods(batch, want_hpos=False)
    e.close()
    m = engine.MultiEngine(engine.cli_params(), (0, 0, 0))
    try:
        multi = m.process(batch, want_pairs=True, want_hpos=False)
        assert multi.pairs.tobytes() == single.pairs.tobytes()
        for field, pos in (("read_off", 0.9), ("read_off", 0.1), ("hap_off", 0.5), ("win_read", 0.7)):
            arr = getattr(batch, field).copy()
            k = int(len(arr) * pos)
            arr[k], arr[k + 1] = arr[k + 1], arr[k]
            bad = engine.Batch(batch.hap_start, batch.win_hap, batch.win_read if field != "win_read" else arr,
                               batch.hap_off if field != "hap_off" else arr, batch.hap_seq,
                               batch.read_off if field != "read_off" else arr, batch.read_seq, batch.read_qual, batch.read_mapq,
                               batch.read_start, batch.read_flags)
            with pytest.raises(RuntimeError, match="inconsistent batch|bad argument"):
                m.process(bad, want_pairs=True, want_hpos=False)
        again = m.process(batch, want_pairs=True, want_hpos=False)   # the scheduler survives the refusals
        assert again.pairs.tobytes() == single.pairs.tobytes()
    finally:
        m.close()


@pytest.mark.gpu
def test_multi_process_two_devices():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("one GPU visible")
    _multi_vs_single(tuple(range(min(torch.cuda.device_count(), 8))))
