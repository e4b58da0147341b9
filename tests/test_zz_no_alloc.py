"""dindel_b200_run_static (include/dindel_b200.h): K1 over the resident batch with the kernel's scratch in the library's
file-scope device array, no allocation inside the call.  Through the C ABI, against the existing entry point
(dindel_b200_run + dindel_b200_download), byte for byte (K1 is FP64 add / compare only: the tolerance of the K1 tests is
zero), on two calls in a row over DIFFERENT batches, so that the second call runs on the scratch the first one left
behind.  Collected last on purpose (the file name).
"""
import ctypes as C

import numpy as np
import pytest

from dindel_tgi_b200 import engine, synth
from tests import util

pytestmark = pytest.mark.gpu


def _reference_results(e, batch):
    e.upload(batch, want_hpos=True)
    e.run()
    return e.download()


def _same(a, b, what):
    util.assert_records_equal(a.pairs, b.pairs, what)
    assert a.pairs.tobytes() == b.pairs.tobytes(), what
    assert np.array_equal(a.hpos, b.hpos), what
    assert np.array_equal(a.win_status, b.win_status), what
    assert np.array_equal(a.read_on_hap, b.read_on_hap), what
    assert np.array_equal(a.win_pair, b.win_pair), what


def test_run_static_matches_run_twice_in_a_row():
    import torch
    e = engine.Engine(util.make_params(engine.Params, "cli"), 0)
    assert e.lib.dindel_b200_static_scratch_bytes() == 16 << 20
    batches = [synth.make_batch(2, 40, seed=501), synth.make_batch(2, 23, seed=502), synth.make_batch(2, 40, seed=501)]
    want = [_reference_results(e, b) for b in batches]
    assert all(len(w.pairs) > 1000 and (w.win_status == 0).all() for w in want)
    free = []
    for k, (b, w) in enumerate(zip(batches, want)):
        e.upload(b, want_hpos=True)           # set-up: may allocate
        torch.cuda.synchronize()
        f0 = torch.cuda.mem_get_info(0)[0]
        got = e.run_static()                  # the call under test: allocates nothing
        free.append((f0, torch.cuda.mem_get_info(0)[0]))
        _same(got, w, "call %d" % k)
        assert got.timing.k1_launches == 2    # the generic kernel + the window summary
    # device memory is untouched by the calls (the first one may load the kernel's module: not counted)
    assert free[1][0] == free[1][1] and free[2][0] == free[2][1], free
    # results left resident (out = NULL), fetched by the existing download
    e.upload(batches[1], want_hpos=True)
    assert e.lib.dindel_b200_run_static(e.h, None) == 0
    _same(e.download(), want[1], "resident")
    e.close()


def test_run_static_refuses_what_does_not_fit():
    """configs[3] shapes (250-base reads x ~320 columns: 81 KB of back-pointers per warp) exceed the 16 MiB array."""
    e = engine.Engine(util.make_params(engine.Params, "cli"), 0)
    b = synth.make_batch(4, 2, seed=7)
    e.upload(b, want_hpos=False)
    o, *_ = e._alloc_out(b, False)
    rc = e.lib.dindel_b200_run_static(e.h, C.byref(o))
    assert rc == -5  # DINDEL_E_UNSUPPORTED
    msg = e.lib.dindel_b200_last_error(e.h).decode()
    assert "too large for the static scratch" in msg and "dindel_b200_run" in msg, msg
    # the context is still usable, through both entry points
    e.run()
    res = e.download()
    small = synth.make_batch(2, 5, seed=9)
    want = _reference_results(e, small)
    e.upload(small, want_hpos=True)
    _same(e.run_static(), want, "after a refusal")
    assert len(res.pairs) == b.n_pairs()
    # no resident batch
    e2 = engine.Engine(util.make_params(engine.Params, "cli"), 0)
    assert e2.lib.dindel_b200_run_static(e2.h, None) == -1
    e2.close()
    e.close()
