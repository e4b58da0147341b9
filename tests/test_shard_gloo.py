"""Multi-GPU path on CPU: the window scheduler (dindel_b200_plan_shards) + the gather of the shard
results, run as a world_size-2 `gloo` job.  Windows are independent (DInDel.cpp:1310-1411), so
the N>1 path is "every rank computes its own contiguous window range, results concatenated in
window order, no data-path collective".  Here each rank's compute leg is the CPU oracle (this is a
test); on the GPU box the same slicing feeds dindel_b200_submit (tests/test_k1_gpu.py)."""
import os
import socket
import sys

import numpy as np
import pytest

from dindel_tgi_b200 import engine, synth
from tests import util

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_plan_shards_covers_and_balances():
    batch = synth.make_batch(2, 64, seed=5)
    for n in (1, 2, 3, 8, 64, 100):
        b = engine.plan_shards(batch, n)
        assert b[0] == 0 and b[-1] == batch.n_windows and np.all(np.diff(b) >= 0)
        cells = [batch.slice_windows(int(b[k]), int(b[k + 1])).n_cells() if b[k + 1] > b[k] else 0 for k in range(n)]
        assert sum(cells) == batch.n_cells()
        if n <= 8:   # near-equal cost: no shard above the mean by more than the largest window
            per_win = max(batch.slice_windows(w, w + 1).n_cells() for w in range(batch.n_windows))
            assert max(cells) <= batch.n_cells() / n + per_win
    assert list(engine.plan_shards(batch.slice_windows(0, 0), 4)) == [0, 0, 0, 0, 0]


def _rank_main(rank, world, port, q):
    import torch.distributed as dist
    from oracle.pyoracle import Oracle, params_from
    sys.path.insert(0, ROOT)
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    batch = synth.make_batch(2, 6, seed=77)           # every rank builds the same job description
    bounds = engine.plan_shards(batch, world)
    mine = batch.slice_windows(int(bounds[rank]), int(bounds[rank + 1]))
    p = params_from(engine.cli_params())
    recs, _ = util.oracle_batch(Oracle("port"), mine, p)   # stand-in for dindel_b200_submit on this rank's GPU
    gathered = [None] * world
    dist.all_gather_object(gathered, (int(bounds[rank]), recs.tobytes(), mine.n_pairs()))
    dist.barrier()
    if rank == 0:
        gathered.sort(key=lambda t: t[0])
        q.put((b"".join(g[1] for g in gathered), sum(g[2] for g in gathered)))
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_rank_shards_equal_single_process():
    import torch.multiprocessing as mp
    from oracle.pyoracle import Oracle, params_from
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_rank_main, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    blob, n_pairs = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    batch = synth.make_batch(2, 6, seed=77)
    want, _ = util.oracle_batch(Oracle("port"), batch, params_from(engine.cli_params()))
    assert n_pairs == batch.n_pairs() == len(want)
    assert blob == want.tobytes()      # byte-identical to the single-process result, in window order
