mkdir -p gpurun_out
cat > /tmp/e2e.py <<'PY'
import sys, os, time, ctypes as C, numpy as np
sys.path.insert(0, os.getcwd())
from dindel_tgi_b200 import engine, synth
import bench
batch = synth.make_batch(2, 20000)
eng = engine.Engine(engine.cli_params(), 0); lib = eng.lib
names = ("hap_start","win_hap","win_read","hap_off","hap_seq","read_off","read_seq","read_qual","read_mapq","read_start","read_flags")
pb = engine.Batch(*[bench.pinned_copy(lib, getattr(batch, n)) for n in names])
n_pairs = batch.n_pairs()
pairs = bench.pinned_copy(lib, np.zeros(n_pairs, dtype=engine.PAIR_DTYPE))
stride = (int(np.max(np.diff(batch.read_off))) + 7) & ~7
hpos = bench.pinned_copy(lib, np.zeros((n_pairs, stride), dtype=np.int16))
ws = bench.pinned_copy(lib, np.zeros(batch.n_windows, dtype=np.int32)); oh = bench.pinned_copy(lib, np.zeros(batch.n_reads, dtype=np.uint8))
wp = np.zeros(batch.n_windows + 1, dtype=np.int64)
o = engine._BatchOut(); o.n_pairs, o.pairs, o.win_pair = n_pairs, pairs.ctypes.data, wp.ctypes.data
o.win_status, o.read_on_hap, o.hpos, o.hpos_stride = ws.ctypes.data, oh.ctypes.data, hpos.ctypes.data, stride
cb = pb._c()
for i in range(5):
    t0 = time.perf_counter(); eng._check(lib.dindel_b200_submit(eng.h, C.byref(cb), C.byref(o)), "submit"); wall = 1e3*(time.perf_counter()-t0)
    t = eng.timing()
print("chunks=%s lanes=%s wall %.1f ms device %.1f ms sum_k1 %.1f ms launches %d" % (os.environ.get("DINDEL_B200_CHUNKS"), os.environ.get("DINDEL_B200_LANES"), wall, t.total_ms, t.k1_ms, t.k1_launches))
PY
python -m pytest tests -m gpu -x -q 2>&1 | tail -n 2
for ch in 4 8 16 0; do for ln in 1 2; do DINDEL_B200_CHUNKS=$ch DINDEL_B200_LANES=$ln python /tmp/e2e.py 2>&1 | tail -n 1; done; done
