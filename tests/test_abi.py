"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every symbol that
include/dindel_b200.h declares, and fails loudly (no CPU fallback) without a CUDA device."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from dindel_tgi_b200 import engine
from tests import util

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "dindel_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(dindel_b200_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = engine.load_library()
    names = _declared_symbols()
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), n
    assert sorted(engine.EXPORTS) == names
    assert lib.dindel_b200_abi_version() == 5


def test_struct_layouts_match_header():
    assert C.sizeof(engine.Params) == 56
    assert engine.PAIR_DTYPE.itemsize == 64
    assert C.sizeof(engine._BatchIn) == 16 + 14 * 8
    assert C.sizeof(engine._BatchOut) == 8 + 8 + 8 + 8 + 8 + 8 + 8
    assert C.sizeof(engine.Timing) == 5 * 4 + 2 * 4 + 4 + 16 + 4 * 4  # 4 bytes padding before the int64s
    assert C.sizeof(engine._Job) == 6 * 8 + 8 + 3 * 8 and C.sizeof(engine._EmOut) == 6 * 8


def test_struct_sizes_against_the_compiled_header(tmp_path):
    """The ctypes mirrors in engine.py against sizeof() of the header itself, compiled as plain C (the boundary is a C ABI)."""
    import subprocess
    names = ["dindel_params", "dindel_batch_in", "dindel_batch_out", "dindel_pair_out", "dindel_timing", "dindel_job",
             "dindel_em_out", "dindel_stats_in", "dindel_read_stats"]
    src = '#include <stdio.h>\n#include "dindel_b200.h"\nint main(void){' + "".join(
        'printf("%%zu\\n", sizeof(%s));' % n for n in names) + "return 0;}\n"
    c = tmp_path / "sizes.c"
    c.write_text(src)
    exe = str(tmp_path / "sizes")
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), str(c), "-o", exe])
    got = [int(x) for x in subprocess.check_output([exe]).split()]
    want = [C.sizeof(engine.Params), C.sizeof(engine._BatchIn), C.sizeof(engine._BatchOut), engine.PAIR_DTYPE.itemsize,
            C.sizeof(engine.Timing), C.sizeof(engine._Job), C.sizeof(engine._EmOut), C.sizeof(engine._StatsIn),
            engine.STATS_DTYPE.itemsize]
    assert got == want, list(zip(names, got, want))


def test_param_presets_follow_reference_defaults():
    d, c = engine.default_params(), engine.cli_params()
    assert (d.pError, d.pMut, d.maxLengthDel, d.padCover, d.bMid) == (1e-4, 1e-4, 10, 5, -1)   # ObservationModel.hpp:39-62
    assert (c.pError, c.pMut, c.maxLengthDel, c.padCover) == (5e-4, 1e-5, 5, 2)               # DInDel.cpp:4122-4157
    assert d.pFirstgLO == 0.01 and d.checkBaseQualThreshold == 0.95 and d.mapQualThreshold == 100.0


def test_count_pairs_and_cells_host_logic():
    batch, _, want, _ = util.load_golden("cfg2_cli")
    lib = engine.load_library()
    cb = batch._c()
    assert lib.dindel_b200_count_pairs(C.byref(cb)) == len(want) == batch.n_pairs()
    assert lib.dindel_b200_count_cells(C.byref(cb)) == batch.n_cells()
    bad = engine.Batch(batch.hap_start, batch.win_hap[::-1].copy(), batch.win_read, batch.hap_off, batch.hap_seq,
                       batch.read_off, batch.read_seq, batch.read_qual, batch.read_mapq, batch.read_start,
                       batch.read_flags)
    assert lib.dindel_b200_count_pairs(C.byref(bad._c())) == -1


def test_large_batches_are_validated_in_slices():
    """Above 200 k reads the CSR checks run on several host threads (validate_parallel in csrc/api.cu): a defect anywhere in
    any slice still refuses the batch.  Host only, through dindel_b200_count_pairs."""
    from dindel_tgi_b200 import synth
    batch = synth.make_batch(2, 4000, seed=12)
    assert batch.n_reads >= 200000
    lib = engine.load_library()
    assert lib.dindel_b200_count_pairs(C.byref(batch._c())) == batch.n_pairs()
    for field in ("read_off", "hap_off", "win_read", "win_hap"):
        for pos in (0.02, 0.26, 0.5, 0.74, 0.98):
            arr = getattr(batch, field).copy()
            k = int(len(arr) * pos)
            arr[k], arr[k + 1] = arr[k + 1], arr[k]
            if arr[k] == arr[k + 1]:
                continue
            kw = dict(win_hap=batch.win_hap, win_read=batch.win_read, hap_off=batch.hap_off, read_off=batch.read_off)
            kw[field] = arr
            bad = engine.Batch(batch.hap_start, kw["win_hap"], kw["win_read"], kw["hap_off"], batch.hap_seq, kw["read_off"],
                               batch.read_seq, batch.read_qual, batch.read_mapq, batch.read_start, batch.read_flags)
            assert lib.dindel_b200_count_pairs(C.byref(bad._c())) == -1, (field, pos)


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    with pytest.raises(RuntimeError, match="no usable CUDA device"):
        engine.Engine(engine.cli_params(), 0)


def test_product_package_never_touches_the_oracle():
    pkg = os.path.join(ROOT, "dindel_tgi_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h")):
                txt = open(os.path.join(dp, f)).read()
                assert "pyoracle" not in txt and "dindel_oracle" not in txt and "libdindel_ref" not in txt, f


def test_synth_is_deterministic_and_shaped():
    from dindel_tgi_b200 import synth
    a, b = synth.make_batch(2, 5), synth.make_batch(2, 5)
    for f in util.FIELDS:
        assert np.array_equal(getattr(a, f), getattr(b, f)), f
    nh = np.diff(a.win_hap)
    assert nh.min() >= 2 and nh.max() <= 8
    assert set(np.diff(a.read_off)) == {100}
    hl = np.diff(a.hap_off)
    assert hl.min() >= 100 and hl.max() <= 140
    c = synth.make_batch(4, 1)
    assert set(np.diff(c.read_off)) == {250} and np.diff(c.win_hap)[0] == 32
