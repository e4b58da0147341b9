"""ctypes binding of libdindel_b200.so (include/dindel_b200.h) + numpy batch containers.

This is a test / bench driver over the C ABI, not a compute path: every number comes from the
CUDA library.  There is no CPU fallback; loading fails loudly when the library is missing and
`Engine()` raises when no CUDA device is usable.

Host-side mirror of the reference interface (same names and argument meaning):
  Engine.compute_likelihoods(batch)  ~  DetInDel::computeLikelihoods  (DInDel.cpp:1707-1739)
  PairResults fields                 ~  MLAlignment                   (MLAlignment.hpp:50-70)
  Params                             ~  ObservationModelParameters    (ObservationModel.hpp:28-99)
"""
import ctypes as C
import time
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# DINDEL_B200_LIB selects another build of the same library (kernel-variant experiments)
LIB_PATH = os.environ.get("DINDEL_B200_LIB") or os.path.join(_HERE, "libdindel_b200.so")

WIN_OK, WIN_HAPSIZE, WIN_NAN, WIN_LL_POSITIVE = 0, 1, 2, 3
HPOS_INS, HPOS_LO, HPOS_RO = -1, -3, -4
READ_UNMAPPED, READ_REVERSE = 1, 2


class Params(C.Structure):
    _fields_ = [(n, C.c_double) for n in
                ("pError", "pMut", "mapQualThreshold", "pFirstgLO", "checkBaseQualThreshold")] + \
               [(n, C.c_int32) for n in ("maxLengthDel", "padCover", "bMid", "forceReadOnHaplotype")]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class _BatchIn(C.Structure):
    _fields_ = [("n_windows", C.c_int32), ("n_haps", C.c_int32), ("n_reads", C.c_int32),
                ("reserved", C.c_int32),
                ("hap_start", C.c_void_p), ("win_hap", C.c_void_p), ("win_read", C.c_void_p),
                ("hap_off", C.c_void_p), ("hap_seq", C.c_void_p), ("read_off", C.c_void_p),
                ("read_seq", C.c_void_p), ("read_qual", C.c_void_p), ("read_mapq", C.c_void_p),
                ("read_start", C.c_void_p), ("read_flags", C.c_void_p),
                ("read_ins_mode", C.c_void_p), ("read_ins_anchor", C.c_void_p), ("read_lib", C.c_void_p)]


class _BatchOut(C.Structure):
    _fields_ = [("n_pairs", C.c_int64), ("pairs", C.c_void_p), ("hpos", C.c_void_p),
                ("hpos_stride", C.c_int32), ("flags", C.c_int32), ("win_pair", C.c_void_p),
                ("win_status", C.c_void_p), ("read_on_hap", C.c_void_p)]


class Timing(C.Structure):
    _fields_ = [("h2d_ms", C.c_float), ("prep_ms", C.c_float), ("k1_ms", C.c_float),
                ("d2h_ms", C.c_float), ("total_ms", C.c_float), ("k1_launches", C.c_int32),
                ("prep_launches", C.c_int32), ("n_pairs", C.c_int64), ("n_cells", C.c_int64),
                ("post_ms", C.c_float), ("host_ms", C.c_float), ("n_ranges", C.c_int32), ("n_chunks", C.c_int32)]


class _PairSumIn(C.Structure):
    _fields_ = [("n_jobs", C.c_int32), ("reserved", C.c_int32), ("job_window", C.c_void_p),
                ("job_group", C.c_void_p), ("job_off", C.c_void_p), ("start", C.c_void_p),
                ("read_group", C.c_void_p)]


class _EmIn(C.Structure):
    _fields_ = [("n_sets", C.c_int32), ("reserved", C.c_int32), ("set_window", C.c_void_p),
                ("set_off", C.c_void_p), ("compatible", C.c_void_p), ("a0", C.c_double), ("tol", C.c_double)]


class _CoverIn(C.Structure):
    _fields_ = [("n_entries", C.c_int32), ("n_keys", C.c_int32), ("maxMismatch", C.c_int32), ("doFilter", C.c_int32),
                ("win_entry", C.c_void_p), ("win_key", C.c_void_p), ("entry_hap", C.c_void_p), ("entry_key", C.c_void_p),
                ("entry_type", C.c_void_p), ("entry_left", C.c_void_p), ("entry_right", C.c_void_p),
                ("read_strand", C.c_void_p)]


class _CoverOut(C.Structure):
    _fields_ = [("hap_filtered", C.c_void_p), ("hap_uncovered", C.c_void_p), ("key_nf", C.c_void_p),
                ("key_nr", C.c_void_p)]


class _CigarIn(C.Structure):
    _fields_ = [("mode", C.c_int32), ("cigar_cap", C.c_int32), ("win_h1", C.c_void_p), ("win_h2", C.c_void_p),
                ("win_ref_start", C.c_void_p), ("hap_n_indels", C.c_void_p), ("hap_ref_pos", C.c_void_p)]


class _CigarOut(C.Structure):
    _fields_ = [("read_hap", C.c_void_p), ("read_ref_pos", C.c_void_p), ("read_n_ops", C.c_void_p), ("ops", C.c_void_p)]


CIGAR_PAIR, CIGAR_BEST = 0, 1
OUT_KEEP_HPOS = 1
VAR_DEL, VAR_INS, VAR_OTHER = 0, 1, 2


class _EmOut(C.Structure):
    _fields_ = [("loglik", C.c_void_p), ("iters", C.c_void_p), ("pi", C.c_void_p), ("freqs", C.c_void_p),
                ("margin", C.c_void_p), ("energy", C.c_void_p)]


class _Job(C.Structure):
    _fields_ = [("in_", C.c_void_p), ("out", C.c_void_p), ("cover", C.c_void_p), ("cover_out", C.c_void_p),
                ("pair_sums", C.c_void_p), ("win_sum_off", C.c_void_p), ("n_groups", C.c_int32), ("reserved", C.c_int32),
                ("read_group", C.c_void_p), ("em", C.c_void_p), ("em_out", C.c_void_p)]


class _StatsIn(C.Structure):
    _fields_ = [("n_jobs", C.c_int32), ("reserved", C.c_int32), ("job_window", C.c_void_p), ("job_mode", C.c_void_p),
                ("job_group", C.c_void_p), ("job_cand", C.c_void_p), ("cand_hap", C.c_void_p), ("cand_start", C.c_void_p),
                ("cand_end", C.c_void_p), ("cand_flags", C.c_void_p), ("read_group", C.c_void_p)]


STATS_DTYPE = np.dtype([
    ("msq", "<f8"), ("allmsq", "<f8"), ("mLogBQ", "<f8"), ("n", "<i4"), ("nf", "<i4"), ("nr", "<i4"), ("numOffBoth", "<i4"),
    ("numIndels", "<i4"), ("nBQT", "<i4"), ("nmmBQT", "<i4"), ("nMMLeft", "<i4"), ("nMMRight", "<i4"),
    ("numUnmappedRealigned", "<i4"), ("n_reads", "<i4"), ("reserved", "<i4")])
assert STATS_DTYPE.itemsize == 72
STATS_SITE, STATS_MAP, STATS_POOL = 0, 1, 2
CAND_PRESENT, CAND_HAS_VAR = 1, 2

PAIR_DTYPE = np.dtype([
    ("ll", "<f8"), ("llOn", "<f8"), ("llOff", "<f8"), ("mLogBQ", "<f8"),
    ("firstBase", "<i4"), ("lastBase", "<i4"),
    ("numIndels", "<u2"), ("numMismatch", "<u2"), ("nBQT", "<u2"), ("nmmBQT", "<u2"),
    ("nMMLeft", "<u2"), ("nMMRight", "<u2"),
    ("offHap", "u1"), ("offHapHMQ", "u1"), ("bMid", "<i2"), ("status", "u1"), ("pad0", "u1"), ("mapState0", "<u2"), ("pad", "u1", (4,)),
])
assert PAIR_DTYPE.itemsize == 64

EXPORTS = [
    "dindel_b200_default_params", "dindel_b200_cli_params", "dindel_b200_abi_version",
    "dindel_b200_strerror", "dindel_b200_last_error", "dindel_b200_create", "dindel_b200_destroy",
    "dindel_b200_count_pairs", "dindel_b200_count_cells", "dindel_b200_submit",
    "dindel_b200_upload", "dindel_b200_run", "dindel_b200_download", "dindel_b200_sync",
    "dindel_b200_get_timing", "dindel_b200_hpos_stride", "dindel_b200_alloc_pinned",
    "dindel_b200_free_pinned", "dindel_b200_measure_fp64_peak", "dindel_b200_pair_sums",
    "dindel_b200_bayes_em", "dindel_b200_plan_shards", "dindel_b200_coverage", "dindel_b200_set_libraries",
    "dindel_b200_last_kernel_ms", "dindel_b200_cigars", "dindel_b200_process", "dindel_b200_count_sums",
    "dindel_b200_sum_error_bound", "dindel_b200_multi_create", "dindel_b200_multi_destroy", "dindel_b200_multi_devices",
    "dindel_b200_multi_set_libraries", "dindel_b200_multi_process", "dindel_b200_multi_get_shard",
    "dindel_b200_multi_last_error", "dindel_b200_load_likelihoods", "dindel_b200_run_static",
    "dindel_b200_static_scratch_bytes", "dindel_b200_read_stats", "dindel_b200_load_records",
]

_lib = None


def load_library():
    """dlopen the CUDA library; raises (no fallback) when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a).  There is no CPU fallback for this path.")
    lib = C.CDLL(LIB_PATH)
    lib.dindel_b200_strerror.restype = C.c_char_p
    lib.dindel_b200_last_error.restype = C.c_char_p
    lib.dindel_b200_last_error.argtypes = [C.c_void_p]
    lib.dindel_b200_create.argtypes = [C.POINTER(Params), C.c_int, C.POINTER(C.c_void_p)]
    lib.dindel_b200_destroy.argtypes = [C.c_void_p]
    lib.dindel_b200_count_pairs.restype = C.c_int64
    lib.dindel_b200_count_pairs.argtypes = [C.POINTER(_BatchIn)]
    lib.dindel_b200_count_cells.restype = C.c_int64
    lib.dindel_b200_count_cells.argtypes = [C.POINTER(_BatchIn)]
    lib.dindel_b200_submit.argtypes = [C.c_void_p, C.POINTER(_BatchIn), C.POINTER(_BatchOut)]
    lib.dindel_b200_upload.argtypes = [C.c_void_p, C.POINTER(_BatchIn), C.c_int]
    lib.dindel_b200_run.argtypes = [C.c_void_p]
    lib.dindel_b200_sync.argtypes = [C.c_void_p]
    lib.dindel_b200_download.argtypes = [C.c_void_p, C.POINTER(_BatchOut)]
    lib.dindel_b200_get_timing.argtypes = [C.c_void_p, C.POINTER(Timing)]
    lib.dindel_b200_last_kernel_ms.argtypes = [C.c_void_p]
    lib.dindel_b200_last_kernel_ms.restype = C.c_float
    lib.dindel_b200_hpos_stride.argtypes = [C.c_void_p]
    lib.dindel_b200_alloc_pinned.restype = C.c_void_p
    lib.dindel_b200_alloc_pinned.argtypes = [C.c_size_t]
    lib.dindel_b200_free_pinned.argtypes = [C.c_void_p]
    lib.dindel_b200_measure_fp64_peak.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
    lib.dindel_b200_pair_sums.argtypes = [C.c_void_p, C.POINTER(_PairSumIn), C.c_void_p]
    lib.dindel_b200_bayes_em.argtypes = [C.c_void_p, C.POINTER(_EmIn), C.POINTER(_EmOut)]
    lib.dindel_b200_coverage.argtypes = [C.c_void_p, C.POINTER(_CoverIn), C.POINTER(_CoverOut)]
    lib.dindel_b200_set_libraries.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.dindel_b200_plan_shards.argtypes = [C.POINTER(_BatchIn), C.c_int, C.c_void_p]
    lib.dindel_b200_process.argtypes = [C.c_void_p, C.POINTER(_Job)]
    lib.dindel_b200_load_likelihoods.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.dindel_b200_count_sums.restype = C.c_int64
    lib.dindel_b200_count_sums.argtypes = [C.POINTER(_BatchIn)]
    lib.dindel_b200_sum_error_bound.restype = C.c_double
    lib.dindel_b200_sum_error_bound.argtypes = [C.c_int64, C.c_double]
    lib.dindel_b200_multi_create.argtypes = [C.POINTER(Params), C.c_void_p, C.c_int, C.POINTER(C.c_void_p)]
    lib.dindel_b200_multi_destroy.argtypes = [C.c_void_p]
    lib.dindel_b200_multi_devices.argtypes = [C.c_void_p]
    lib.dindel_b200_multi_process.argtypes = [C.c_void_p, C.POINTER(_Job)]
    lib.dindel_b200_multi_get_shard.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(Timing)]
    lib.dindel_b200_multi_last_error.restype = C.c_char_p
    lib.dindel_b200_multi_last_error.argtypes = [C.c_void_p]
    lib.dindel_b200_multi_set_libraries.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.dindel_b200_read_stats.argtypes = [C.c_void_p, C.POINTER(_StatsIn), C.c_void_p]
    lib.dindel_b200_load_records.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.dindel_b200_run_static.argtypes = [C.c_void_p, C.POINTER(_BatchOut)]
    lib.dindel_b200_static_scratch_bytes.restype = C.c_size_t
    _lib = lib
    return lib


def plan_shards(batch, n_shards):
    """Window scheduler (dindel_b200_plan_shards): boundaries of n_shards contiguous window ranges of
    near-equal K1 cost; shard k = batch.slice_windows(b[k], b[k+1]).  Host only."""
    out = np.zeros(n_shards + 1, dtype=np.int32)
    cb = batch._c()
    rc = load_library().dindel_b200_plan_shards(C.byref(cb), n_shards, out.ctypes.data)
    if rc != 0:
        raise RuntimeError("dindel_b200_plan_shards failed: %s" % load_library().dindel_b200_strerror(rc).decode())
    return out


def default_params():
    p = Params()
    load_library().dindel_b200_default_params(C.byref(p))
    return p


def cli_params():
    p = Params()
    load_library().dindel_b200_cli_params(C.byref(p))
    return p


class Batch:
    """A batch of realignment windows in the C ABI's SoA/CSR layout (numpy, host memory)."""

    def __init__(self, hap_start, win_hap, win_read, hap_off, hap_seq, read_off, read_seq,
                 read_qual, read_mapq, read_start, read_flags):
        self.hap_start = np.ascontiguousarray(hap_start, dtype=np.uint32)
        self.win_hap = np.ascontiguousarray(win_hap, dtype=np.int32)
        self.win_read = np.ascontiguousarray(win_read, dtype=np.int32)
        self.hap_off = np.ascontiguousarray(hap_off, dtype=np.int64)
        self.hap_seq = np.ascontiguousarray(hap_seq, dtype=np.uint8)
        self.read_off = np.ascontiguousarray(read_off, dtype=np.int64)
        self.read_seq = np.ascontiguousarray(read_seq, dtype=np.uint8)
        self.read_qual = np.ascontiguousarray(read_qual, dtype=np.uint8)
        self.read_mapq = np.ascontiguousarray(read_mapq, dtype=np.uint8)
        self.read_start = np.ascontiguousarray(read_start, dtype=np.uint32)
        self.read_flags = np.ascontiguousarray(read_flags, dtype=np.uint8)

    @property
    def n_windows(self):
        return len(self.hap_start)

    @property
    def n_haps(self):
        return len(self.hap_off) - 1

    @property
    def n_reads(self):
        return len(self.read_off) - 1

    def pair_offsets(self):
        nh = np.diff(self.win_hap).astype(np.int64)
        nr = np.diff(self.win_read).astype(np.int64)
        return np.concatenate([[0], np.cumsum(nh * nr)]).astype(np.int64)

    def n_pairs(self):
        return int(self.pair_offsets()[-1])

    def n_cells(self):
        """sum over pairs of R*(H+2): the GCUPS numerator."""
        cols = (self.hap_off[self.win_hap[1:]] - self.hap_off[self.win_hap[:-1]]) + \
            2 * np.diff(self.win_hap).astype(np.int64)
        rows = self.read_off[self.win_read[1:]] - self.read_off[self.win_read[:-1]]
        return int(np.sum(cols * rows))

    def input_bytes(self):
        return sum(a.nbytes for a in (self.hap_start, self.win_hap, self.win_read, self.hap_off,
                                      self.hap_seq, self.read_off, self.read_seq, self.read_qual,
                                      self.read_mapq, self.read_start, self.read_flags))

    def _c(self, ins=None):
        """ins: optional (read_ins_mode u8, read_ins_anchor i32, read_lib i32) arrays (insert-size prior)."""
        b = _BatchIn()
        b.n_windows, b.n_haps, b.n_reads = self.n_windows, self.n_haps, self.n_reads
        for n in ("hap_start", "win_hap", "win_read", "hap_off", "hap_seq", "read_off", "read_seq",
                  "read_qual", "read_mapq", "read_start", "read_flags"):
            setattr(b, n, getattr(self, n).ctypes.data)
        if ins is not None:
            self._ins = (np.ascontiguousarray(ins[0], dtype=np.uint8), np.ascontiguousarray(ins[1], dtype=np.int32),
                         np.ascontiguousarray(ins[2], dtype=np.int32))
            b.read_ins_mode, b.read_ins_anchor, b.read_lib = (a.ctypes.data for a in self._ins)
        return b

    def window(self, w):
        """(haps list[bytes], reads list[(seq bytes, qual array, mapq, start, flags)], hapStart)."""
        haps = [bytes(self.hap_seq[self.hap_off[h]:self.hap_off[h + 1]])
                for h in range(self.win_hap[w], self.win_hap[w + 1])]
        reads = []
        for r in range(self.win_read[w], self.win_read[w + 1]):
            a, b = self.read_off[r], self.read_off[r + 1]
            reads.append((bytes(self.read_seq[a:b]), self.read_qual[a:b].copy(), int(self.read_mapq[r]),
                          int(self.read_start[r]), int(self.read_flags[r])))
        return haps, reads, int(self.hap_start[w])

    def slice_windows(self, w0, w1):
        """Sub-batch of windows [w0, w1) (copies)."""
        h0, h1 = int(self.win_hap[w0]), int(self.win_hap[w1])
        r0, r1 = int(self.win_read[w0]), int(self.win_read[w1])
        hb0, hb1 = int(self.hap_off[h0]), int(self.hap_off[h1])
        rb0, rb1 = int(self.read_off[r0]), int(self.read_off[r1])
        return Batch(self.hap_start[w0:w1], self.win_hap[w0:w1 + 1] - h0, self.win_read[w0:w1 + 1] - r0,
                     self.hap_off[h0:h1 + 1] - hb0, self.hap_seq[hb0:hb1], self.read_off[r0:r1 + 1] - rb0,
                     self.read_seq[rb0:rb1], self.read_qual[rb0:rb1], self.read_mapq[r0:r1],
                     self.read_start[r0:r1], self.read_flags[r0:r1])

    @staticmethod
    def from_windows(windows):
        """windows: iterable of (hapStart, [hap bytes...], [(seq, qual, mapq, start, flags)...])."""
        hs, wh, wr, ho, hseq, ro, rseq, rq, mq, rs, rf = [], [0], [0], [0], [], [0], [], [], [], [], []
        for hap_start, haps, reads in windows:
            hs.append(hap_start)
            for h in haps:
                hseq.append(np.frombuffer(bytes(h), dtype=np.uint8))
                ho.append(ho[-1] + len(h))
            for seq, qual, mapq, start, flags in reads:
                rseq.append(np.frombuffer(bytes(seq), dtype=np.uint8))
                rq.append(np.asarray(qual, dtype=np.uint8))
                assert len(seq) == len(rq[-1])
                ro.append(ro[-1] + len(seq))
                mq.append(mapq)
                rs.append(start & 0xffffffff)
                rf.append(flags)
            wh.append(len(ho) - 1)
            wr.append(len(ro) - 1)
        cat = lambda l: np.concatenate(l) if l else np.zeros(0, dtype=np.uint8)
        return Batch(hs, wh, wr, ho, cat(hseq), ro, cat(rseq), cat(rq), mq, rs, rf)


class PairResults:
    def __init__(self, pairs, win_pair, win_status, read_on_hap, hpos, timing):
        self.pairs, self.win_pair, self.win_status = pairs, win_pair, win_status
        self.read_on_hap, self.hpos, self.timing = read_on_hap, hpos, timing


class Engine:
    """One context on one device (dindel_b200_create)."""

    def __init__(self, params=None, device=0):
        self.lib = load_library()
        self.params = params if params is not None else cli_params()
        h = C.c_void_p()
        rc = self.lib.dindel_b200_create(C.byref(self.params), device, C.byref(h))
        if rc != 0:
            raise RuntimeError("dindel_b200_create failed: %s" % self.lib.dindel_b200_strerror(rc).decode())
        self.h = h
        self._resident = None

    def close(self):
        if getattr(self, "h", None):
            self.lib.dindel_b200_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc, what):
        if rc != 0:
            raise RuntimeError("%s failed: %s (%s)" % (what, self.lib.dindel_b200_strerror(rc).decode(),
                                                       self.lib.dindel_b200_last_error(self.h).decode()))

    def _alloc_out(self, batch, want_hpos, stride=None):
        n = batch.n_pairs()
        pairs = np.zeros(n, dtype=PAIR_DTYPE)
        win_pair = np.zeros(batch.n_windows + 1, dtype=np.int64)
        win_status = np.zeros(batch.n_windows, dtype=np.int32)
        on_hap = np.zeros(batch.n_reads, dtype=np.uint8)
        hpos = None
        o = _BatchOut()
        o.n_pairs = n
        o.pairs = pairs.ctypes.data
        o.win_pair = win_pair.ctypes.data
        o.win_status = win_status.ctypes.data
        o.read_on_hap = on_hap.ctypes.data
        if want_hpos:
            if stride is None:
                maxr = int(np.max(np.diff(batch.read_off))) if batch.n_reads else 0
                stride = max(8, (maxr + 7) & ~7)
            hpos = np.full((n, stride), -32768, dtype=np.int16)
            o.hpos = hpos.ctypes.data
            o.hpos_stride = stride
        return o, pairs, win_pair, win_status, on_hap, hpos

    def timing(self):
        t = Timing()
        self.lib.dindel_b200_get_timing(self.h, C.byref(t))
        return t

    def set_libraries(self, probs, p95):
        """probs: list of Library::probs arrays; p95: list of getNinetyFifthPctProb() (dindel_b200_set_libraries)."""
        off = np.concatenate([[0], np.cumsum([len(p) for p in probs])]).astype(np.int64)
        flat = np.ascontiguousarray(np.concatenate(probs) if probs else np.zeros(0), dtype=np.float64)
        p95 = np.ascontiguousarray(p95, dtype=np.float64)
        self._check(self.lib.dindel_b200_set_libraries(self.h, len(probs), off.ctypes.data_as(C.c_void_p),
                                                       flat.ctypes.data_as(C.c_void_p), p95.ctypes.data_as(C.c_void_p)),
                    "set_libraries")

    def compute_likelihoods(self, batch, want_hpos=True, keep_hpos=False, ins=None):
        """Host buffers in, host buffers out: dindel_b200_submit.  keep_hpos: the hpos rows are produced and stay
        resident in HBM for coverage() without being copied to the host (DINDEL_OUT_KEEP_HPOS)."""
        o, pairs, win_pair, win_status, on_hap, hpos = self._alloc_out(batch, want_hpos)
        if keep_hpos:
            o.flags = OUT_KEEP_HPOS
        cb = batch._c(ins)
        self._check(self.lib.dindel_b200_submit(self.h, C.byref(cb), C.byref(o)), "submit")
        self._resident, self._n_haps = (batch, want_hpos or keep_hpos), batch.n_haps
        return PairResults(pairs, win_pair, win_status, on_hap, hpos, self.timing())

    # resident-batch API (bench: kernel-only timing with inputs already in HBM)
    def upload(self, batch, want_hpos=False):
        cb = batch._c()
        self._check(self.lib.dindel_b200_upload(self.h, C.byref(cb), int(want_hpos)), "upload")
        self._resident = (batch, want_hpos)

    def run(self, sync=True):
        self._check(self.lib.dindel_b200_run(self.h), "run")
        if sync:
            self._check(self.lib.dindel_b200_sync(self.h), "sync")
        return self.timing()

    def download(self):
        batch, want_hpos = self._resident
        stride = self.lib.dindel_b200_hpos_stride(self.h) if want_hpos else None
        o, pairs, win_pair, win_status, on_hap, hpos = self._alloc_out(batch, want_hpos, stride)
        self._check(self.lib.dindel_b200_download(self.h, C.byref(o)), "download")
        return PairResults(pairs, win_pair, win_status, on_hap, hpos, self.timing())

    def run_static(self):
        """dindel_b200_run_static: K1 over the resident batch with the kernel's scratch in the library's static device
        array (no allocation inside the call), results copied into host arrays allocated HERE before the call."""
        batch, want_hpos = self._resident
        stride = self.lib.dindel_b200_hpos_stride(self.h) if want_hpos else None
        o, pairs, win_pair, win_status, on_hap, hpos = self._alloc_out(batch, want_hpos, stride)
        self._check(self.lib.dindel_b200_run_static(self.h, C.byref(o)), "run_static")
        return PairResults(pairs, win_pair, win_status, on_hap, hpos, self.timing())

    def measure_fp64_peak(self):
        v = C.c_double()
        self._check(self.lib.dindel_b200_measure_fp64_peak(self.h, C.byref(v)), "measure_fp64_peak")
        return v.value

    def load_likelihoods(self, rls, nhs):
        """dindel_b200_load_likelihoods: rls = list of flat r-major ll matrices (one per window), nhs = haplotypes per
        window.  Afterwards pair_sums / bayes_em run on them."""
        wh = np.concatenate([[0], np.cumsum(nhs)]).astype(np.int32)
        wr = np.concatenate([[0], np.cumsum([len(x) // h for x, h in zip(rls, nhs)])]).astype(np.int32)
        ll = np.ascontiguousarray(np.concatenate(rls), dtype=np.float64)
        self._check(self.lib.dindel_b200_load_likelihoods(self.h, len(rls), wh.ctypes.data, wr.ctypes.data, ll.ctypes.data),
                    "load_likelihoods")
        self._resident = None
        self._n_haps, self._n_reads = int(wh[-1]), int(wr[-1])

    def load_records(self, recs, nhs, mapq=None, flags=None):
        """dindel_b200_load_records: recs = list of PAIR_DTYPE arrays (one per window, r-major), nhs = haplotypes per
        window, mapq / flags = per-read bytes over the whole batch."""
        wh = np.concatenate([[0], np.cumsum(nhs)]).astype(np.int32)
        wr = np.concatenate([[0], np.cumsum([len(x) // h for x, h in zip(recs, nhs)])]).astype(np.int32)
        flat = np.ascontiguousarray(np.concatenate(recs), dtype=PAIR_DTYPE)
        mq = None if mapq is None else np.ascontiguousarray(mapq, dtype=np.uint8)
        fl = None if flags is None else np.ascontiguousarray(flags, dtype=np.uint8)
        self._check(self.lib.dindel_b200_load_records(self.h, len(recs), wh.ctypes.data, wr.ctypes.data, flat.ctypes.data,
                                                      mq.ctypes.data if mq is not None else None,
                                                      fl.ctypes.data if fl is not None else None), "load_records")
        self._resident = None
        self._n_haps, self._n_reads = int(wh[-1]), int(wr[-1])

    def read_stats(self, jobs, read_group=None):
        """K6 (dindel_b200_read_stats) over the resident batch.  jobs: list of (window, mode, group, cands) with
        cands = [(hap, startRead, endRead, flags)]; returns a STATS_DTYPE array, one row per job."""
        jw = np.array([j[0] for j in jobs], dtype=np.int32)
        jm = np.array([j[1] for j in jobs], dtype=np.int32)
        jg = np.array([j[2] for j in jobs], dtype=np.int32)
        jc = np.concatenate([[0], np.cumsum([len(j[3]) for j in jobs])]).astype(np.int64)
        flat = [c for j in jobs for c in j[3]]
        ch = np.array([c[0] for c in flat], dtype=np.int32)
        cs = np.array([c[1] for c in flat], dtype=np.int32)
        ce = np.array([c[2] for c in flat], dtype=np.int32)
        cf = np.array([c[3] for c in flat], dtype=np.uint8)
        return self.read_stats_flat(jw, jm, jg, jc, ch, cs, ce, cf, read_group)

    def read_stats_flat(self, jw, jm, jg, jc, ch, cs, ce, cf, read_group=None):
        """The same on the C ABI's own flat arrays (dindel_stats_in)."""
        nj = len(jw)
        jw, jm, jg = (np.ascontiguousarray(x, dtype=np.int32) for x in (jw, jm, jg))
        jc = np.ascontiguousarray(jc, dtype=np.int64)
        ch, cs, ce = (np.ascontiguousarray(x, dtype=np.int32) for x in (ch, cs, ce))
        cf = np.ascontiguousarray(cf, dtype=np.uint8)
        rg = None if read_group is None else np.ascontiguousarray(read_group, dtype=np.int32)
        a = _StatsIn()
        a.n_jobs = nj
        a.job_window, a.job_mode, a.job_group, a.job_cand = jw.ctypes.data, jm.ctypes.data, jg.ctypes.data, jc.ctypes.data
        a.cand_hap, a.cand_start, a.cand_end, a.cand_flags = ch.ctypes.data, cs.ctypes.data, ce.ctypes.data, cf.ctypes.data
        a.read_group = rg.ctypes.data if rg is not None else None
        out = np.zeros(nj, dtype=STATS_DTYPE)
        t0 = time.perf_counter()
        self._check(self.lib.dindel_b200_read_stats(self.h, C.byref(a), out.ctypes.data), "read_stats")
        self.last_call_ms = 1e3 * (time.perf_counter() - t0)
        self.last_kernel_ms = float(self.lib.dindel_b200_last_kernel_ms(self.h))
        return out

    # ---- inference kernels over the resident likelihood matrix ----
    def pair_sums(self, job_window, starts, job_group=None, read_group=None):
        """K2 (diploidGLF / pooled genotype sums).  starts: list of float64 arrays, one per job, each
        nh*(nh+1)/2 long (h1 <= h2 row-major).  Returns the list of summed arrays."""
        nj = len(job_window)
        jw = np.ascontiguousarray(job_window, dtype=np.int32)
        jg = np.full(nj, -1, dtype=np.int32) if job_group is None else np.ascontiguousarray(job_group, dtype=np.int32)
        off = np.concatenate([[0], np.cumsum([len(x) for x in starts])]).astype(np.int64)
        zero_start = all(x is None or not np.any(x) for x in starts) and getattr(self, "null_start", False)
        st = np.ascontiguousarray(np.concatenate(starts) if nj else np.zeros(0), dtype=np.float64)
        out = np.zeros_like(st)
        rg = None if read_group is None else np.ascontiguousarray(read_group, dtype=np.int32)
        a = _PairSumIn()
        a.n_jobs, a.job_window, a.job_group, a.job_off, a.start = nj, jw.ctypes.data, jg.ctypes.data, off.ctypes.data, (None if zero_start else st.ctypes.data)
        a.read_group = rg.ctypes.data if rg is not None else None
        t0 = time.perf_counter()
        self._check(self.lib.dindel_b200_pair_sums(self.h, C.byref(a), out.ctypes.data), "pair_sums")
        self.last_call_ms = 1e3 * (time.perf_counter() - t0)  # the C call alone (H2D, kernel, D2H), no Python packing
        self.last_kernel_ms = float(self.lib.dindel_b200_last_kernel_ms(self.h))
        return [out[off[j]:off[j + 1]] for j in range(nj)]

    def coverage(self, win_entries, win_nkeys, read_strand, max_mismatch=2, do_filter=False):
        """K4 (DetInDel::filterHaplotypes, DInDel.cpp:1932-2100) over the resident batch.  win_entries: per
        window a list of (hap, key, type, left, right) sorted by hap (map order inside a hap); win_nkeys: keys per
        window.  Returns (hap_filtered[n_haps], hap_uncovered[n_haps], [nf arrays per window], [nr arrays])."""
        nw = len(win_entries)
        we = np.concatenate([[0], np.cumsum([len(e) for e in win_entries])]).astype(np.int64)
        wk = np.concatenate([[0], np.cumsum(win_nkeys)]).astype(np.int64)
        flat = [x for e in win_entries for x in e]
        e = np.array(flat, dtype=np.int64).reshape(len(flat), 5) if flat else np.zeros((0, 5), dtype=np.int64)
        eh, ek, el, er = (np.ascontiguousarray(e[:, i], dtype=np.int32) for i in (0, 1, 3, 4))
        et = np.ascontiguousarray(e[:, 2], dtype=np.uint8)
        rs = np.ascontiguousarray(read_strand, dtype=np.uint8)
        batch = self._resident[0] if getattr(self, "_resident", None) else None
        nh_tot = batch.n_haps if batch is not None else int(self._n_haps)
        filt, unc = np.zeros(max(nh_tot, 1), dtype=np.uint8), np.zeros(max(nh_tot, 1), dtype=np.uint8)
        nf, nr = np.zeros(max(int(wk[-1]), 1), dtype=np.int32), np.zeros(max(int(wk[-1]), 1), dtype=np.int32)
        a, o = _CoverIn(), _CoverOut()
        a.n_entries, a.n_keys, a.maxMismatch, a.doFilter = len(flat), int(wk[-1]), int(max_mismatch), int(bool(do_filter))
        a.win_entry, a.win_key = we.ctypes.data, wk.ctypes.data
        a.entry_hap, a.entry_key, a.entry_type = eh.ctypes.data, ek.ctypes.data, et.ctypes.data
        a.entry_left, a.entry_right, a.read_strand = el.ctypes.data, er.ctypes.data, rs.ctypes.data
        o.hap_filtered, o.hap_uncovered, o.key_nf, o.key_nr = filt.ctypes.data, unc.ctypes.data, nf.ctypes.data, nr.ctypes.data
        self._check(self.lib.dindel_b200_coverage(self.h, C.byref(a), C.byref(o)), "coverage")
        return (filt[:nh_tot], unc[:nh_tot], [nf[wk[w]:wk[w + 1]] for w in range(nw)],
                [nr[wk[w]:wk[w + 1]] for w in range(nw)])

    def cigars(self, mode, win_ref_start, hap_ref_pos, win_h1=None, win_h2=None, hap_n_indels=None, cigar_cap=16,
               n_reads=None, out=None):
        """K5 (the haplotype choice of DInDel.cpp:503-509 / :597-606 + DetInDel::getCIGAR, DInDel.cpp:728-882) over
        the resident batch.  hap_ref_pos: flat int32 laid out like hap_seq.  Returns (read_hap[n_reads],
        read_ref_pos[n_reads], read_n_ops[n_reads], ops[n_reads, cigar_cap]) with ops = len << 4 | op.
        out: optional caller-owned (read_hap, read_ref_pos, read_n_ops, ops) arrays of those shapes (e.g. pinned)."""
        batch = self._resident[0] if getattr(self, "_resident", None) else None
        nr = int(n_reads) if n_reads is not None else (batch.n_reads if batch is not None else int(self._n_reads))
        i32 = lambda x: None if x is None else np.ascontiguousarray(x, dtype=np.int32)
        rs, rp, h1, h2, ni = i32(win_ref_start), i32(hap_ref_pos), i32(win_h1), i32(win_h2), i32(hap_n_indels)
        ptr = lambda x: None if x is None else x.ctypes.data
        if out is not None:
            rh, rr, rn, ops = out
            if any(x.dtype != np.int32 or x.size < max(nr, 1) or not x.flags.c_contiguous for x in (rh, rr, rn)) \
                    or ops.dtype != np.uint32 or ops.ndim != 2 or ops.shape[1] != int(cigar_cap) or ops.shape[0] < max(nr, 1) \
                    or not ops.flags.c_contiguous:
                raise ValueError("cigars: out arrays must be contiguous int32[n_reads] x3 and uint32[n_reads, cigar_cap]")
        else:
            rh, rr, rn = (np.zeros(max(nr, 1), dtype=np.int32) for _ in range(3))
            ops = np.zeros((max(nr, 1), int(cigar_cap)), dtype=np.uint32)
        a, o = _CigarIn(), _CigarOut()
        a.mode, a.cigar_cap = int(mode), int(cigar_cap)
        a.win_h1, a.win_h2, a.win_ref_start, a.hap_n_indels, a.hap_ref_pos = ptr(h1), ptr(h2), ptr(rs), ptr(ni), ptr(rp)
        o.read_hap, o.read_ref_pos, o.read_n_ops, o.ops = rh.ctypes.data, rr.ctypes.data, rn.ctypes.data, ops.ctypes.data
        t0 = time.perf_counter()
        self._check(self.lib.dindel_b200_cigars(self.h, C.byref(a), C.byref(o)), "cigars")
        self.last_call_ms = 1e3 * (time.perf_counter() - t0)
        self.last_kernel_ms = float(self.lib.dindel_b200_last_kernel_ms(self.h))
        return rh[:nr], rr[:nr], rn[:nr], ops[:nr]

    def bayes_em(self, set_window, compatible, a0=0.001, tol=1e-4):
        """K3 (estimateHaplotypeFrequenciesBayesEM core).  compatible: list of uint8 arrays (nh each).
        Returns (loglik[n], iters[n], [pi arrays], [freq arrays])."""
        ns = len(set_window)
        sw = np.ascontiguousarray(set_window, dtype=np.int32)
        off = np.concatenate([[0], np.cumsum([len(x) for x in compatible])]).astype(np.int64)
        comp = np.ascontiguousarray(np.concatenate(compatible) if ns else np.zeros(0), dtype=np.uint8)
        ll, it = np.zeros(ns), np.zeros(ns, dtype=np.int32)
        pi, fr = np.zeros(len(comp)), np.zeros(len(comp))
        self.last_em_margin, self.last_em_energy = np.zeros(ns), np.zeros(ns)
        a, o = _EmIn(), _EmOut()
        a.n_sets, a.set_window, a.set_off, a.compatible, a.a0, a.tol = ns, sw.ctypes.data, off.ctypes.data, comp.ctypes.data, a0, tol
        o.loglik, o.iters, o.pi, o.freqs = ll.ctypes.data, it.ctypes.data, pi.ctypes.data, fr.ctypes.data
        o.margin, o.energy = self.last_em_margin.ctypes.data, self.last_em_energy.ctypes.data
        t0 = time.perf_counter()
        self._check(self.lib.dindel_b200_bayes_em(self.h, C.byref(a), C.byref(o)), "bayes_em")
        self.last_call_ms = 1e3 * (time.perf_counter() - t0)
        self.last_kernel_ms = float(self.lib.dindel_b200_last_kernel_ms(self.h))
        return ll, it, [pi[off[j]:off[j + 1]] for j in range(ns)], [fr[off[j]:off[j + 1]] for j in range(ns)]


class JobResults:
    """What one dindel_b200_process / dindel_b200_multi_process call returned."""
    pass


def _build_job(batch, want_pairs, want_hpos, cover, pair_sums, n_groups, read_group, em, ins=None, pinned=None):
    """Allocates the outputs of a job and fills the C structs.  cover: (win_entries, win_nkeys, read_strand,
    max_mismatch, do_filter) as Engine.coverage takes them; em: (set_window, [compatible arrays], a0, tol).
    Returns (job struct, keep-alive list, JobResults with numpy views)."""
    keep, res = [], JobResults()
    lib = load_library()
    cb = batch._c(ins)
    keep.append(cb)
    n = batch.n_pairs()
    o = _BatchOut()
    res.win_pair = np.zeros(batch.n_windows + 1, dtype=np.int64)
    res.win_status = np.zeros(max(batch.n_windows, 1), dtype=np.int32)
    res.read_on_hap = np.zeros(max(batch.n_reads, 1), dtype=np.uint8)
    res.pairs = res.hpos = None
    if want_pairs:
        res.pairs = np.zeros(n, dtype=PAIR_DTYPE) if pinned is None else pinned
        o.n_pairs, o.pairs = n, res.pairs.ctypes.data
    if want_hpos:
        maxr = int(np.max(np.diff(batch.read_off))) if batch.n_reads else 0
        stride = max(8, (maxr + 7) & ~7)
        res.hpos = np.full((n, stride), -32768, dtype=np.int16)
        o.hpos, o.hpos_stride = res.hpos.ctypes.data, stride
    o.win_pair, o.win_status, o.read_on_hap = res.win_pair.ctypes.data, res.win_status.ctypes.data, res.read_on_hap.ctypes.data
    keep.append(o)
    job = _Job()
    job.in_, job.out = C.addressof(cb), C.addressof(o)
    if cover is not None:
        win_entries, win_nkeys, read_strand, max_mismatch, do_filter = cover
        we = np.concatenate([[0], np.cumsum([len(e) for e in win_entries])]).astype(np.int64)
        wk = np.concatenate([[0], np.cumsum(win_nkeys)]).astype(np.int64)
        flat = [x for e in win_entries for x in e]
        e = np.array(flat, dtype=np.int64).reshape(len(flat), 5) if flat else np.zeros((0, 5), dtype=np.int64)
        eh, ek, el, er = (np.ascontiguousarray(e[:, i], dtype=np.int32) for i in (0, 1, 3, 4))
        et = np.ascontiguousarray(e[:, 2], dtype=np.uint8)
        rs = np.ascontiguousarray(read_strand, dtype=np.uint8)
        res.hap_filtered, res.hap_uncovered = np.zeros(max(batch.n_haps, 1), dtype=np.uint8), np.zeros(max(batch.n_haps, 1), dtype=np.uint8)
        res.key_nf, res.key_nr = np.zeros(max(int(wk[-1]), 1), dtype=np.int32), np.zeros(max(int(wk[-1]), 1), dtype=np.int32)
        res.win_key = wk
        a, co = _CoverIn(), _CoverOut()
        a.n_entries, a.n_keys, a.maxMismatch, a.doFilter = len(flat), int(wk[-1]), int(max_mismatch), int(bool(do_filter))
        a.win_entry, a.win_key = we.ctypes.data, wk.ctypes.data
        a.entry_hap, a.entry_key, a.entry_type = eh.ctypes.data, ek.ctypes.data, et.ctypes.data
        a.entry_left, a.entry_right, a.read_strand = el.ctypes.data, er.ctypes.data, rs.ctypes.data
        co.hap_filtered, co.hap_uncovered, co.key_nf, co.key_nr = (res.hap_filtered.ctypes.data, res.hap_uncovered.ctypes.data,
                                                                  res.key_nf.ctypes.data, res.key_nr.ctypes.data)
        keep += [we, wk, eh, ek, el, er, et, rs, a, co]
        job.cover, job.cover_out = C.addressof(a), C.addressof(co)
    if pair_sums:
        ns = int(lib.dindel_b200_count_sums(C.byref(cb)))
        g = max(int(n_groups), 1)
        res.pair_sums = np.zeros(max(ns * g, 1), dtype=np.float64)
        res.win_sum_off = np.zeros(batch.n_windows + 1, dtype=np.int64)
        job.pair_sums, job.win_sum_off, job.n_groups = res.pair_sums.ctypes.data, res.win_sum_off.ctypes.data, int(n_groups)
        if n_groups:
            rg = np.ascontiguousarray(read_group, dtype=np.int32)
            keep.append(rg)
            job.read_group = rg.ctypes.data
    if em is not None:
        set_window, compatible, a0, tol = em
        nsets = len(set_window)
        sw = np.ascontiguousarray(set_window, dtype=np.int32)
        off = np.concatenate([[0], np.cumsum([len(x) for x in compatible])]).astype(np.int64)
        comp = np.ascontiguousarray(np.concatenate(compatible) if nsets else np.zeros(0), dtype=np.uint8)
        res.em_loglik, res.em_iters = np.zeros(max(nsets, 1)), np.zeros(max(nsets, 1), dtype=np.int32)
        res.em_pi, res.em_freqs = np.zeros(max(len(comp), 1)), np.zeros(max(len(comp), 1))
        res.em_margin, res.em_energy, res.em_off = np.zeros(max(nsets, 1)), np.zeros(max(nsets, 1)), off
        a, eo = _EmIn(), _EmOut()
        a.n_sets, a.set_window, a.set_off, a.compatible, a.a0, a.tol = nsets, sw.ctypes.data, off.ctypes.data, comp.ctypes.data, a0, tol
        eo.loglik, eo.iters, eo.pi, eo.freqs = res.em_loglik.ctypes.data, res.em_iters.ctypes.data, res.em_pi.ctypes.data, res.em_freqs.ctypes.data
        eo.margin, eo.energy = res.em_margin.ctypes.data, res.em_energy.ctypes.data
        keep += [sw, off, comp, a, eo]
        job.em, job.em_out = C.addressof(a), C.addressof(eo)
    return job, keep, res


def _engine_process(self, batch, want_pairs=True, want_hpos=False, cover=None, pair_sums=False, n_groups=0,
                    read_group=None, em=None, ins=None):
    """dindel_b200_process: K1 and, on the still-resident results, K4 / K2 / K3, streamed through the device in
    ranges of windows (any batch size).  See _build_job for the arguments."""
    job, keep, res = _build_job(batch, want_pairs, want_hpos, cover, pair_sums, n_groups, read_group, em, ins)
    self._check(self.lib.dindel_b200_process(self.h, C.byref(job)), "process")
    res.timing = self.timing()
    self._resident, self._n_haps = (batch, want_hpos or cover is not None), batch.n_haps
    return res


Engine.process = _engine_process


def sum_error_bound(n_terms, magnitude):
    return float(load_library().dindel_b200_sum_error_bound(int(n_terms), float(magnitude)))


class MultiEngine:
    """The window scheduler over several GPUs of one box (dindel_b200_multi_*): one context + one host worker
    thread per device inside the library, windows sharded by cost, results in window order."""

    def __init__(self, params=None, devices=(0,)):
        self.lib = load_library()
        self.params = params if params is not None else cli_params()
        dev = np.ascontiguousarray(devices, dtype=np.int32)
        h = C.c_void_p()
        rc = self.lib.dindel_b200_multi_create(C.byref(self.params), dev.ctypes.data, len(dev), C.byref(h))
        if rc != 0:
            raise RuntimeError("dindel_b200_multi_create failed: %s" % self.lib.dindel_b200_strerror(rc).decode())
        self.h, self.n_devices = h, len(dev)

    def close(self):
        if getattr(self, "h", None):
            self.lib.dindel_b200_multi_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def process(self, batch, want_pairs=True, want_hpos=False, cover=None, pair_sums=False, n_groups=0, read_group=None,
                em=None, ins=None, pinned=None):
        job, keep, res = _build_job(batch, want_pairs, want_hpos, cover, pair_sums, n_groups, read_group, em, ins, pinned)
        t0 = time.perf_counter()
        rc = self.lib.dindel_b200_multi_process(self.h, C.byref(job))
        res.wall_ms = 1e3 * (time.perf_counter() - t0)
        if rc != 0:
            raise RuntimeError("multi_process failed: %s (%s)" % (self.lib.dindel_b200_strerror(rc).decode(),
                                                                  self.lib.dindel_b200_multi_last_error(self.h).decode()))
        res.shards = []
        for k in range(self.n_devices):
            a, b, t = C.c_int32(), C.c_int32(), Timing()
            self.lib.dindel_b200_multi_get_shard(self.h, k, C.byref(a), C.byref(b), C.byref(t))
            res.shards.append((a.value, b.value, t))
        return res
