"""oracle/pyoracle.py — TEST INFRASTRUCTURE ONLY: ctypes loader for the two CPU checkers.

  Oracle("port")      -> oracle/libdindel_oracle.so    (plain-C restatement, dindel_oracle.c)
  Oracle("reference") -> oracle/_ref/libdindel_ref.so  (the reference's own ObservationModelFB.cpp)
Both export the same entry points (prefix ora_ / ref_) and the same POD layouts.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_SO = os.path.join(HERE, "libdindel_oracle.so")
REF_SO = os.path.join(HERE, "_ref", "libdindel_ref.so")


class OraParams(C.Structure):
    _fields_ = [(n, C.c_double) for n in
                ("pError", "pMut", "mapQualThreshold", "pFirstgLO", "checkBaseQualThreshold")] + \
               [(n, C.c_int32) for n in ("maxLengthDel", "padCover", "bMid", "forceReadOnHaplotype")]


ORA_DTYPE = np.dtype([
    ("ll", "<f8"), ("llOn", "<f8"), ("llOff", "<f8"), ("mLogBQ", "<f8"),
    ("firstBase", "<i4"), ("lastBase", "<i4"), ("numIndels", "<i4"), ("numMismatch", "<i4"),
    ("nBQT", "<i4"), ("nmmBQT", "<i4"), ("nMMLeft", "<i4"), ("nMMRight", "<i4"),
    ("offHap", "<i4"), ("offHapHMQ", "<i4"), ("bMid", "<i4"), ("status", "<i4")])
assert ORA_DTYPE.itemsize == 80


STATS_DTYPE = np.dtype([
    ("msq", "<f8"), ("allmsq", "<f8"), ("mLogBQ", "<f8"), ("n", "<i4"), ("nf", "<i4"), ("nr", "<i4"), ("numOffBoth", "<i4"),
    ("numIndels", "<i4"), ("nBQT", "<i4"), ("nmmBQT", "<i4"), ("nMMLeft", "<i4"), ("nMMRight", "<i4"),
    ("numUnmappedRealigned", "<i4"), ("n_reads", "<i4"), ("reserved", "<i4")])
assert STATS_DTYPE.itemsize == 72


def build(which="port"):
    subprocess.check_call(["make", "-s", "-C", HERE, "oracle" if which == "port" else "ref"])


def params_from(p):
    """engine.Params (or any object with the same fields) -> OraParams."""
    o = OraParams()
    for n, _ in OraParams._fields_:
        setattr(o, n, getattr(p, n))
    return o


class Oracle:
    def __init__(self, kind="port"):
        self.kind = kind
        path = PORT_SO if kind == "port" else REF_SO
        if kind == "port" and not os.path.exists(path):
            build("port")
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        self.lib = C.CDLL(path)
        pre = "ora_" if kind == "port" else "ref_"
        self._pair = getattr(self.lib, pre + "pair")
        self._window = getattr(self.lib, pre + "window")
        self._p2p = getattr(self.lib, pre + "phred_to_prob")
        self._p2p.restype = C.c_double
        assert getattr(self.lib, pre + "sizeof_params")() == C.sizeof(OraParams)
        assert getattr(self.lib, pre + "sizeof_pair_out")() == ORA_DTYPE.itemsize

    def phred_to_prob(self, ph):
        return self._p2p(int(ph))

    def pair(self, hap, read, qual, map_qual, pos_stat_first, unmapped, hap_start, params):
        """qual: probabilities (float64 array); map_qual: probability."""
        hap, read = bytes(hap), bytes(read)
        R = len(read)
        q = np.ascontiguousarray(qual, dtype=np.float64)
        out = np.zeros(1, dtype=ORA_DTYPE)
        hpos = np.zeros(max(R, 1), dtype=np.int32)
        ms = np.zeros(max(R, 1), dtype=np.int32)
        self._pair(hap, len(hap), read, R, q.ctypes.data_as(C.c_void_p), C.c_double(map_qual),
                   C.c_double(pos_stat_first), int(unmapped), C.c_uint32(hap_start & 0xffffffff),
                   C.byref(params), out.ctypes.data_as(C.c_void_p), hpos.ctypes.data_as(C.c_void_p),
                   ms.ctypes.data_as(C.c_void_p))
        return out[0], hpos[:R], ms[:R]

    def library(self, counts):
        """reference only: Library(counts) -> (probs[maxins], p95)   (Library.hpp:31-129)"""
        counts = np.ascontiguousarray(counts, dtype=np.float64)
        probs = np.zeros(len(counts), dtype=np.float64)
        p95 = C.c_double()
        n = self.lib.ref_library(counts.ctypes.data_as(C.c_void_p), len(counts), probs.ctypes.data_as(C.c_void_p), C.byref(p95))
        return probs[:n].copy(), p95.value

    def pair_ins(self, hap, read, qual, map_qual, pos_stat_first, unmapped, hap_start, params, ins_mode, ins_anchor,
                 probs=None, p95=None, counts=None):
        """pair() with the insert-size prior (ObservationModelFB.cpp:279-293).  The port takes the library's probs
        and p95; the reference builds its own Library from the raw histogram `counts`."""
        hap, read = bytes(hap), bytes(read)
        R = len(read)
        q = np.ascontiguousarray(qual, dtype=np.float64)
        out = np.zeros(1, dtype=ORA_DTYPE)
        hpos = np.zeros(max(R, 1), dtype=np.int32)
        ms = np.zeros(max(R, 1), dtype=np.int32)
        vp = lambda a: a.ctypes.data_as(C.c_void_p)
        if self.kind == "port":
            pr = np.ascontiguousarray(probs, dtype=np.float64)
            self.lib.ora_pair_ins(hap, len(hap), read, R, vp(q), C.c_double(map_qual), C.c_double(pos_stat_first),
                                  int(unmapped), C.c_uint32(hap_start & 0xffffffff), C.byref(params), vp(out), vp(hpos),
                                  vp(ms), int(ins_mode), int(ins_anchor), vp(pr), len(pr), C.c_double(p95))
        else:
            cn = np.ascontiguousarray(counts, dtype=np.float64)
            self.lib.ref_pair_ins(hap, len(hap), read, R, vp(q), C.c_double(map_qual), C.c_double(pos_stat_first),
                                  int(unmapped), C.c_uint32(hap_start & 0xffffffff), C.byref(params), vp(out), vp(hpos),
                                  vp(ms), int(ins_mode), int(ins_anchor), vp(cn), len(cn))
        return out[0], hpos[:R], ms[:R]

    def window(self, batch, w, params, want_hpos=True):
        """All pairs of window w of an engine.Batch -> (records r-major [nr*nh], hpos [nr*nh, stride])."""
        h0, h1 = int(batch.win_hap[w]), int(batch.win_hap[w + 1])
        r0, r1 = int(batch.win_read[w]), int(batch.win_read[w + 1])
        nh, nr = h1 - h0, r1 - r0
        hb, rb = int(batch.hap_off[h0]), int(batch.read_off[r0])
        hap_off = (batch.hap_off[h0:h1 + 1] - hb).astype(np.int32)
        read_off = (batch.read_off[r0:r1 + 1] - rb).astype(np.int32)
        hap_seq = np.ascontiguousarray(batch.hap_seq[hb:int(batch.hap_off[h1])])
        read_seq = np.ascontiguousarray(batch.read_seq[rb:int(batch.read_off[r1])])
        read_qual = np.ascontiguousarray(batch.read_qual[rb:int(batch.read_off[r1])])
        mapq = np.ascontiguousarray(batch.read_mapq[r0:r1])
        start = np.ascontiguousarray(batch.read_start[r0:r1])
        flags = np.ascontiguousarray(batch.read_flags[r0:r1])
        out = np.zeros(nh * nr, dtype=ORA_DTYPE)
        stride = int(np.max(np.diff(read_off))) if nr else 1
        stride = max(stride, 1)
        hpos = np.full((max(nh * nr, 1), stride), -32768, dtype=np.int32) if want_hpos else None
        vp = lambda a: a.ctypes.data_as(C.c_void_p)
        n = self._window(vp(hap_seq), vp(hap_off), nh, vp(read_seq), vp(read_qual), vp(read_off), vp(mapq),
                         vp(start), vp(flags), nr, C.c_uint32(int(batch.hap_start[w])), C.byref(params),
                         vp(out), vp(hpos) if want_hpos else None, stride)
        return n, out, hpos


    # ---- K2 / K3 restatements (port only) ----
    def pair_sums(self, rl, nh, start, read_group=None, group=-1):
        rl = np.ascontiguousarray(rl, dtype=np.float64)
        nr = len(rl) // nh
        start = np.ascontiguousarray(start, dtype=np.float64)
        out = np.zeros(nh * (nh + 1) // 2, dtype=np.float64)
        rg = None if read_group is None else np.ascontiguousarray(read_group, dtype=np.int32)
        vp = lambda a: a.ctypes.data_as(C.c_void_p)
        self.lib.ora_pair_sums(vp(rl), nr, nh, vp(rg) if rg is not None else None, int(group), vp(start), vp(out))
        return out

    def bayes_em(self, rl, nh, compatible, a0=0.001, tol=1e-4):
        rl = np.ascontiguousarray(rl, dtype=np.float64)
        nr = len(rl) // nh
        comp = np.ascontiguousarray(compatible, dtype=np.uint8)
        ll = C.c_double()
        pi = np.zeros(nh)
        fr = np.zeros(nh)
        vp = lambda a: a.ctypes.data_as(C.c_void_p)
        it = self.lib.ora_bayes_em(vp(rl), nr, nh, vp(comp), C.c_double(a0), C.c_double(tol), C.byref(ll), vp(pi), vp(fr))
        return ll.value, it, pi, fr

    # ---- K6: read statistics of a called site (port only) ----
    def read_stats(self, mode, recs, nh, map_qual, reverse, unmapped, cands, pad, read_group=None, group=-1):
        """recs: ORA_DTYPE records r-major [nr*nh]; cands: list of (hap, startRead, endRead, flags)."""
        recs = np.ascontiguousarray(recs, dtype=ORA_DTYPE)
        nr = len(recs) // nh
        mqv = np.ascontiguousarray(map_qual, dtype=np.float64)
        rev = np.ascontiguousarray(reverse, dtype=np.uint8)
        unm = np.ascontiguousarray(unmapped, dtype=np.uint8)
        rg = None if read_group is None else np.ascontiguousarray(read_group, dtype=np.int32)
        ch = np.array([c[0] for c in cands], dtype=np.int32)
        cs = np.array([c[1] for c in cands], dtype=np.int32)
        ce = np.array([c[2] for c in cands], dtype=np.int32)
        cf = np.array([c[3] for c in cands], dtype=np.uint8)
        out = np.zeros(1, dtype=STATS_DTYPE)
        vp = lambda a: a.ctypes.data_as(C.c_void_p)
        self.lib.ora_read_stats(int(mode), int(nh), int(nr), vp(recs), vp(mqv), vp(rev), vp(unm), vp(rg) if rg is not None else None,
                                int(group), len(cands), vp(ch), vp(cs), vp(ce), vp(cf), int(pad), vp(out))
        return out[0]

    def addlogs(self, a, b):
        self.lib.ora_addlogs.restype = C.c_double
        return self.lib.ora_addlogs(C.c_double(a), C.c_double(b))

    def digamma(self, x):
        self.lib.ora_digamma.restype = C.c_double
        return self.lib.ora_digamma(C.c_double(x))

    # ---- a13: filterHaplotypes restatement (port only) ----
    def filter_haplotypes(self, haps, reads, hpos, off_hap_hmq, num_indels, strand, entries, nk, max_mismatch,
                          do_filter):
        """One window.  haps/reads: lists of bytes; hpos: int32 [nr*nh, stride] (pair r*nh+h); off_hap_hmq /
        num_indels: [nr*nh]; strand: [nr]; entries: list of (hap, key, type, left, right) sorted by hap.
        -> (filtered[nh], uncovered[nh], nf[nk], nr[nk])"""
        nh, nr = len(haps), len(reads)
        hap_off = np.concatenate([[0], np.cumsum([len(h) for h in haps])]).astype(np.int32)
        read_off = np.concatenate([[0], np.cumsum([len(r) for r in reads])]).astype(np.int32)
        hs, rs = b"".join(haps) + b"\0", b"".join(reads) + b"\0"
        hpos = np.ascontiguousarray(hpos, dtype=np.int32)
        ohq = np.ascontiguousarray(off_hap_hmq, dtype=np.uint8)
        ni = np.ascontiguousarray(num_indels, dtype=np.int32)
        st = np.ascontiguousarray(strand, dtype=np.uint8)
        ne = len(entries)
        e = np.array(entries, dtype=np.int64).reshape(ne, 5) if ne else np.zeros((0, 5), dtype=np.int64)
        eh, ek, el, er = (np.ascontiguousarray(e[:, i], dtype=np.int32) for i in (0, 1, 3, 4))
        et = np.ascontiguousarray(e[:, 2], dtype=np.uint8)
        filt, unc = np.zeros(max(nh, 1), dtype=np.uint8), np.zeros(max(nh, 1), dtype=np.uint8)
        nf, nrv = np.zeros(max(nk, 1), dtype=np.int32), np.zeros(max(nk, 1), dtype=np.int32)
        vp = lambda a: a.ctypes.data_as(C.c_void_p)
        self.lib.ora_filter_haplotypes(nh, nr, hs, vp(hap_off), rs, vp(read_off), vp(hpos), int(hpos.shape[1]) if hpos.ndim == 2 else 1,
                                       vp(ohq), vp(ni), vp(st), ne, vp(eh), vp(ek), vp(et), vp(el), vp(er), int(nk),
                                       int(max_mismatch), int(do_filter), vp(filt), vp(unc), vp(nf), vp(nrv))
        return filt[:nh], unc[:nh], nf[:nk], nrv[:nk]

    def choose_hap(self, mode, ll, n_indels, h1=-1, h2=-1):
        """Haplotype a read is realigned to (DInDel.cpp:597-606 mode 0 / :503-509 mode 1)."""
        ll = np.ascontiguousarray(ll, dtype=np.float64)
        ni = np.ascontiguousarray(n_indels, dtype=np.int32)
        return int(self.lib.ora_choose_hap(int(mode), len(ll), ll.ctypes.data_as(C.c_void_p), ni.ctypes.data_as(C.c_void_p),
                                           int(h1), int(h2)))

    def get_cigar(self, hpos, hml, ref_start, cap=64):
        """DetInDel::getCIGAR (DInDel.cpp:728-882) -> (n_ops or -error, refPos, [(op, len), ...])."""
        hpos = np.ascontiguousarray(hpos, dtype=np.int32)
        hml = np.ascontiguousarray(hml, dtype=np.int32)
        ops = np.zeros(2 * cap, dtype=np.int32)
        ref_pos = C.c_int32(0)
        n = int(self.lib.ora_get_cigar(len(hpos), hpos.ctypes.data_as(C.c_void_p), len(hml), hml.ctypes.data_as(C.c_void_p),
                                       int(ref_start), C.byref(ref_pos), ops.ctypes.data_as(C.c_void_p), cap))
        k = max(0, min(n, cap))
        return n, int(ref_pos.value), [(int(ops[2 * i]), int(ops[2 * i + 1])) for i in range(k)]
