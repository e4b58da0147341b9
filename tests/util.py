"""Shared helpers for the parity tests."""
import os

import numpy as np

from dindel_tgi_b200 import engine

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "k1_golden.npz")
FIELDS = ("hap_start", "win_hap", "win_read", "hap_off", "hap_seq", "read_off", "read_seq", "read_qual",
          "read_mapq", "read_start", "read_flags")
# ObservationModel.hpp:39-62 (struct defaults) and DInDel.cpp:4122-4157 (CLI defaults)
PSETS = {"cli": (5e-4, 1e-5, 100.0, 0.01, 0.95, 5, 2, -1, 0),
         "default": (1e-4, 1e-4, 100.0, 0.01, 0.95, 10, 5, -1, 0)}
CMP_FIELDS = ("ll", "llOn", "llOff", "mLogBQ", "firstBase", "lastBase", "numIndels", "numMismatch", "nBQT",
              "nmmBQT", "nMMLeft", "nMMRight", "offHap", "offHapHMQ", "bMid", "status")


def make_params(cls, pset):
    return cls(*PSETS[pset]) if isinstance(pset, str) else cls(*pset)


def golden_cases():
    z = np.load(GOLDEN)
    return sorted({k.split("/")[0] for k in z.files})


def load_golden(name):
    z = np.load(GOLDEN)
    b = engine.Batch(*[z[f"{name}/{f}"] for f in FIELDS])
    return b, str(z[f"{name}/pset"]), z[f"{name}/records"], z[f"{name}/hpos"]


def assert_records_equal(got, want, what=""):
    """Bit-exact comparison of pair records (GPU 64-byte records or oracle 80-byte records)."""
    assert len(got) == len(want), f"{what}: {len(got)} vs {len(want)} pairs"
    for f in CMP_FIELDS:
        a, b = np.asarray(got[f]), np.asarray(want[f])
        if a.dtype.kind == "f":
            same = a.view(np.int64) == b.astype(np.float64).view(np.int64)
        else:
            same = a.astype(np.int64) == b.astype(np.int64)
        if not same.all():
            i = int(np.nonzero(~same)[0][0])
            raise AssertionError(f"{what}: field {f} differs at pair {i}: got {a[i]!r} want {b[i]!r} "
                                 f"({int((~same).sum())} of {len(a)} pairs differ)")


def assert_hpos_equal(got, want, batch, what=""):
    """hpos rows compared over each read's own length."""
    lens = np.diff(batch.read_off)
    nh = np.diff(batch.win_hap)
    nr = np.diff(batch.win_read)
    # read length of every pair, r-major inside each window
    plen = np.concatenate([np.repeat(lens[batch.win_read[w]:batch.win_read[w + 1]], nh[w])
                           for w in range(batch.n_windows)]) if batch.n_windows else np.zeros(0, int)
    assert len(plen) == len(got) == len(want), what
    stride = min(got.shape[1], want.shape[1])
    assert plen.max(initial=0) <= stride
    mask = np.arange(stride)[None, :] < plen[:, None]
    diff = (got[:, :stride].astype(np.int32) != want[:, :stride].astype(np.int32)) & mask
    if diff.any():
        i = int(np.nonzero(diff.any(axis=1))[0][0])
        raise AssertionError(f"{what}: hpos differs at pair {i}: got {got[i, :plen[i]]} want {want[i, :plen[i]]}")


def oracle_batch(oracle, batch, params):
    """Run a CPU checker over every window of a batch -> (records, hpos) in the engine's pair order."""
    recs, hps = [], []
    for w in range(batch.n_windows):
        n, o, hp = oracle.window(batch, w, params)
        recs.append(o)
        hps.append(hp[:len(o)])
    if not recs:
        return np.zeros(0, dtype=recs_dtype()), np.zeros((0, 1), np.int32)
    stride = max(h.shape[1] for h in hps)
    allh = np.full((sum(len(o) for o in recs), stride), -32768, dtype=np.int32)
    i = 0
    for o, h in zip(recs, hps):
        allh[i:i + len(o), :h.shape[1]] = h
        i += len(o)
    return np.concatenate(recs), allh


def recs_dtype():
    from oracle.pyoracle import ORA_DTYPE
    return ORA_DTYPE


def load_cover_golden():
    """tests/golden/cover_golden.json.gz (reference's own filterHaplotypes, make_cover_golden.py) as a list of
    windows: dict(haps, reads, strand, hpos [nr*nh, stride] int32 (r-major), ohq, ni, entries [(h, key id, type,
    left, right)], keys [str], nk, maxMismatch, doFilter, filtered, cov {key str: (nf, nr)})."""
    import gzip
    import json
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "cover_golden.json.gz")
    out = []
    for w in json.load(gzip.open(path, "rt")):
        nh, nr = w["nh"], w["nr"]
        stride = max(len(r) for r in w["reads"])
        hpos = np.full((nr * nh, stride), -32768, dtype=np.int32)
        ohq, ni = np.zeros(nr * nh, dtype=np.uint8), np.zeros(nr * nh, dtype=np.int32)
        for p in w["pairs"]:
            h, r = p[0], p[1]
            ohq[r * nh + h], ni[r * nh + h] = p[2], p[3]
            hpos[r * nh + h, :len(p) - 4] = p[4:]
        keys, entries = [], []
        for h, key, ty, left, right in w["entries"]:   # already sorted by hap, map order inside a hap
            if key not in keys:
                keys.append(key)
            entries.append((h, keys.index(key), ty, left, right))
        out.append(dict(haps=[s.encode() for s in w["haps"]], reads=[s.encode() for s in w["reads"]],
                        strand=np.array(w["strand"], dtype=np.uint8), hpos=hpos, ohq=ohq, ni=ni, entries=entries,
                        keys=keys, nk=len(keys), maxMismatch=w["maxMismatch"], doFilter=w["doFilter"],
                        filtered=np.array(w["filtered"], dtype=np.uint8), cov={k: (a, b) for k, a, b in w["cov"]}))
    return out


def load_cigar_golden():
    """tests/golden/cigar_golden.json.gz (the reference's own getCIGAR call sites, make_cigar_golden.py): list of
    per-read dicts mode ('D' diploid / 'B' pooled), h1, h2, nh, onHap, hmax, refStart, refPos, ll[nh], nInd[nh],
    hpos[R], hml[H], cigar [op, len, op, len, ...]."""
    import gzip
    import json
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "cigar_golden.json.gz")
    return json.load(gzip.open(path, "rt"))
