#!/usr/bin/env python
"""bench.py — read-haplotype likelihoods/s and HMM GCUPS of the K1 path on B200 (BASELINE.json).

A "step" is one pass of the hot path (DetInDel::computeLikelihoods for every window) over one
batch of synthetic windows of BASELINE configs[1]: 100k diploid windows, 30x 100-bp reads, up to 8
candidate haplotypes (dindel_tgi_b200/synth.py, cfg 2).  Per GPU the batch is fixed, so N GPUs
process N shards ("weak" scaling, no data-path collective: windows are independent).

  value : pairs/s with the batch already resident in HBM (dindel_b200_upload once, then
          dindel_b200_run per step), device time from CUDA events on the library's stream.
  e2e   : the same through the C ABI with pinned HOST buffers, every step ONE dindel_b200_process call: H2D of every
          input array, K1, K4 (filterHaplotypes on the hpos rows where K1 left them in HBM) and K2 (the read sums of
          diploidGLF for every haplotype pair), D2H of the 64-byte pair records, filtered[] / coverage counts and the
          sums: the diploid hot path as the patched reference driver runs it per batch of windows.
          --hpos-d2h copies the hpos rows to the host instead of running K4 on them.
  extra_configs : configs[2] (pooled 1000x, 16 haplotypes: K1 + K3 EM + per-pool K2) and configs[3] (250-bp reads, 32
          haplotypes, maxLengthDel 10) measured the same way at a size that keeps the default run within minutes,
          each with a parity gate (sampled windows bit-equal to oracle/_ref, outside the timed region).
  strong : rank 0 drives every visible GPU through dindel_b200_multi_process (the in-process window scheduler) on ONE
          fixed job (the rank-0 batch tiled --strong-tiles times), at 1 device and at N: wall time, pairs/s, efficiency
          and the per-device H2D / kernel / D2H / host split; buffers byte-identical between the two.
  roofline : K1 is bound by the FP64 add/compare issue rate, not HBM: achieved =
          cells/s x W_cell (5J+10 FP64 pipe ops per cell), peak = DADD rate measured on this GPU.
  cpu_baseline / --impl reference : the reference's own ObservationModelFBMaxErr (oracle/_ref,
          built from /root/reference) -- or the C port when that is absent -- on the host cores,
          windows split round-robin across one process per core like makeWindows.py's file split.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "read-haplotype likelihoods/s"
UNIT = "pairs/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--windows", type=int, default=100000, help="windows per GPU (configs[1]: 100k)")
    ap.add_argument("--cfg", type=int, default=2)
    ap.add_argument("--no-hpos", action="store_true", help="no per-base hpos rows at all (K1 only, no K4)")
    ap.add_argument("--hpos-d2h", action="store_true",
                    help="e2e copies the hpos rows to the host (host-side filterHaplotypes) instead of running K4 on them")
    ap.add_argument("--cpu-windows", type=int, default=0, help="cpu_baseline sample (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip extra_configs (configs[2] / configs[3])")
    ap.add_argument("--extra-windows", default="2000,10000", help="windows of the configs[2], configs[3] extra runs")
    ap.add_argument("--no-strong", action="store_true", help="skip the strong-scaling sub-record (multi-device call)")
    ap.add_argument("--strong-tiles", type=int, default=3, help="strong job = the rank-0 batch tiled this many times")
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------
# clocks
# ---------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------
# CPU baseline: reference code on the host cores
# ---------------------------------------------------------------------------------------------
def _cpu_worker(args):
    kind, fields, wins, pvals = args
    from dindel_tgi_b200 import engine
    from oracle.pyoracle import Oracle, OraParams
    ora = Oracle(kind)
    if wins and wins[0] == 0:   # one worker says which shared object does the arithmetic (the driver cannot see forked workers' maps)
        maps = [l.split()[-1] for l in open("/proc/self/maps") if "libdindel_ref.so" in l or "libdindel_oracle.so" in l]
        sys.stderr.write("cpu arm worker %d (kind %s) maps: %s\n" % (os.getpid(), kind, sorted(set(maps))))
    batch = engine.Batch(*fields)
    p = OraParams(*pvals)
    n = 0
    t0 = time.perf_counter()
    for w in wins:
        k, _, _ = ora.window(batch, w, p, want_hpos=False)
        n += max(k, 0)
    return n, time.perf_counter() - t0


def cpu_baseline(batch, params, n_windows, cores=None):
    """Reference implementation on `cores` host processes over the first n_windows windows."""
    import multiprocessing as mp
    from oracle.pyoracle import REF_SO
    kind = "reference" if os.path.exists(REF_SO) else "port"
    cores = cores or (len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else os.cpu_count())
    n_windows = max(1, min(n_windows, batch.n_windows))
    cores = max(1, min(cores, n_windows))
    sub = batch.slice_windows(0, n_windows)
    fields = [getattr(sub, f) for f in ("hap_start", "win_hap", "win_read", "hap_off", "hap_seq", "read_off",
                                        "read_seq", "read_qual", "read_mapq", "read_start", "read_flags")]
    pvals = tuple(getattr(params, n) for n, _ in params._fields_)
    jobs = [(kind, fields, list(range(c, n_windows, cores)), pvals) for c in range(cores)]
    ctx = mp.get_context("fork")
    t0 = time.perf_counter()
    with ctx.Pool(cores) as pool:
        res = pool.map(_cpu_worker, jobs)
    wall = time.perf_counter() - t0
    pairs = sum(r[0] for r in res)
    busy = max(r[1] for r in res)
    return {"pairs": pairs, "cells": sub.n_cells(), "wall_s": wall, "busy_s": busy, "cores": cores, "kind": kind,
            "windows": n_windows}


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU path on all host threads, same config and metric."""
    if rank != 0:
        return
    from dindel_tgi_b200 import engine, synth
    params = engine.Params(5e-4, 1e-5, 100.0, 0.01, 0.95, 5, 2, -1, 0) if args.cfg != 4 else \
        engine.Params(5e-4, 1e-5, 100.0, 0.01, 0.95, 10, 2, -1, 0)
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    # bounded sample per step: ~60 windows per core (about 16k pairs = 4-5 s per core-step at ~4k
    # pairs/s/core), so that process start-up is amortised and K+W steps still end within minutes
    nwin = args.cpu_windows or max(8, int(60 * cores))
    batch = synth.make_batch(args.cfg, nwin)
    times, pairs, kind = [], 0, "port"
    for i in range(args.warmup + args.steps):
        r = cpu_baseline(batch, params, nwin, cores)
        kind = r["kind"]
        if i >= args.warmup:
            times.append(r["wall_s"])
            pairs = r["pairs"]
    ms = 1e3 * float(np.mean(times))
    val = pairs / (ms * 1e-3)
    sample = "%d windows (%d pairs, %d cells) of cfg %d per step, round-robin over %d processes" % (
        nwin, pairs, batch.n_cells(), args.cfg, cores)
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(args), "bounded_sample": sample},
            "gcups": batch.n_cells() / (ms * 1e-3) / 1e9,
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


def workload_name(args):
    return {2: "configs[1]: %dk synthetic diploid windows/GPU, 30x 100 bp reads, <=8 candidate haplotypes" % (args.windows // 1000),
            3: "configs[2]: pooled 1000x windows, 16 haplotypes", 4: "configs[3]: 250 bp reads, 32 haplotypes"}[args.cfg] \
        if args.windows >= 1000 else "cfg %d, %d windows/GPU (reduced)" % (args.cfg, args.windows)


# ---------------------------------------------------------------------------------------------
def pinned_copy(lib, arr):
    """numpy array backed by cudaHostAlloc memory holding a copy of arr."""
    import ctypes as C
    n = max(arr.nbytes, 1)
    p = lib.dindel_b200_alloc_pinned(n)
    if not p:
        raise RuntimeError("cudaHostAlloc failed for %d bytes" % n)
    buf = (C.c_uint8 * n).from_address(p)
    out = np.frombuffer(buf, dtype=arr.dtype, count=arr.size).reshape(arr.shape)
    out[...] = arr
    return out


_JSON_FD = None


def claim_stdout():
    """stdout carries the ONE JSON line and nothing else: everything any library writes to fd 1 from here on
    (NCCL's version banner under torchrun, for one) goes to stderr; emit() writes to the original stdout."""
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def _cpu_records_worker(args):
    kind, fields, wins, pvals = args
    from dindel_tgi_b200 import engine
    from oracle.pyoracle import Oracle, OraParams
    ora = Oracle(kind)
    batch = engine.Batch(*fields)
    p = OraParams(*pvals)
    return [(w, ora.window(batch, w, p, want_hpos=False)[1]) for w in wins]


PARITY_FIELDS = ("ll", "llOn", "llOff", "mLogBQ", "firstBase", "lastBase", "numIndels", "numMismatch", "nBQT", "nmmBQT",
                 "nMMLeft", "nMMRight", "offHap", "offHapHMQ", "status")


def parity_gate(batch, params, pairs, win_pair, n_windows):
    """n sampled windows of the batch through oracle/_ref (the reference's own class) on the host cores, compared bit
    for bit with the GPU's pair records.  Outside every timed region."""
    import multiprocessing as mp
    from oracle.pyoracle import REF_SO
    kind = "reference" if os.path.exists(REF_SO) else "port"
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    wins = sorted(set(np.linspace(0, batch.n_windows - 1, min(n_windows, batch.n_windows)).astype(int).tolist()))
    fields = [getattr(batch, f) for f in ("hap_start", "win_hap", "win_read", "hap_off", "hap_seq", "read_off",
                                          "read_seq", "read_qual", "read_mapq", "read_start", "read_flags")]
    pvals = tuple(getattr(params, n) for n, _ in params._fields_)
    k = max(1, min(cores, len(wins)))
    jobs = [(kind, fields, wins[c::k], pvals) for c in range(k)]
    with mp.get_context("fork").Pool(k) as pool:
        res = pool.map(_cpu_records_worker, jobs)
    bad = n = 0
    for part in res:
        for w, recs in part:
            got = pairs[int(win_pair[w]):int(win_pair[w + 1])]
            n += len(got)
            for f in PARITY_FIELDS:
                a, b = np.asarray(got[f]), np.asarray(recs[f])
                same = (a.view(np.int64) == b.astype(np.float64).view(np.int64)) if a.dtype.kind == "f" else (a.astype(np.int64) == b.astype(np.int64))
                bad += int((~same).sum())
    return {"windows": len(wins), "pairs": n, "field_mismatches": bad, "checker": "oracle/_ref (%s)" % kind}


def tile_batch(engine, b, k):
    """the batch repeated k times (CSR offsets shifted): a bigger job of the same distribution without regenerating it"""
    if k == 1:
        return b
    nh, nr = b.n_haps, b.n_reads
    hb, rb = int(b.hap_off[-1]), int(b.read_off[-1])
    cat = np.concatenate
    return engine.Batch(np.tile(b.hap_start, k),
                        cat([b.win_hap[:-1] + i * nh for i in range(k)] + [[k * nh]]),
                        cat([b.win_read[:-1] + i * nr for i in range(k)] + [[k * nr]]),
                        cat([b.hap_off[:-1] + i * hb for i in range(k)] + [[k * hb]]), np.tile(b.hap_seq, k),
                        cat([b.read_off[:-1] + i * rb for i in range(k)] + [[k * rb]]), np.tile(b.read_seq, k),
                        np.tile(b.read_qual, k), np.tile(b.read_mapq, k), np.tile(b.read_start, k), np.tile(b.read_flags, k))


def cover_inputs(engine, lib, batch):
    """K4 inputs (filterHaplotypes, DInDel.cpp:1932-2100): per haplotype one *REF-like entry and, for the non-reference
    haplotypes, one candidate-indel entry over the middle of the haplotype (+-padCover), keyed per haplotype; strands
    from the read flags.  Built once (host string logic in the reference), copied every step."""
    nh_w = np.diff(batch.win_hap).astype(np.int64)
    H = np.diff(batch.hap_off).astype(np.int64)
    hidx = np.arange(batch.n_haps) - np.repeat(batch.win_hap[:-1], nh_w)       # hap index inside its window
    alt = hidx > 0
    ne_h = 1 + alt.astype(np.int64)                                             # entries per haplotype
    e_hap = np.repeat(hidx, ne_h).astype(np.int32)
    first = np.concatenate([[0], np.cumsum(ne_h)[:-1]])
    e_type = np.full(int(ne_h.sum()), engine.VAR_OTHER, dtype=np.uint8)
    e_key = np.zeros(len(e_type), dtype=np.int32)
    e_left = np.zeros(len(e_type), dtype=np.int32)
    e_right = np.zeros(len(e_type), dtype=np.int32)
    ai = first[alt] + 1
    e_type[ai] = np.where(hidx[alt] % 2 == 1, engine.VAR_INS, engine.VAR_DEL)
    e_key[ai] = hidx[alt]
    e_left[ai] = (H[alt] // 2 - 6).astype(np.int32)
    e_right[ai] = (H[alt] // 2 + 6).astype(np.int32)
    ne_w = np.add.reduceat(ne_h, batch.win_hap[:-1].astype(np.int64)) if batch.n_windows else np.zeros(0, np.int64)
    win_entry = np.concatenate([[0], np.cumsum(ne_w)]).astype(np.int64)
    win_key = np.concatenate([[0], np.cumsum(nh_w)]).astype(np.int64)
    strand = ((batch.read_flags >> 1) & 1).astype(np.uint8)
    ci, co = engine._CoverIn(), engine._CoverOut()
    keep = [pinned_copy(lib, x) for x in (win_entry, win_key, e_hap, e_key, e_type, e_left, e_right, strand)]
    ci.n_entries, ci.n_keys, ci.maxMismatch, ci.doFilter = len(e_type), int(win_key[-1]), 2, 0
    (ci.win_entry, ci.win_key, ci.entry_hap, ci.entry_key, ci.entry_type, ci.entry_left, ci.entry_right,
     ci.read_strand) = [x.ctypes.data for x in keep]
    filt = pinned_copy(lib, np.zeros(batch.n_haps, dtype=np.uint8))
    knf = pinned_copy(lib, np.zeros(int(win_key[-1]), dtype=np.int32))
    knr = pinned_copy(lib, np.zeros(int(win_key[-1]), dtype=np.int32))
    co.hap_filtered, co.key_nf, co.key_nr = filt.ctypes.data, knf.ctypes.data, knr.ctypes.data
    return {"ci": ci, "co": co, "keep": keep + [filt], "h2d": sum(x.nbytes for x in keep),
            "d2h": filt.nbytes + knf.nbytes + knr.nbytes, "knf": knf, "knr": knr}


class HostJob:
    """One batch with pinned host buffers and the dindel_job describing the pass over it (what a C++ caller holds)."""

    def __init__(self, engine, lib, batch, cover=True, hpos_d2h=False, sums=True, n_groups=0, em=None, want_hpos=True):
        import ctypes as C
        names = ("hap_start", "win_hap", "win_read", "hap_off", "hap_seq", "read_off", "read_seq", "read_qual",
                 "read_mapq", "read_start", "read_flags")
        self.batch = batch
        self.pb = engine.Batch(*[pinned_copy(lib, getattr(batch, n)) for n in names])
        n_pairs = batch.n_pairs()
        self.pairs = pinned_copy(lib, np.zeros(n_pairs, dtype=engine.PAIR_DTYPE))
        self.win_pair = np.zeros(batch.n_windows + 1, dtype=np.int64)
        self.win_status = pinned_copy(lib, np.zeros(batch.n_windows, dtype=np.int32))
        self.on_hap = pinned_copy(lib, np.zeros(batch.n_reads, dtype=np.uint8))
        stride = (int(np.max(np.diff(batch.read_off))) + 7) & ~7
        self.hpos = pinned_copy(lib, np.zeros((n_pairs, stride), dtype=np.int16)) if hpos_d2h else None
        o = engine._BatchOut()
        self.cov = cover_inputs(engine, lib, batch) if (cover and want_hpos and not hpos_d2h) else None
        o.n_pairs, o.pairs, o.win_pair = n_pairs, self.pairs.ctypes.data, self.win_pair.ctypes.data
        o.win_status, o.read_on_hap = self.win_status.ctypes.data, self.on_hap.ctypes.data
        if self.hpos is not None:
            o.hpos, o.hpos_stride = self.hpos.ctypes.data, stride
        self.o, self.cb = o, self.pb._c()
        job = engine._Job()
        job.in_, job.out = C.addressof(self.cb), C.addressof(o)
        self.h2d = batch.input_bytes() + 2 * 8 * (batch.n_windows + 1)
        self.d2h = self.pairs.nbytes + self.win_status.nbytes + self.on_hap.nbytes + (self.hpos.nbytes if self.hpos is not None else 0)
        if self.cov:
            job.cover, job.cover_out = C.addressof(self.cov["ci"]), C.addressof(self.cov["co"])
            self.h2d += self.cov["h2d"]; self.d2h += self.cov["d2h"]
        self.sums = None
        if sums:
            ns = int(lib.dindel_b200_count_sums(C.byref(self.cb)))
            self.sums = pinned_copy(lib, np.zeros(max(ns * max(n_groups, 1), 1), dtype=np.float64))
            job.pair_sums, job.n_groups = self.sums.ctypes.data, int(n_groups)
            if n_groups:
                self.rg = pinned_copy(lib, np.zeros(batch.n_reads, dtype=np.int32))
                job.read_group = self.rg.ctypes.data
                self.h2d += self.rg.nbytes
            self.d2h += self.sums.nbytes
        self.em = None
        if em is not None:
            sets, comps, a0, tol = em
            sw = np.ascontiguousarray(sets, dtype=np.int32)
            off = np.concatenate([[0], np.cumsum([len(x) for x in comps])]).astype(np.int64)
            comp = np.ascontiguousarray(np.concatenate(comps), dtype=np.uint8)
            self.em = {"ll": np.zeros(len(sw)), "it": np.zeros(len(sw), dtype=np.int32), "pi": np.zeros(len(comp)),
                       "fr": np.zeros(len(comp)), "margin": np.zeros(len(sw)), "keep": (sw, off, comp)}
            a, eo = engine._EmIn(), engine._EmOut()
            a.n_sets, a.set_window, a.set_off, a.compatible, a.a0, a.tol = len(sw), sw.ctypes.data, off.ctypes.data, comp.ctypes.data, a0, tol
            eo.loglik, eo.iters, eo.pi, eo.freqs = (self.em[k].ctypes.data for k in ("ll", "it", "pi", "fr"))
            eo.margin = self.em["margin"].ctypes.data
            self.em["structs"] = (a, eo)
            job.em, job.em_out = C.addressof(a), C.addressof(eo)
            self.h2d += sw.nbytes + off.nbytes + comp.nbytes
            self.d2h += sum(self.em[k].nbytes for k in ("ll", "it", "pi", "fr", "margin"))
        self.job = job

    def path(self):
        st = ["K1"] + (["K4 (filterHaplotypes on the resident hpos rows)"] if self.cov else []) + \
             (["K2 (read sums of every haplotype pair)"] if self.sums is not None else []) + (["K3 (Bayes EM)"] if self.em else [])
        return "one dindel_b200_process call: H2D, " + " + ".join(st) + ", D2H of the 64-byte pair records and the stage outputs" + \
               ("; hpos rows copied to the host" if self.hpos is not None else "")


def measure(args, eng, engine, lib, cfg, n_windows, params, rank, steps, warmup, barrier, max_over_ranks, sum_over_ranks, world,
            fp64_peak, sampler=None, pooled=False):
    """resident-input K1 throughput + the end-to-end pass of one workload; returns the fields of a bench record."""
    import ctypes as C
    from dindel_tgi_b200 import synth
    t0 = time.time()
    batch = synth.make_batch(cfg, n_windows, seed=synth.SEED0 + cfg + 1000 * rank)
    gen_s = time.time() - t0
    n_pairs, n_cells = batch.n_pairs(), batch.n_cells()
    J = params.maxLengthDel + 1
    w_cell = 5 * J + 10
    tot_pairs, tot_cells = sum_over_ranks(float(n_pairs)), sum_over_ranks(float(n_cells))
    eng.upload(batch, want_hpos=not args.no_hpos)
    for _ in range(warmup):
        eng.run()
    if sampler:
        sampler.start()
    barrier()
    k1_ms, launches = [], 0
    wall0 = time.perf_counter()
    for _ in range(steps):
        t = eng.run()  # CUDA events on the library's stream bracket the step's kernels
        k1_ms.append(t.k1_ms)
        launches += t.k1_launches
    barrier()
    wall_ms = 1e3 * (time.perf_counter() - wall0)
    clocks = sampler.stop() if sampler else None
    dev_ms = max_over_ranks(float(np.sum(k1_ms)))
    ms_per_step = dev_ms / steps
    # ---- end to end: one dindel_b200_process call per step, pinned host buffers ----
    em = None
    if pooled:
        rng = np.random.default_rng(cfg)
        nh_w = np.diff(batch.win_hap)
        sets, comps = [], []
        for w in range(batch.n_windows):       # 16 active-variant sets per window (singlevariant: one per haplotype's variant set)
            for k in range(16):
                c = np.ones(nh_w[w], dtype=np.uint8) if k == 0 else (rng.random(nh_w[w]) < 0.7).astype(np.uint8)
                c[0] = 1
                sets.append(w); comps.append(c)
        em = (sets, comps, 0.001, 1e-4)
    hj = HostJob(engine, lib, batch, cover=not args.no_hpos, hpos_d2h=args.hpos_d2h, sums=True, n_groups=1 if pooled else 0,
                 em=em, want_hpos=not args.no_hpos)

    def e2e_step():
        eng._check(lib.dindel_b200_process(eng.h, C.byref(hj.job)), "process")
        return eng.timing()

    for _ in range(max(1, warmup - 1)):
        e2e_step()
    barrier()
    e0 = time.perf_counter()
    post_ms = dev_total = 0.0
    for _ in range(steps):
        t = e2e_step()
        post_ms += t.post_ms
        dev_total += t.total_ms
    barrier()
    e2e_wall = 1e3 * (time.perf_counter() - e0)
    e2e_ms = max_over_ranks(e2e_wall) / steps  # host wall clock around the call (includes host indexing)
    rec = {
        "value": tot_pairs / (ms_per_step * 1e-3), "ms_per_step": ms_per_step, "gcups": tot_cells / (ms_per_step * 1e-3) / 1e9,
        "windows_per_gpu": batch.n_windows, "pairs_per_gpu": n_pairs, "cells_per_gpu": n_cells, "gen_s": round(gen_s, 1),
        "wall_ms_per_step": wall_ms / steps, "gpu_launches": int(launches), "clocks": clocks,
        "e2e": {"value": tot_pairs / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(hj.h2d), "d2h_bytes_per_step": int(hj.d2h),
                "ms_per_step": e2e_ms, "device_ms_per_step": dev_total / steps, "gcups": tot_cells / (e2e_ms * 1e-3) / 1e9,
                "path": hj.path(), "stages_ms_per_step": post_ms / steps,
                "k4_covering_reads": int(hj.cov["knf"].sum() + hj.cov["knr"].sum()) if hj.cov else None,
                "pair_sums_checksum": float(hj.sums.sum()) if hj.sums is not None else None,
                "hpos_d2h": hj.hpos is not None, "ll_checksum": float(hj.pairs["ll"].sum()),
                "pcie_gbs": (hj.h2d + hj.d2h) / (e2e_ms * 1e-3) / 1e9},
        "roofline": {"bound": "fp64-issue", "achieved": (tot_cells / world * w_cell / (ms_per_step * 1e-3)) / 1e12,
                     "peak": fp64_peak / 1e12, "unit": "TOP/s",
                     "frac": (tot_cells / world * w_cell / (ms_per_step * 1e-3)) / fp64_peak,
                     "w_cell": w_cell, "peak_source": "measured here: dindel_b200_measure_fp64_peak (register-resident DADD chains)"},
    }
    if em is not None:
        rec["e2e"]["k3_sets_per_step"] = len(em[0])
        rec["e2e"]["k3_mean_passes"] = float(np.mean(hj.em["it"]))
        rec["e2e"]["k3_loglik_checksum"] = float(hj.em["ll"].sum())
    return rec, batch, hj


def strong_record(args, engine, lib, params, batch, n_dev):
    """rank 0 alone: ONE fixed job (the batch tiled) through dindel_b200_multi_process on 1 device, then on n_dev."""
    import ctypes as C
    import zlib
    job_batch = tile_batch(engine, batch, max(1, args.strong_tiles))
    hj = HostJob(engine, lib, job_batch, cover=True, sums=True)
    out = {"windows": job_batch.n_windows, "pairs": job_batch.n_pairs(), "cells": job_batch.n_cells(),
           "job": "the rank-0 batch tiled x%d, one dindel_b200_multi_process call (K1 + K4 + K2), pinned host buffers" % max(1, args.strong_tiles)}
    base = None
    for nd in sorted({1, n_dev}):
        m = engine.MultiEngine(params, tuple(range(nd)))
        try:
            def call():
                t0 = time.perf_counter()
                rc = lib.dindel_b200_multi_process(m.h, C.byref(hj.job))
                if rc != 0:
                    raise RuntimeError("multi_process: %s" % lib.dindel_b200_multi_last_error(m.h).decode())
                return 1e3 * (time.perf_counter() - t0)
            call()                                   # warm-up: buffers grow once
            ms = min(call(), call())
            shards = []
            for k in range(nd):
                a, b, t = C.c_int32(), C.c_int32(), engine.Timing()
                lib.dindel_b200_multi_get_shard(m.h, k, C.byref(a), C.byref(b), C.byref(t))
                shards.append({"device": k, "windows": b.value - a.value, "pairs": int(t.n_pairs), "h2d_busy_ms": t.h2d_ms,
                               "kernel_span_ms": t.k1_ms, "d2h_busy_ms": t.d2h_ms, "stages_ms": t.post_ms,
                               "device_total_ms": t.total_ms, "host_ms": t.host_ms, "ranges": t.n_ranges, "chunks": t.n_chunks})
            crc = zlib.crc32(hj.pairs.view(np.uint8)[: min(hj.pairs.nbytes, 1 << 30)]) ^ zlib.crc32(hj.sums.view(np.uint8))
            rec = {"devices": nd, "ms": ms, "pairs_per_s": out["pairs"] / (ms * 1e-3), "gcups": out["cells"] / (ms * 1e-3) / 1e9,
                   "per_device": shards, "crc32": crc}
            if base is None:
                base = rec
            out["n%d" % nd] = rec
        finally:
            m.close()
    if n_dev > 1:
        out["efficiency_vs_n1"] = base["ms"] / (n_dev * out["n%d" % n_dev]["ms"])
        out["buffers_identical_to_n1"] = bool(base["crc32"] == out["n%d" % n_dev]["crc32"])
    return out


def main():
    args = parse()
    claim_stdout()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    import torch
    from dindel_tgi_b200 import engine
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the K1 path has no CPU fallback")
    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    # ranks waiting for rank 0's single-process sections wait on the HOST (gloo), so that no NCCL kernel spins on the
    # GPUs rank 0 is driving through the multi-device call
    host_group = dist.new_group(backend="gloo") if dist is not None else None

    def host_barrier():
        if dist is not None:
            dist.barrier(group=host_group)

    def max_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    params = engine.cli_params()
    if args.cfg == 4:
        params.maxLengthDel = 10
    eng = engine.Engine(params, local)
    lib = eng.lib
    fp64_peak = eng.measure_fp64_peak()
    # shards differ per rank (seeded by rank): the job's units are the sum over ranks, not world x rank 0's
    rec, batch, hj = measure(args, eng, engine, lib, args.cfg, args.windows, params, rank, args.steps, args.warmup, barrier,
                             max_over_ranks, sum_over_ranks, world, fp64_peak, sampler=ClockSampler(local), pooled=args.cfg == 3)

    # ---- K5 over the hpos rows the e2e step left resident (reported beside the step, not part of it) ----
    infer = None
    if world == 1 and hj.cov and eng.timing().n_ranges == 1:
        nh_w = np.diff(batch.win_hap).astype(np.int64)
        hml = (np.arange(len(batch.hap_seq), dtype=np.int64) - np.repeat(batch.hap_off[:-1], np.diff(batch.hap_off))).astype(np.int32)
        k5a = (engine.CIGAR_PAIR, np.zeros(batch.n_windows, np.int32), hml, np.zeros(batch.n_windows, np.int32),
               (nh_w - 1).astype(np.int32), np.zeros(batch.n_haps, np.int32))
        # called the way the adapter calls it (include/dindel_b200.hpp): 6 operation slots per read in pinned
        # buffers, one exact-capacity second pass only if some read needs more
        cap = 6
        k5o = [pinned_copy(lib, np.zeros(batch.n_reads, np.int32)) for _ in range(3)] + \
              [pinned_copy(lib, np.zeros((batch.n_reads, cap), np.uint32))]
        eng.cigars(*k5a, cigar_cap=cap, n_reads=batch.n_reads, out=k5o)   # warm-up (buffers grow once)
        _, _, k5n, _ = eng.cigars(*k5a, cigar_cap=cap, n_reads=batch.n_reads, out=k5o)
        k5_ms, k5_kernel_ms, k5_max = eng.last_call_ms, eng.last_kernel_ms, int(k5n.max()) if len(k5n) else 0
        if k5_max > cap:                                        # second pass at the exact capacity
            _, _, k5n, _ = eng.cigars(*k5a, cigar_cap=k5_max, n_reads=batch.n_reads)
            k5_ms, k5_kernel_ms = k5_ms + eng.last_call_ms, k5_kernel_ms + eng.last_kernel_ms
        infer = {"k5_cigars_ms": k5_ms, "k5_kernel_ms": k5_kernel_ms, "k5_max_ops": k5_max,
                 "k5_reads_per_s": batch.n_reads / (k5_ms * 1e-3), "k5_frac_of_k1_step": k5_ms / rec["ms_per_step"],
                 "k5_ops_checksum": int(k5n.sum())}
        # ---- K6: the read statistics of three called sites per window (one dip.map site on the first / last haplotype,
        # two variant positions on haplotype pairs) for the whole resident batch in ONE call ----
        nwb = batch.n_windows
        H0 = np.diff(batch.hap_off)[batch.win_hap[:-1]].astype(np.int32)       # length of each window's first haplotype
        jw = np.repeat(np.arange(nwb, dtype=np.int32), 3)
        jm = np.tile(np.array([engine.STATS_MAP, engine.STATS_SITE, engine.STATS_SITE], np.int32), nwb)
        jc = np.arange(3 * nwb + 1, dtype=np.int64) * 2
        last = (nh_w - 1).astype(np.int32)
        ch = np.stack([np.zeros(nwb, np.int32), last, np.zeros(nwb, np.int32), np.minimum(1, last), np.minimum(1, last), last], 1).ravel()
        cs = np.repeat(np.stack([H0 // 2, H0 // 3, 2 * H0 // 3], 1).ravel(), 2).astype(np.int32)
        k6 = eng.read_stats_flat(jw, jm, np.full(3 * nwb, -1, np.int32), jc, ch, cs, cs + 2, np.ones(6 * nwb, np.uint8))   # warm-up
        k6 = eng.read_stats_flat(jw, jm, np.full(3 * nwb, -1, np.int32), jc, ch, cs, cs + 2, np.ones(6 * nwb, np.uint8))
        infer.update({"k6_read_stats_ms": eng.last_call_ms, "k6_kernel_ms": eng.last_kernel_ms, "k6_jobs": int(3 * nwb),
                      "k6_jobs_per_s": 3 * nwb / (eng.last_call_ms * 1e-3), "k6_frac_of_k1_step": eng.last_call_ms / rec["ms_per_step"],
                      "k6_covered_checksum": int(k6["n"].sum()), "k6_reads_checksum": int(k6["n_reads"].sum())})
    parity = None
    if rank == 0 and not args.no_cpu_baseline:
        parity = parity_gate(batch, params, hj.pairs, hj.win_pair, 32 if args.cfg == 2 else 8)

    # ---- extra_configs: configs[2] and configs[3] (rank 0's GPU, after the headline's timed regions) ----
    extra = None
    if rank == 0 and args.cfg == 2 and not args.no_extra and args.windows >= 1000:
        extra = {}
        nw3, nw4 = (int(x) for x in args.extra_windows.split(","))
        for name, cfg, nw, mld in (("configs[2]", 3, nw3, 5), ("configs[3]", 4, nw4, 10)):
            if nw <= 0:
                continue
            p2 = engine.cli_params()
            p2.maxLengthDel = mld
            e2 = engine.Engine(p2, local)
            try:
                noop = lambda: torch.cuda.synchronize()
                r2, b2, hj2 = measure(args, e2, engine, lib, cfg, nw, p2, 0, 2, 3, noop, lambda x: x, lambda x: x, 1, fp64_peak,
                                      pooled=cfg == 3)
                r2["workload"] = {3: "configs[2]: pooled 1000x windows, 16 haplotypes (K1 + K3 EM, 16 sets per window + per-pool K2)",
                                  4: "configs[3]: 250 bp reads, homopolymer-rich, 32 haplotypes, maxLengthDel 10"}[cfg]
                r2["steps"], r2["warmup"] = 2, 3
                if not args.no_cpu_baseline:
                    r2["parity_gate"] = parity_gate(b2, p2, hj2.pairs, hj2.win_pair, 16)
                r2.pop("clocks", None)
                extra[name] = r2
                del hj2, b2
            finally:
                e2.close()
    host_barrier()

    # ---- strong scaling through the in-process window scheduler: rank 0 drives every GPU, the others wait ----
    strong = None
    if not args.no_strong and args.cfg == 2:
        if rank == 0:
            try:
                strong = strong_record(args, engine, lib, params, batch, min(world, torch.cuda.device_count()))
            except Exception as e:      # reported, never fatal for the headline line
                strong = {"error": str(e)}
        host_barrier()

    if rank == 0:
        line = {
            "metric": METRIC, "value": rec["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": rec["ms_per_step"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(args), "windows_per_gpu": rec["windows_per_gpu"],
                       "pairs_per_gpu": rec["pairs_per_gpu"], "cells_per_gpu": rec["cells_per_gpu"], "maxLengthDel": params.maxLengthDel,
                       "l2": "inputs (%.0f MB/GPU) and outputs exceed the 126 MB L2; no flush needed" % (batch.input_bytes() / 1e6),
                       "gen_s": rec["gen_s"], "wall_ms_per_step": rec["wall_ms_per_step"],
                       "value_excludes": "K0 (homopolymer codes) and the work-unit lists are built once by dindel_b200_upload, "
                                         "outside the timed dindel_b200_run steps (0.1 % of a step); e2e includes them every step"},
            "gcups": rec["gcups"],
            "e2e": rec["e2e"],
            "gpu_launches": rec["gpu_launches"],
            "clocks": rec["clocks"],
            "roofline": dict(rec["roofline"], traffic=TRAFFIC.get(args.cfg) and int(rec["cells_per_gpu"] * TRAFFIC[args.cfg]) or None,
                             traffic_source=TRAFFIC_SOURCE if TRAFFIC.get(args.cfg) else None),
        }
        if infer:
            line["inference_kernels"] = infer
        if parity:
            line["parity_gate"] = parity
        if extra:
            line["extra_configs"] = extra
        if strong:
            line["strong"] = strong
        if world == 1 and not args.no_cpu_baseline:
            cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
            nwin = args.cpu_windows or max(8, int(100 * cores))
            r = cpu_baseline(batch, params, nwin, cores)
            line["cpu_baseline"] = {"value": r["pairs"] / r["wall_s"], "unit": UNIT, "cores": r["cores"],
                                    "kind": r["kind"],
                                    "sample": "first %d windows of the workload (%d pairs, %.1f s wall, one process per core)" % (
                                        r["windows"], r["pairs"], r["wall_s"]),
                                    "gcups": r["cells"] / r["wall_s"] / 1e9}
        emit(line)
    eng.close()
    if dist is not None:
        dist.destroy_process_group()


# DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum) per cell of the dominant K1 launch, from one `ncu --set full`
# capture of `python bench.py --windows 3000 --steps 2 --warmup 3 --no-cpu-baseline --no-extra --no-strong` (hpos rows
# included): 118.1 MB read + 185.5 MB written for the launch's 8.9 G cells (round 1's capture: 302 MB).
TRAFFIC = {2: 303.6e6 / 8.9e9}
TRAFFIC_SOURCE = "profiles/r2_k1_j6_v5_ncu_raw.csv (k1_tiled_kernel<4,6>, ncu --set full, round 2), scaled to this step's cells"


if __name__ == "__main__":
    main()
