#!/usr/bin/env python
"""bench.py — read-haplotype likelihoods/s and HMM GCUPS of the K1 path on B200 (BASELINE.json).

A "step" is one pass of the hot path (DetInDel::computeLikelihoods for every window) over one
batch of synthetic windows of BASELINE configs[1]: 100k diploid windows, 30x 100-bp reads, up to 8
candidate haplotypes (dindel_tgi_b200/synth.py, cfg 2).  Per GPU the batch is fixed, so N GPUs
process N shards ("weak" scaling, no data-path collective: windows are independent).

  value : pairs/s with the batch already resident in HBM (dindel_b200_upload once, then
          dindel_b200_run per step), device time from CUDA events on the library's stream.
  e2e   : the same through the C ABI with pinned HOST buffers, every step: dindel_b200_submit (H2D of every
          input array, kernels, D2H of the 64-byte pair records) followed by dindel_b200_coverage (K4 =
          filterHaplotypes on the hpos rows where K1 left them in HBM; H2D of the variant entries, D2H of
          filtered[] and the coverage counts).  --hpos-d2h copies the hpos rows to the host instead.
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


def main():
    args = parse()
    claim_stdout()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    import ctypes as C
    import torch
    from dindel_tgi_b200 import engine, synth
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the K1 path has no CPU fallback")
    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    params = engine.cli_params()
    if args.cfg == 4:
        params.maxLengthDel = 10
    eng = engine.Engine(params, local)
    lib = eng.lib
    t0 = time.time()
    batch = synth.make_batch(args.cfg, args.windows, seed=synth.SEED0 + args.cfg + 1000 * rank)
    gen_s = time.time() - t0
    n_pairs, n_cells = batch.n_pairs(), batch.n_cells()
    J = params.maxLengthDel + 1
    w_cell = 5 * J + 10

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

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

    # shards differ per rank (seeded by rank): the job's units are the sum over ranks, not world x rank 0's
    tot_pairs, tot_cells = sum_over_ranks(float(n_pairs)), sum_over_ranks(float(n_cells))
    fp64_peak = eng.measure_fp64_peak()

    # ---- resident-input throughput (value) ----
    eng.upload(batch, want_hpos=not args.no_hpos)
    for _ in range(args.warmup):
        eng.run()
    sampler = ClockSampler(local)
    sampler.start()
    barrier()
    k1_ms, launches = [], 0
    wall0 = time.perf_counter()
    for _ in range(args.steps):
        t = eng.run()  # CUDA events on the library's stream bracket the step's kernels
        k1_ms.append(t.k1_ms)
        launches += t.k1_launches
    barrier()
    wall_ms = 1e3 * (time.perf_counter() - wall0)
    clocks = sampler.stop()
    dev_ms = max_over_ranks(float(np.sum(k1_ms)))
    ms_per_step = dev_ms / args.steps
    value = tot_pairs / (ms_per_step * 1e-3)
    gcups = tot_cells / (ms_per_step * 1e-3) / 1e9

    # ---- end to end through dindel_b200_submit with pinned host buffers ----
    names = ("hap_start", "win_hap", "win_read", "hap_off", "hap_seq", "read_off", "read_seq", "read_qual",
             "read_mapq", "read_start", "read_flags")
    pb = engine.Batch(*[pinned_copy(lib, getattr(batch, n)) for n in names])
    pairs = pinned_copy(lib, np.zeros(n_pairs, dtype=engine.PAIR_DTYPE))
    win_pair = np.zeros(batch.n_windows + 1, dtype=np.int64)
    win_status = pinned_copy(lib, np.zeros(batch.n_windows, dtype=np.int32))
    on_hap = pinned_copy(lib, np.zeros(batch.n_reads, dtype=np.uint8))
    stride = (int(np.max(np.diff(batch.read_off))) + 7) & ~7
    use_k4 = not args.no_hpos and not args.hpos_d2h
    hpos = pinned_copy(lib, np.zeros((n_pairs, stride), dtype=np.int16)) if args.hpos_d2h else None
    o = engine._BatchOut()
    if use_k4:
        o.flags = engine.OUT_KEEP_HPOS
    # K4 inputs (filterHaplotypes, DInDel.cpp:1932-2100): per haplotype one *REF-like entry and, for the
    # non-reference haplotypes, one candidate-indel entry over the middle of the haplotype (+-padCover), keyed per
    # haplotype; strands from the read flags.  Built once (host string logic in the reference), copied every step.
    cov = None
    if use_k4:
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
        cov = (ci, co, sum(x.nbytes for x in keep), filt.nbytes + knf.nbytes + knr.nbytes, knf, knr)
    o.n_pairs, o.pairs, o.win_pair = n_pairs, pairs.ctypes.data, win_pair.ctypes.data
    o.win_status, o.read_on_hap = win_status.ctypes.data, on_hap.ctypes.data
    if hpos is not None:
        o.hpos, o.hpos_stride = hpos.ctypes.data, stride
    cb = pb._c()
    h2d = batch.input_bytes() + 2 * 8 * (batch.n_windows + 1) + (cov[2] if cov else 0)
    d2h = pairs.nbytes + win_status.nbytes + on_hap.nbytes + (hpos.nbytes if hpos is not None else 0) + (cov[3] if cov else 0)

    def e2e_step():
        eng._check(lib.dindel_b200_submit(eng.h, C.byref(cb), C.byref(o)), "submit")
        dev = eng.timing().total_ms
        t0 = time.perf_counter()
        if cov:
            eng._check(lib.dindel_b200_coverage(eng.h, C.byref(cov[0]), C.byref(cov[1])), "coverage")
        return dev, 1e3 * (time.perf_counter() - t0)

    for _ in range(max(1, args.warmup - 1)):
        e2e_step()
    barrier()
    e0 = time.perf_counter()
    e2e_dev, k4_ms = 0.0, 0.0
    for _ in range(args.steps):
        d, k = e2e_step()
        e2e_dev += d
        k4_ms += k
    barrier()
    e2e_wall = 1e3 * (time.perf_counter() - e0)
    e2e_ms = max_over_ranks(e2e_wall) / args.steps  # host wall clock around submit (includes host indexing)
    e2e_val = tot_pairs / (e2e_ms * 1e-3)
    checksum = float(pairs["ll"].sum())

    # ---- K2 / K3 over the resident likelihood matrix (reported beside the step, not part of it) ----
    infer = None
    if world == 1:
        nh_w = np.diff(batch.win_hap).astype(np.int64)
        jw = np.arange(batch.n_windows, dtype=np.int32)
        starts = [np.zeros(int(n * (n + 1) // 2)) for n in nh_w]
        eng.pair_sums(jw, starts)                                   # warm-up at full size (buffers grow once)
        sums = eng.pair_sums(jw, starts)                            # diploidGLF core, one job per window
        k2_ms = eng.last_call_ms                                    # wall time of the C call (H2D + kernel + D2H)
        infer = {"k2_pair_sums_ms": k2_ms, "k2_kernel_ms": eng.last_kernel_ms, "k2_windows_per_s": batch.n_windows / (k2_ms * 1e-3),
                 "k2_frac_of_k1_step": k2_ms / ms_per_step, "k2_checksum": float(sum(float(x.sum()) for x in sums[:1000]))}
        if cov:   # K5 over the hpos rows the e2e step left resident: realigned-read CIGARs, diploid choice among (0, nh-1)
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
            eng.last_call_ms, eng.last_kernel_ms = k5_ms, k5_kernel_ms
            infer.update({"k5_cigars_ms": eng.last_call_ms, "k5_kernel_ms": eng.last_kernel_ms, "k5_max_ops": k5_max,
                          "k5_reads_per_s": batch.n_reads / (eng.last_call_ms * 1e-3), "k5_frac_of_k1_step": eng.last_call_ms / ms_per_step,
                          "k5_ops_checksum": int(k5n.sum())})
        if args.cfg == 3:
            comp = [np.ones(int(n), dtype=np.uint8) for n in nh_w]
            eng.bayes_em(jw, comp)                                  # warm-up at full size
            ll, it, _, _ = eng.bayes_em(jw, comp)                   # Bayes EM core, one active-variant set per window
            k3_ms = eng.last_call_ms
            infer.update({"k3_bayes_em_ms": k3_ms, "k3_kernel_ms": eng.last_kernel_ms, "k3_sets_per_s": batch.n_windows / (k3_ms * 1e-3),
                          "k3_frac_of_k1_step": k3_ms / ms_per_step, "k3_mean_iters": float(np.mean(it))})

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(args), "windows_per_gpu": batch.n_windows,
                       "pairs_per_gpu": n_pairs, "cells_per_gpu": n_cells, "maxLengthDel": params.maxLengthDel,
                       "l2": "inputs (%.0f MB/GPU) and outputs exceed the 126 MB L2; no flush needed" % (batch.input_bytes() / 1e6),
                       "gen_s": round(gen_s, 1), "wall_ms_per_step": wall_ms / args.steps},
            "gcups": gcups,
            "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "ms_per_step": e2e_ms, "device_ms_per_step": e2e_dev / args.steps, "gcups": tot_cells / (e2e_ms * 1e-3) / 1e9,
                    "path": ("K1 + K4: submit (H2D, kernels, D2H of the 64-byte pair records) then dindel_b200_coverage "
                             "(filterHaplotypes on the hpos rows where K1 left them in HBM)") if cov else
                            ("K1, hpos rows copied to the host" if hpos is not None else "K1 only, no hpos rows"),
                    "k4_ms_per_step": (k4_ms / args.steps) if cov else None,
                    "k4_covering_reads": int(cov[4].sum() + cov[5].sum()) if cov else None,
                    "hpos_d2h": hpos is not None, "ll_checksum": checksum},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": "fp64-issue", "achieved": (tot_cells / world * w_cell / (ms_per_step * 1e-3)) / 1e12,
                         "peak": fp64_peak / 1e12, "unit": "TOP/s",
                         "frac": (tot_cells / world * w_cell / (ms_per_step * 1e-3)) / fp64_peak,
                         # dram read+write of k1_tiled_kernel<4,6> from one `ncu --set full` capture (profiles/
                         # r1_k1_tiled_v4_k4.md: 302 MB for 8.9 G cells, hpos rows included), scaled to this step's cells
                         "traffic": int(n_cells * (301.8e6 / 8.9e9)) if args.cfg == 2 and not args.no_hpos else None,
                         "w_cell": w_cell, "peak_source": "measured here: dindel_b200_measure_fp64_peak (register-resident DADD chains)",
                         "hbm_gbs": (h2d + d2h) / (ms_per_step * 1e-3) / 1e9},
        }
        if infer:
            line["inference_kernels"] = infer
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


if __name__ == "__main__":
    main()
