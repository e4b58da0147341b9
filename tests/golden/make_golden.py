"""Regenerates tests/golden/k1_golden.npz from the REFERENCE'S OWN code (oracle/_ref, built from
/root/reference by `make -C oracle ref`).  Run in the build container only:
    python tests/golden/make_golden.py
Stores the inputs (engine.Batch arrays), the parameter sets, and per pair the reference's
MLAlignment image + hpos, so the GPU box (which has no /root/reference) can check against them.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from dindel_tgi_b200 import synth  # noqa: E402
from oracle.pyoracle import Oracle, OraParams  # noqa: E402

PARAM_SETS = {
    "cli": OraParams(5e-4, 1e-5, 100.0, 0.01, 0.95, 5, 2, -1, 0),       # DInDel.cpp:4122-4157
    "default": OraParams(1e-4, 1e-4, 100.0, 0.01, 0.95, 10, 5, -1, 0),  # ObservationModel.hpp:39-62
}
CASES = [("cfg2_cli", 2, 6, "cli"), ("cfg2_default", 2, 3, "default"), ("cfg4_default", 4, 1, "default")]
FIELDS = ("hap_start", "win_hap", "win_read", "hap_off", "hap_seq", "read_off", "read_seq", "read_qual",
          "read_mapq", "read_start", "read_flags")


def main():
    ref = Oracle("reference")
    out = {}
    for name, cfg, nw, pset in CASES:
        b = synth.make_batch(cfg, nw, seed=777 + cfg)
        if cfg == 4:  # keep the fixture small: 6 haplotypes, 12 reads
            b = synth.Batch.from_windows([(hs, haps[:6], reads[:12]) for hs, haps, reads in
                                          [(b.window(0)[2], b.window(0)[0], b.window(0)[1])]])
        recs, hposes = [], []
        for w in range(b.n_windows):
            n, o, hp = ref.window(b, w, PARAM_SETS[pset])
            assert n == len(o)
            recs.append(o)
            hposes.append(hp.astype(np.int16))
        for f in FIELDS:
            out[f"{name}/{f}"] = getattr(b, f)
        out[f"{name}/records"] = np.concatenate(recs)
        stride = max(h.shape[1] for h in hposes)
        hp_all = np.full((sum(len(h) for h in hposes), stride), -32768, dtype=np.int16)
        i = 0
        for h in hposes:
            hp_all[i:i + len(h), :h.shape[1]] = h
            i += len(h)
        out[f"{name}/hpos"] = hp_all
        out[f"{name}/pset"] = np.array(pset)
    np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "k1_golden.npz"), **out)
    print({k: v.shape for k, v in out.items() if k.endswith("records")})


if __name__ == "__main__":
    main()
