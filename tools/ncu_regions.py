#!/usr/bin/env python
"""tools/ncu_regions.py — attributes the executed instructions and warp-stall samples of one profiled kernel to source
regions.  Inputs: the SASS page of an ncu report (`ncu -i X.ncu-rep --page source --csv --kernel-name ... > sass.csv`) and
`nvdisasm -g -fun <symbol index> <cubin>` of the same kernel (line info from -lineinfo); the two list the kernel's
instructions in the same order.  Regions are line ranges of csrc/k1_tiled.cu + whole included files.

    python tools/ncu_regions.py sass.csv dis.txt [table index]
"""
import collections
import csv
import re
import sys


def main():
    sass, dis, which = sys.argv[1], sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 0
    cur, inst = None, []
    pat = re.compile(r'^\s+/\*([0-9a-f]{4,6})\*/\s+(.*?);')
    for l in open(dis):
        m = re.match(r'\s*//## File "(.*?)", line (\d+)', l)
        if m:
            cur = (m.group(1).split('/')[-1], int(m.group(2)))
            continue
        m = pat.match(l)
        if m and '.byte' not in l:
            inst.append((m.group(2).strip(), cur))
    rows = list(csv.reader(open(sass)))
    starts = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
    hdr = rows[starts[which]]
    end = starts[which + 1] - 1 if which + 1 < len(starts) else len(rows)
    body = [r for r in rows[starts[which] + 1:end] if len(r) == len(hdr)]
    assert len(body) == len(inst), (len(body), len(inst))
    ie, si = hdr.index("Instructions Executed"), hdr.index("# Samples")
    stall = {k: hdr.index(k) for k in ("stall_wait", "stall_no_inst", "stall_short_sb", "stall_long_sb", "stall_math", "stall_not_selected", "stall_selected")}
    src = open("dindel_tgi_b200/csrc/k1_tiled.cu").read().splitlines()
    mark = {}
    for name, text in (("left", "=================== left half"), ("ao", "alpha[bMid] + obs[bMid] (first add"), ("right", "=================== right half"),
                       ("meet", "=================== meeting row"), ("unit", "Units of this class, sorted by RO slot")):
        hits = [i + 1 for i, l in enumerate(src) if text in l]
        mark[name] = hits[0] if hits else None

    def region(loc):
        if not loc:
            return "none"
        f, l = loc
        if f != "k1_tiled.cu":
            return f
        if l < mark["unit"]:
            return "prologue (tables, smem carve)"
        if l < mark["left"]:
            return "unit / pair setup"
        if l < mark["ao"]:
            return "left rows"
        if l < mark["right"]:
            return "alpha+obs store"
        # the meeting row starts after the right-half lambda call
        m = mark["meet"] or 10 ** 9
        return "right rows" if l < m else "meeting row + scans + tail call"
    tot = collections.Counter()
    per = collections.defaultdict(collections.Counter)
    for (txt, loc), r in zip(inst, body):
        k = region(loc)
        per[k]["inst"] += int(r[ie] or 0)
        per[k]["samples"] += int(r[si] or 0)
        per[k]["sass"] += 1
        for s, i in stall.items():
            per[k][s] += int(r[i] or 0)
    ti = sum(v["inst"] for v in per.values())
    ts = sum(v["samples"] for v in per.values())
    print("kernel:", rows[starts[which] - 1][1] if starts[which] else "")
    print("%-34s %7s %7s %8s   %s" % ("region", "inst %", "smpl %", "SASS", "stall samples: wait / no_inst / short_sb / long_sb / math / not_sel / selected"))
    for k, v in sorted(per.items(), key=lambda kv: -kv[1]["samples"]):
        print("%-34s %7.2f %7.2f %8d   %s" % (k, 100.0 * v["inst"] / ti, 100.0 * v["samples"] / ts, v["sass"],
                                              " / ".join("%d" % v[s] for s in stall)))
    print("marks:", mark)


if __name__ == "__main__":
    main()
