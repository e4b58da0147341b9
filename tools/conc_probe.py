#!/usr/bin/env python
"""tools/conc_probe.py WORK N ARM -- N concurrent stage-3 processes of one arm on window file 1; wall/user/sys of each
(os.wait4 rusage).  Measurement aid for DESIGN.md section 9 (where the pipeline's wall time goes); not part of the product."""
import os
import sys
import time

work, n, arm = sys.argv[1], int(sys.argv[2]), sys.argv[3]
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
binp = os.path.join(root, "oracle", "_ref", "dindel_b200" if arm == "gpu" else "dindel_ref")
os.chdir(work)
pids = {}
t0 = time.time()
for i in range(n):
    pid = os.fork()
    if pid == 0:
        fd = os.open("probe.%s.%d.%d.log" % (arm, n, i), os.O_WRONLY | os.O_CREAT | os.O_TRUNC)
        os.dup2(fd, 1)
        os.dup2(fd, 2)
        os.execv(binp, [binp, "--analysis", "indels", "--doDiploid", "--bamFile", "sample.bam", "--ref", "ref.fa", "--varFile",
                        "win.1.txt", "--libFile", "s1.libraries.txt", "--outputFile", "probe.%s.%d.%d" % (arm, n, i)])
    pids[pid] = i
res = []
while pids:
    pid, st, ru = os.wait4(-1, 0)
    res.append((pids.pop(pid), time.time() - t0, ru.ru_utime, ru.ru_stime, ru.ru_nvcsw, ru.ru_nivcsw, st))
res.sort()
w = [r[1] for r in res]
print("%s n=%d wall min/max %.2f/%.2f  user mean %.2f  sys mean %.2f  vol-cs mean %d  invol-cs mean %d  rc %s" % (
    arm, n, min(w), max(w), sum(r[2] for r in res) / n, sum(r[3] for r in res) / n, sum(r[4] for r in res) / n,
    sum(r[5] for r in res) / n, sorted(set(r[6] for r in res))))
