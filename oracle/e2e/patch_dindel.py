#!/usr/bin/env python
"""oracle/e2e/patch_dindel.py — TEST INFRASTRUCTURE ONLY.

Applies INTEGRATION.md's two-line patch to the reference's driver at BUILD time:
    python oracle/e2e/patch_dindel.py /root/reference/DInDel.cpp oracle/_ref/DInDel_b200.cpp
The output (git-ignored, under oracle/_ref/) is DInDel.cpp with (1) `#include "dindel_b200.hpp"` after
its own includes and (2) the body of DetInDel::computeLikelihoods (DInDel.cpp:1707-1739) preceded by a
forward to dindel_b200::Engine::computeLikelihoods + return.  Nothing else changes, so the binary built
from it (oracle/_ref/dindel_b200) is "the reference with the GPU engine dropped in": candidate
haplotypes, filterHaplotypes, diploidGLF, glf writer etc. are the reference's own code.
No reference source is stored in the repository."""
import re
import sys

src, dst = sys.argv[1], sys.argv[2]
text = open(src, encoding="latin-1").read()

inc = '#include "Faster.hpp"\n'
assert text.count(inc) == 1
text = text.replace(inc, inc + '#include "dindel_b200.hpp"\n')

sig = re.compile(r"\nvoid DetInDel::computeLikelihoods\(const vector<Haplotype> &haps, const vector<Read> & reads, "
                 r"vector<vector<MLAlignment> > & liks, uint32_t leftPos, uint32_t rightPos, vector<int> & onHap\)\n\{\n")
hits = list(sig.finditer(text))
assert len(hits) == 2, len(hits)          # the first one sits inside a /* ... */ block (DInDel.cpp:1652)
m = hits[-1]
fwd = ("\t{ // dindel_b200 drop-in (INTEGRATION.md section 0)\n"
       "\t\tstatic dindel_b200::Engine *gpu = 0;\n"
       "\t\tif (!gpu) gpu = new dindel_b200::Engine(params.obsParams, getenv(\"DINDEL_B200_DEVICE\") ? atoi(getenv(\"DINDEL_B200_DEVICE\")) : 0);\n"
       "\t\tgpu->computeLikelihoods(haps, reads, liks, leftPos, rightPos, onHap);\n"
       "\t\treturn;\n"
       "\t}\n")
text = text[:m.end()] + fwd + text[m.end():]
open(dst, "w", encoding="latin-1").write(text)
