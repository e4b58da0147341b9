#!/usr/bin/env python
"""oracle/e2e/patch_dindel.py — TEST INFRASTRUCTURE ONLY.

Applies INTEGRATION.md's four-line patch to the reference's driver at BUILD time:
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
check = len(sys.argv) > 3 and sys.argv[3] == "--check"   # differential build, see CHECK below
dump = len(sys.argv) > 3 and sys.argv[3] == "--dump"     # stock CPU driver + the DUMP hook below
text = open(src, encoding="latin-1").read()

inc = '#include "Faster.hpp"\n'
assert text.count(inc) == 1
if not (len(sys.argv) > 3 and sys.argv[3] == "--dump"):
    text = text.replace(inc, inc + '#include "dindel_b200_driver.hpp"\nstatic dindel_b200::Engine *dindel_b200_gpu = 0;\n'
                        'static dindel_b200::Inference dindel_b200_inf; // K2 / K3 with the guard band (INTEGRATION.md section 0b)\n'
                        '#define DINDEL_B200_INF (dindel_b200_replay ? dindel_b200_replay->inf : dindel_b200_inf)\n'
                        '// K6: the read statistics of called sites from the GPU when the policy says so (dindel_b200::deviceStats); false = the loop below it runs\n'
                        '#define DINDEL_B200_STATS(JOBS, NREADS, OUT) (dindel_b200_replay ? dindel_b200::deviceStats(*dindel_b200_replay, JOBS, NREADS, OUT) : (dindel_b200_gpu ? dindel_b200::deviceStats(*dindel_b200_gpu, JOBS, NREADS, OUT) : false))\n'
                        'static std::string dindel_b200_bam_arg; static int dindel_b200_multiple_files = 0; // main() -> batched driver\n'
                        'static struct dindel_b200_report_t { ~dindel_b200_report_t() { if (dindel_b200_inf.nWindows) '
                        'fprintf(stderr, "dindel_b200 inference: windows %ld rechecked_on_host %ld em_sets %ld em_set_mismatch %ld em_unsupported_windows %ld\\n", '
                        'dindel_b200_inf.nWindows, dindel_b200_inf.nRecheck, dindel_b200_inf.nEmSets, dindel_b200_inf.nEmMismatch, dindel_b200_inf.nEmUnsupported); '
                        'if (dindel_b200_gpu && dindel_b200_gpu->nStatsCalls) fprintf(stderr, "dindel_b200 read stats: device_calls %ld jobs %ld\\n", dindel_b200_gpu->nStatsCalls, dindel_b200_gpu->nStatsJobs); } } dindel_b200_report;\n'
                        '// the one engine of the process; K2 / K3 (the read sums of diploidGLF / the EM of estimateHaplotypeFrequenciesBayesEM) on the GPU too\n'
                        'static dindel_b200::Engine *dindel_b200_engine(const DetInDel::Parameters & params, size_t nBams)\n{\n'
                        '\tif (!dindel_b200_gpu) {\n'
                        '\t\tdindel_b200_gpu = new dindel_b200::Engine(params.obsParams, getenv("DINDEL_B200_DEVICE") ? atoi(getenv("DINDEL_B200_DEVICE")) : 0);\n'
                        '\t\tconst bool infer = !getenv("DINDEL_B200_HOST_INFERENCE") && (params.doDiploid != params.estimateHapFreqs);\n'
                        '\t\tdindel_b200_gpu->setInference(infer, params.estimateHapFreqs ? int(nBams) : 0);\n'
                        '\t\tdindel_b200_inf.enable(infer);\n'
                        '\t\t// $DINDEL_B200_DEVICES=0,1,..: the batched driver shards every batch over these GPUs (dindel_b200_multi_process);\n'
                        '\t\t// the pooled path keeps one device (its per-window K3 calls need the batch resident on the engine\'s own)\n'
                        '\t\tif (getenv("DINDEL_B200_DEVICES") && !params.estimateHapFreqs) {\n'
                        '\t\t\tstd::vector<int> devs; std::string l(getenv("DINDEL_B200_DEVICES")); size_t a = 0;\n'
                        '\t\t\twhile (a <= l.size()) { size_t b = l.find(\',\', a); if (b == std::string::npos) b = l.size(); if (b > a) devs.push_back(atoi(l.substr(a, b - a).c_str())); a = b + 1; }\n'
                        '\t\t\tdindel_b200_gpu->enableMulti(devs);\n'
                        '\t\t}\n'
                        '\t}\n\treturn dindel_b200_gpu;\n}\n'
                        '// the batched driver creates the engine on a helper thread (context creation under the first prepare phase)\n'
                        'static dindel_b200::Engine *dindel_b200_engine_factory(void *det) { return dindel_b200_engine(((DetInDel *)det)->params, ((DetInDel *)det)->getMyBams().size()); }\n')
    # the class gets one friend (a patched copy of DInDel.hpp next to the patched DInDel.cpp; the class layout is unchanged)
    hdr = open(src[:-len("DInDel.cpp")] + "DInDel.hpp", encoding="latin-1").read()
    cls = "class DetInDel\n{\npublic:\n"
    assert hdr.count(cls) == 1
    hdr = hdr.replace(cls, cls + "\tfriend struct dindel_b200_driver; // the batched window loop (include/dindel_b200_driver.hpp)\n")
    open(dst[:-len(".cpp")] + ".hpp", "w", encoding="latin-1").write(hdr)
    own = '#include "DInDel.hpp"\n'
    assert text.count(own) == 1
    text = text.replace(own, '#include "%s"\n' % (dst[:-len(".cpp")].split("/")[-1] + ".hpp"))

sig = re.compile(r"\nvoid DetInDel::computeLikelihoods\(const vector<Haplotype> &haps, const vector<Read> & reads, "
                 r"vector<vector<MLAlignment> > & liks, uint32_t leftPos, uint32_t rightPos, vector<int> & onHap\)\n\{\n")
hits = list(sig.finditer(text))
assert len(hits) == 2, len(hits)          # the first one sits inside a /* ... */ block (DInDel.cpp:1652)
m = hits[-1]
fwd = ("\t{ // dindel_b200 drop-in (INTEGRATION.md section 0)\n"
       "\t\tif (dindel_b200_replay) { // batched driver: the window's results are in the flushed batch\n"
       "\t\t\tif (dindel_b200_replay->w < 0) { liks = vector<vector<MLAlignment> >(haps.size(), vector<MLAlignment>(reads.size())); onHap = vector<int>(reads.size(), 0); return; }\n"
       "\t\t\tdindel_b200_replay->batch->result(dindel_b200_replay->w, haps, reads, liks, onHap);\n"
       "\t\t\treturn;\n"
       "\t\t}\n"
       "\t\tdindel_b200::Engine *gpu = dindel_b200_engine(params, myBams.size());\n"
       "\t\tgpu->computeLikelihoods(haps, reads, liks, leftPos, rightPos, onHap);\n"
       "\t\treturn;\n"
       "\t}\n")
# --check: oracle/_ref/dindel_b200_check.  After the GPU call the reference's own CPU loop runs on the same
# window and every MLAlignment field is compared pair by pair (a differential harness on the inputs the real
# pipeline produces: candidate haplotypes from HaplotypeDistribution, reads as they come out of the BAM);
# one "B200CHECK" line per window on stderr, the GPU result is what the driver continues with.
CHECK = r"""
		{
			static long nPairs = 0, nBad = 0;
			long bad0 = nBad;
			for (size_t hidx=0;hidx<haps.size();hidx++) for (size_t r=0;r<reads.size();r++) {
				ObservationModelFBMaxErr oms(haps[hidx], reads[r], leftPos, params.obsParams);
				MLAlignment c = oms.calcLikelihood();
				const MLAlignment & g = liks[hidx][r];
				bool same = c.ll==g.ll && c.llOn==g.llOn && c.llOff==g.llOff && c.offHap==g.offHap && c.offHapHMQ==g.offHapHMQ
					&& c.firstBase==g.firstBase && c.lastBase==g.lastBase && c.numIndels==g.numIndels && c.numMismatch==g.numMismatch
					&& c.nBQT==g.nBQT && c.nmmBQT==g.nmmBQT && c.mLogBQ==g.mLogBQ && c.nMMLeft==g.nMMLeft && c.nMMRight==g.nMMRight
					&& c.hpos==g.hpos && c.align==g.align && c.hapIndelCovered==g.hapIndelCovered && c.hapSNPCovered==g.hapSNPCovered
					&& c.indels.size()==g.indels.size() && c.snps.size()==g.snps.size();
				if (same) {
					map<int, AlignedVariant>::const_iterator a=c.indels.begin(), b=g.indels.begin();
					for (;a!=c.indels.end();a++,b++) if (a->first!=b->first || a->second.getString()!=b->second.getString()) same=false;
					a=c.snps.begin(); b=g.snps.begin();
					for (;a!=c.snps.end();a++,b++) if (a->first!=b->first || a->second.getString()!=b->second.getString()) same=false;
				}
				nPairs++;
				if (!same) { nBad++; cerr << "B200CHECK mismatch leftPos " << leftPos << " hap " << hidx << " read " << r << " " << bam1_qname(reads[r].getBam()) << " ll " << c.ll << " vs " << g.ll << endl; }
			}
			cerr << "B200CHECK window leftPos " << leftPos << " pairs " << haps.size()*reads.size() << " mismatches " << (nBad-bad0) << " total_pairs " << nPairs << " total_mismatches " << nBad << endl;
		}
"""
if check:
    fwd = fwd.replace("\t\treturn;\n", CHECK + "\t\treturn;\n")
if not dump:
    text = text[:m.end()] + fwd + text[m.end():]
    # third line of the patch: DetInDel::filterHaplotypes (DInDel.cpp:1932-2100), the only consumer of
    # MLAlignment::hpos, forwards to the GPU where K1 left the hpos rows (K4, dindel_b200_coverage)
    fsig = ("void DetInDel::filterHaplotypes(const vector<Haplotype> & haps, const vector<Read> & reads, const vector<vector<MLAlignment> > & liks,"
            "  vector<int> & filtered, map<pair<int, AlignedVariant>, VariantCoverage> & varCoverage, bool doFilter)\n{\n")
    assert text.count(fsig) == 1
    ffwd = ("\tif (dindel_b200_replay && dindel_b200_replay->w >= 0) { // batched driver: K4 ran with the batch\n"
            "\t\tint nf = dindel_b200_replay->batch->filterResult(dindel_b200_replay->w, filtered, varCoverage);\n"
            "\t\tcout << \"Filtered \" << nf << \" haplotypes.\" << endl;\n"
            "\t\treturn;\n"
            "\t}\n"
            "\tif (params.slower && dindel_b200_gpu && !dindel_b200_replay && !getenv(\"DINDEL_B200_HOST_FILTER\")) { // dindel_b200 drop-in (INTEGRATION.md section 0)\n"
            "\t\tint nf = dindel_b200_gpu->filterHaplotypes(haps, reads, liks, filtered, varCoverage, doFilter, params.obsParams.maxMismatch);\n"
            "\t\tcout << \"Filtered \" << nf << \" haplotypes.\" << endl;\n"
            "\t\treturn;\n"
            "\t}\n")
    text = text.replace(fsig, fsig + ffwd)
    # fourth line of the patch: the per-read loops around DetInDel::getCIGAR (--outputRealignedBAM, DInDel.cpp:500-513
    # pooled / :596-611 diploid) take the CIGARs K5 (dindel_b200_cigars) derived on the GPU from the resident ll / hpos
    # rows; $DINDEL_B200_HOST_CIGAR keeps the reference's own loop (used to compare the two inside one binary)
    decl = "\t\t\t\tvector < CIGAR > cigars(reads.size());\n"
    assert text.count(decl) == 2
    gcall = ("\t\t\t\tvector < CIGAR > gpuCigars; // dindel_b200 drop-in (INTEGRATION.md section 0)\n"
             "\t\t\t\tif (dindel_b200_gpu && !getenv(\"DINDEL_B200_HOST_CIGAR\")) dindel_b200_gpu->realignedCigars(haps, reads.size(), refSeqPos, H1, H2, gpuCigars);\n")
    i1 = text.index(decl)
    text = text[:i1] + decl + gcall.replace("H1", "-1").replace("H2", "-1") + text[i1 + len(decl):]
    i2 = text.index(decl, i1 + len(decl))
    text = text[:i2] + decl + gcall.replace("H1", "hpl.h1").replace("H2", "hpl.h2") + text[i2 + len(decl):]
    for hx in ("hidx", "hmax"):
        call = "\t\t\t\t\t\tcigars[r]=getCIGAR(haps[%s], reads[r], liks[%s][r], refSeqPos);\n" % (hx, hx)
        assert text.count(call) == 1
        text = text.replace(call, "\t\t\t\t\t\tif (!gpuCigars.empty()) cigars[r]=gpuCigars[r]; else\n" + call)

    # fifth part of the patch: the numeric cores of diploidGLF and of the Bayes EM take their
    # read sums / EM results from K2 / K3; every comparison and printed field that depends on them goes through
    # dindel_b200::Inference's guard band, and DINDEL_B200_TWO_PASS re-runs an undecided window with the reference's
    # own CPU loops, so glf / VCF stay byte-identical.  Each entry: (anchor, how many times it occurs, which occurrence,
    # "before" / "after" / "replace", new text).
    INFER = [
        # ---- diploidGLF (DInDel.cpp:2933-3662) ----
        ("void DetInDel::diploidGLF(const vector<Haplotype> & haps, const vector<Read> & reads, const vector<vector<MLAlignment> > & liks, vector<double> & hapFreqs, vector <HapEstResult > & posteriors,  uint32_t candPos, uint32_t leftPos, uint32_t rightPos,  OutputData & glfData, int index, const AlignedCandidates & candidateVariants, string program=\"all\")\n\n{\n",
         1, 0, "after",
         "\tif (dindel_b200_replay) dindel_b200_replay->batch->bindSums(dindel_b200_replay->w, dindel_b200_replay->inf);\n"
         "\telse if (dindel_b200_gpu && dindel_b200_gpu->wantSums()) DINDEL_B200_INF.setSums(dindel_b200_gpu->windowSumsPtr(), dindel_b200_gpu->windowHaps(), dindel_b200_gpu->windowReads(), 0, 0);\n"
         "\tDINDEL_B200_TWO_PASS(DINDEL_B200_INF, glfData, ;, diploidGLF(haps, reads, liks, hapFreqs, posteriors, candPos, leftPos, rightPos, glfData, index, candidateVariants, program))\n"),
        ("\t\tdouble ll=0.0;\n\t\tfor (size_t r=0;r<reads.size();r++){\n\t\t\tll+=log(0.5)+addLogs(rl[r*nh+h1],rl[r*nh+h2]);\n", 1, 0, "replace",
         "\t\tdouble ll=0.0;\n\t\tif (DINDEL_B200_INF.gpu()) ll = DINDEL_B200_INF.sum(h1, h2, nh, ll); else\n\t\tfor (size_t r=0;r<reads.size();r++){\n\t\t\tll+=log(0.5)+addLogs(rl[r*nh+h1],rl[r*nh+h2]);\n"),
        ("\t\tif (pairs_posterior[h1*nh+h2]>max_ll_indel && (", 1, 0, "replace",
         "\t\tif (DINDEL_B200_INF.gt(pairs_posterior[h1*nh+h2], max_ll_indel) && ("),
        ("\t\tif (pairs_posterior[h1*nh+h2]>max_ll_noindel && (", 1, 0, "replace",
         "\t\tif (DINDEL_B200_INF.gt(pairs_posterior[h1*nh+h2], max_ll_noindel) && ("),
        ("\tqual = - 10.0*( ll_ref - addLogs(max_ll_indel, ll_ref) )/log(10.0);\n", 1, 0, "after",
         "\tDINDEL_B200_INF.checkPhred(ll_ref, max_ll_indel);\n"),
        ("\t\t\tgenoqual = - 10.0*( max_ll_altgeno - addLogs(max_ll_indel, max_ll_altgeno) )/log(10.0);\n", 1, 0, "after",
         "\t\t\tDINDEL_B200_INF.checkPhred(max_ll_altgeno, max_ll_indel);\n"),
        ("\t\t\tdouble ll=pr;\n\t\t\tfor (size_t r=0;r<reads.size();r++){\n", 1, 0, "replace",
         "\t\t\tdouble ll=pr;\n\t\t\tif (DINDEL_B200_INF.gpu()) ll = DINDEL_B200_INF.sum(h1, h2, nh, ll); else\n\t\t\tfor (size_t r=0;r<reads.size();r++){\n"),
        ("\t\t\tif (ll>maxll) {\n", 2, 1, "replace", "\t\t\tif (DINDEL_B200_INF.gt(ll, maxll)) {\n"),   # the first one (:297) is in an unused function
        ("\t\t\to << str << a1 << \"/\" << a2 << \":\" << git->second;\n", 1, 0, "replace",
         "\t\t\to << str << a1 << \"/\" << a2 << \":\" << DINDEL_B200_INF.printed(git->second);\n"),
        ("\t\t\tline.set(\"logZ\", maxll );\n", 1, 0, "replace", "\t\t\tline.set(\"logZ\", DINDEL_B200_INF.printed(maxll) );\n"),
        # ---- K6: the per-read aggregation loops behind a called site ----
        # dip.map sites (DInDel.cpp:3126-3170): every site of the window in one K6 call before the site loop
        ("\t\tfor (map<int, std::set<AlignedVariant> >::const_iterator it = indels.begin();it!=indels.end();it++) {\n\t\t\t\n\t\t\tdouble msq = 0;\n", 1, 0, "before",
         "\t\tstd::vector<dindel_read_stats> b200_ms; bool b200_map = false; size_t b200_mi = 0;\n"
         "\t\t{\n"
         "\t\t\tstd::vector<dindel_b200::StatsJob> jobs_;\n"
         "\t\t\tconst int m_ = (max_indel_pair[0]==max_indel_pair[1]) ? 1 : 2;\n"
         "\t\t\tfor (map<int, std::set<AlignedVariant> >::const_iterator it = indels.begin();it!=indels.end();it++) {\n"
         "\t\t\t\tjobs_.push_back(dindel_b200::StatsJob(DINDEL_STATS_MAP, -1));\n"
         "\t\t\t\tfor (int i=0;i<2;i++) jobs_.back().cands.push_back(dindel_b200::mapCand(haps[max_indel_pair[i]], max_indel_pair[i], it->first, i<m_));\n"
         "\t\t\t}\n"
         "\t\t\tb200_map = DINDEL_B200_STATS(jobs_, reads.size(), b200_ms);\n"
         "\t\t\tif (b200_map && !b200_ms.empty()) numUnmappedRealigned = b200_ms[0].numUnmappedRealigned;\n"
         "\t\t}\n"),
        ("\t\t\tfor (int i=0;i<m;i++) {\n\t\t\t\tint h=max_indel_pair[i];\n\t\t\t\tIt iter = haps[h].indels.find(it->first);\n", 1, 0, "before",
         "\t\t\tif (b200_map) { const dindel_read_stats & s_ = b200_ms[b200_mi++]; msq = s_.msq; numf = s_.nf; numr = s_.nr; n = s_.n; } else\n"),
        # variant positions (:3396-3404, :3504-3554): one K6 call per position, once its arg-max pair is known
        ("\t\tint numUnmappedRealigned = 0;\n\t\tfor (size_t r=0;r<reads.size();r++) {\n", 1, 0, "replace",
         "\t\tint numUnmappedRealigned = 0;\n"
         "\t\tstd::vector<dindel_read_stats> b200_ss; bool b200_site = false;\n"
         "\t\t{\n"
         "\t\t\tstd::vector<dindel_b200::StatsJob> jobs_(1, dindel_b200::StatsJob(DINDEL_STATS_SITE, -1));\n"
         "\t\t\tjobs_[0].cands.push_back(dindel_b200::siteCand(haps[hx1], hx1, pos)); jobs_[0].cands.push_back(dindel_b200::siteCand(haps[hx2], hx2, pos));\n"
         "\t\t\tb200_site = DINDEL_B200_STATS(jobs_, reads.size(), b200_ss);\n"
         "\t\t\tif (b200_site) numUnmappedRealigned = b200_ss[0].numUnmappedRealigned;\n"
         "\t\t}\n"
         "\t\tif (!b200_site) for (size_t r=0;r<reads.size();r++) {\n"),
        ("\t\tBOOST_FOREACH(int r, readidx) {\n\t\t\tdouble mq=-10*log10(1.0-reads[r].mapQual);\n\t\t\tallmsq+=(mq*mq);\n", 1, 0, "before",
         "\t\tif (b200_site) { const dindel_read_stats & s_ = b200_ss[0]; allmsq = s_.allmsq; numMappedIndels = s_.numIndels; nBQT = s_.nBQT; nmmBQT = s_.nmmBQT; "
         "mLogBQ = s_.mLogBQ; nMMLeft = s_.nMMLeft; nMMRight = s_.nMMRight; numOffBoth = s_.numOffBoth; nf = s_.nf; nr = s_.nr; msq = s_.msq; n = s_.n; } else\n"),
        # pooled variants (:2697-2742): every pool member of the variant in one K6 call
        ("\t\tif (1) for (size_t b=0;b<myBams.size();b++) {\n\t\t\tdouble msq=0.0;\n", 1, 0, "before",
         "\t\tstd::vector<dindel_read_stats> b200_ps; std::vector<int> b200_pj(myBams.size(), -1); bool b200_pool = false;\n"
         "\t\t{\n"
         "\t\t\tstd::vector<dindel_b200::StatsJob> jobs_;\n"
         "\t\t\tsize_t most_ = 0;\n"
         "\t\t\tfor (size_t b=0;b<myBams.size();b++) if (readidx[b].size()) {\n"
         "\t\t\t\tb200_pj[b] = int(jobs_.size());\n"
         "\t\t\t\tjobs_.push_back(dindel_b200::StatsJob(DINDEL_STATS_POOL, int(b)));\n"
         "\t\t\t\tfor (size_t h=0;h<nh;h++) jobs_.back().cands.push_back(dindel_b200::poolCand(haps[h], int(h), pav.first, pav.second, hapHasVar[h*nv+idx]!=0));\n"
         "\t\t\t\tif (readidx[b].size() > most_) most_ = readidx[b].size();\n"
         "\t\t\t}\n"
         "\t\t\tb200_pool = DINDEL_B200_STATS(jobs_, most_, b200_ps);\n"
         "\t\t}\n"),
        ("\t\t\t\tBOOST_FOREACH(int r, readidx[b]) {\n\t\t\t\t\t//cout << \" \" << 10*log10(1-reads[r].mapQual);\n\n\t\t\t\t\tsize_t mlidx;", 1, 0, "before",
         "\t\t\t\tif (b200_pool && b200_pj[b] >= 0) { const dindel_read_stats & s_ = b200_ps[b200_pj[b]]; msq = s_.msq; n = s_.n; nf = s_.nf; nr = s_.nr; } else\n"),
        # ---- estimateHaplotypeFrequenciesBayesEM (DInDel.cpp:2103-2930) ----
        ("void DetInDel::estimateHaplotypeFrequenciesBayesEM(const vector<Haplotype> & haps, const vector<Read> & reads, const vector<vector<MLAlignment> > & liks, vector<double> & hapFreqs, vector <HapEstResult > & posteriors,  uint32_t candPos, uint32_t leftPos,   uint32_t rightPos, OutputData & glfData, int index, const AlignedCandidates & candidateVariants, string program=\"all\")\n\n{\n",
         1, 0, "after",
         "\tif (dindel_b200_replay) dindel_b200_replay->batch->bindSums(dindel_b200_replay->w, dindel_b200_replay->inf);\n"
         "\telse if (dindel_b200_gpu && dindel_b200_gpu->wantSums()) DINDEL_B200_INF.setSums(dindel_b200_gpu->windowSumsPtr(), dindel_b200_gpu->windowHaps(), dindel_b200_gpu->windowReads(), dindel_b200_gpu->windowPools(), dindel_b200_gpu->windowPoolSizes().empty() ? 0 : &dindel_b200_gpu->windowPoolSizes()[0]);\n"
         "\tDINDEL_B200_TWO_PASS(DINDEL_B200_INF, glfData, posteriors.clear(), estimateHaplotypeFrequenciesBayesEM(haps, reads, liks, hapFreqs, posteriors, candPos, leftPos, rightPos, glfData, index, candidateVariants, program))\n"),
        ("\n\tfor (int th=0;th<nav;th++) {\n", 1, 0, "before",
         "\n\tif (DINDEL_B200_INF.gpu()) { // every EM run of the window in one K3 launch: compatible[] of each set as at :2394-2408 (checked again set by set in em())\n"
         "\t\tstd::vector<uint8_t> b200c(size_t(nav)*nh, 0);\n"
         "\t\tfor (int th=0;th<nav;th++) for (size_t h=0;h<nh;h++) {\n"
         "\t\t\tint c = filtered[h]==0;\n"
         "\t\t\tif (c) for (It it=haps[h].indels.begin();it!=haps[h].indels.end();it++)\n"
         "\t\t\t\tif (!it->second.isRef() && !(it->second.isSNP() && it->second.getString()[3]=='D') && activeVariants[th].find(PAV(it->first,it->second))==activeVariants[th].end()) { c=0; break; }\n"
         "\t\t\tb200c[size_t(th)*nh+h]=c;\n"
         "\t\t}\n"
         "\t\tif (dindel_b200_replay) dindel_b200_replay->inf.emBegin(*dindel_b200_replay, b200c, nav, params.bayesa0, params.EMtol);\n"
         "\t\telse DINDEL_B200_INF.emBegin(*dindel_b200_gpu, b200c, nav, params.bayesa0, params.EMtol);\n"
         "\t}\n"),
        ("\t\twhile (!converged) {\n", 1, 0, "before",
         "\t\tif (DINDEL_B200_INF.em(th, compatible, tol, loglik, pi, iter)) converged = true; // K3 ran this EM\n"),
        ("\t\t\t\t\tBOOST_FOREACH(int r, readidx[b]) {\n\t\t\t\t\t\tll+=log5+addLogs(rl[r*nh+h1],rl[r*nh+h2]);\n", 1, 0, "before",
         "\t\t\t\t\tif (DINDEL_B200_INF.gpu()) ll = DINDEL_B200_INF.sumGroup(int(b), h1, h2, nh, ll); else\n"),
        ("\t\t\t\tline.set(\"post_prob_variant\", exp(logp));\n\t\t\t\tline.set(\"est_freq\", freq);\n\t\t\t\tline.set(\"logZ\", logz);\n", 1, 0, "replace",
         "\t\t\t\tline.set(\"post_prob_variant\", DINDEL_B200_INF.printedLin(exp(logp)));\n\t\t\t\tline.set(\"est_freq\", DINDEL_B200_INF.printedLin(freq));\n\t\t\t\tline.set(\"logZ\", DINDEL_B200_INF.printed(logz));\n"),
        ("\t\t\t\t\t\tif (hapFreqs[h]>1.0/double(2*reads.size())) {\n", 1, 0, "replace",
         "\t\t\t\t\t\tif (DINDEL_B200_INF.gtLin(hapFreqs[h], 1.0/double(2*reads.size()))) {\n"),
        ("\t\t\t\t\t\t\tos << \":\" << hapFreqs[h];\n", 1, 0, "replace",
         "\t\t\t\t\t\t\tos << \":\" << DINDEL_B200_INF.printedLin(hapFreqs[h]);\n"),
        ("\t\t\t\t\to << lik[n];\n", 1, 0, "replace", "\t\t\t\t\to << DINDEL_B200_INF.printed(lik[n]);\n"),
    ]
    if not check:
        INFER += [
            ("bool DetInDel::getHaplotypes(vector<Haplotype> &haps, const vector<Read> & reads,uint32_t pos, uint32_t & leftPos, uint32_t & rightPos, const AlignedCandidates & candidateVariants)\n{\n", 1, 0, "after",
             "\tif (dindel_b200_replay) { // batched driver: prepared before the batch was flushed\n"
             "\t\thaps = *dindel_b200_replay->haps; leftPos = dindel_b200_replay->leftPos; rightPos = dindel_b200_replay->rightPos;\n"
             "\t\treturn dindel_b200_replay->skip;\n"
             "\t}\n"),
            ("void DetInDel::detectIndels(const string & variantsFileName)\n{\n", 1, 0, "after",
             "\t// dindel_b200: the window loop batched (prepare on host threads -> one GPU pass per batch -> finish on host threads)\n"
             "\tif (params.slower && params.inferenceMethod==\"empirical\" && !params.outputRealignedBAM && !params.outputPooledLikelihoods && !params.showHapAlignments && !getenv(\"DINDEL_B200_NO_BATCH\")) {\n"
             "\t\tconst bool infer = !getenv(\"DINDEL_B200_HOST_INFERENCE\") && (params.doDiploid != params.estimateHapFreqs);\n"
             "\t\tdindel_b200_driver drv(*this, dindel_b200_engine_factory, this, dindel_b200_bam_arg, dindel_b200_multiple_files, infer);\n"
             "\t\tint nt = getenv(\"DINDEL_B200_THREADS\") ? atoi(getenv(\"DINDEL_B200_THREADS\")) : int(sysconf(_SC_NPROCESSORS_ONLN));\n"
             "\t\tif (nt < 1) nt = 1; if (nt > 64) nt = 64;\n"
             "\t\tdrv.run(variantsFileName, nt, getenv(\"DINDEL_B200_BATCH\") ? size_t(atol(getenv(\"DINDEL_B200_BATCH\"))) : size_t(4096));\n"
             "\t\treturn;\n"
             "\t}\n"),
            ("\t\t\tdetInDel.detectIndels(varFile);\n", 1, 0, "before",
             "\t\t\tdindel_b200_bam_arg = file; dindel_b200_multiple_files = multipleFiles; // the batched driver opens one DetInDel per worker thread\n"),
        ]
    for anchor, count, which, where, code in INFER:
        assert text.count(anchor) == count, (anchor, text.count(anchor))
        i = -1
        for _ in range(which + 1):
            i = text.index(anchor, i + 1)
        if where == "before":
            text = text[:i] + code + text[i:]
        elif where == "after":
            text = text[:i + len(anchor)] + code + text[i + len(anchor):]
        else:
            text = text[:i] + code + text[i + len(anchor):]

# --dump: oracle/_ref/dindel_ref_dump = the STOCK CPU driver whose filterHaplotypes (DInDel.cpp:1932-2100)
# appends, at its end, its inputs and outputs as text to the file named by $DINDEL_B200_DUMP_COVER.  This is
# how tests/golden/cover_golden.json.gz is produced (tests/golden/make_cover_golden.py): golden vectors of
# the reference's own filterHaplotypes, against which the a13 oracle restatement and the K4 kernel are pinned.
DUMP = r"""
	if (getenv("DINDEL_B200_DUMP_COVER")) {
		FILE *df = fopen(getenv("DINDEL_B200_DUMP_COVER"), "a");
		fprintf(df, "W %d %d %d %d %d\n", int(haps.size()), int(reads.size()), params.obsParams.padCover, params.obsParams.maxMismatch, int(doFilter));
		for (size_t h=0;h<haps.size();h++) fprintf(df, "H %s\n", haps[h].seq.c_str());
		for (size_t r=0;r<reads.size();r++) {
			int strand = 0;
			if (reads[r].isUnmapped()) { if (!reads[r].mateIsReverse()) strand = 1; } else { if (reads[r].isReverse()) strand = 1; }
			fprintf(df, "R %s %d\n", reads[r].seq.seq.c_str(), strand);
		}
		for (size_t h=0;h<haps.size();h++) for (size_t r=0;r<reads.size();r++) {
			fprintf(df, "P %d %d %d %d", int(h), int(r), int(liks[h][r].offHapHMQ), liks[h][r].numIndels);
			for (size_t b=0;b<liks[h][r].hpos.size();b++) fprintf(df, " %d", liks[h][r].hpos[b]);
			fprintf(df, "\n");
		}
		for (size_t h=0;h<haps.size();h++) for (map<int, AlignedVariant>::const_iterator it = haps[h].indels.begin();it!= haps[h].indels.end();it++) {
			const AlignedVariant & av = it->second;
			int ty = av.getType() == Variant::DEL ? 0 : (av.getType() == Variant::INS ? 1 : 2);
			fprintf(df, "E %d %d|%d|%s %d %d %d\n", int(h), it->first, av.getStartHap(), av.getString().c_str(), ty, av.getLeftFlankRead() - params.obsParams.padCover, av.getRightFlankRead() + params.obsParams.padCover);
		}
		fprintf(df, "F"); for (size_t h=0;h<haps.size();h++) fprintf(df, " %d", filtered[h]); fprintf(df, "\n");
		for (map<PAV, VariantCoverage>::const_iterator it = varCoverage.begin(); it != varCoverage.end(); it++)
			fprintf(df, "C %d|%d|%s %d %d\n", it->first.first, it->first.second.getStartHap(), it->first.second.getString().c_str(), it->second.nf, it->second.nr);
		fclose(df);
	}
"""
# --dump, second hook: the two call sites of DetInDel::getCIGAR (DInDel.cpp:511 pooled / :610 diploid, the
# --outputRealignedBAM option) append, per read, the inputs of the haplotype choice (:502-509, :597-606) and of
# getCIGAR (:728-882: the read's hpos, the haplotype's hap.ml.hpos, refSeqPos) and the resulting CIGAR to the file
# named by $DINDEL_B200_DUMP_CIGAR: golden vectors for the a-next row "realigned-BAM CIGARs" (K5,
# tests/golden/make_cigar_golden.py).
DUMP_CIG = r"""
						if (getenv("DINDEL_B200_DUMP_CIGAR")) {
							FILE *df = fopen(getenv("DINDEL_B200_DUMP_CIGAR"), "a");
							const int hx = HX;
							fprintf(df, "G MODE %d %d %d %d %d %d %d", H1, H2, int(haps.size()), int(onHap[r]), hx, refSeqPos, cigars[r].refPos);
							fprintf(df, " L"); for (size_t h=0;h<haps.size();h++) fprintf(df, " %.17g", liks[h][r].ll);
							fprintf(df, " N"); for (size_t h=0;h<haps.size();h++) fprintf(df, " %d", haps[h].countIndels());
							fprintf(df, " P"); for (size_t b=0;b<liks[hx][r].hpos.size();b++) fprintf(df, " %d", liks[hx][r].hpos[b]);
							fprintf(df, " M"); for (size_t b=0;b<haps[hx].ml.hpos.size();b++) fprintf(df, " %d", haps[hx].ml.hpos[b]);
							fprintf(df, " C"); for (size_t c=0;c<cigars[r].size();c++) fprintf(df, " %d %d", cigars[r][c].first, cigars[r][c].second);
							fprintf(df, "\n");
							fclose(df);
						}
"""
# --dump, third hook: the numeric cores of DetInDel::diploidGLF (DInDel.cpp:3085-3120, :3240-3266, :3349-3390) and of
# DetInDel::estimateHaplotypeFrequenciesBayesEM (:2431-2545, :2683-2694) append their inputs (the rl matrix, priors,
# filtered[], compatible[]) and their intermediate results (per-pair sums, arg-max pairs, qual / genoqual, per-position
# genotype likelihoods; per active-variant set: iterations, loglik, pi, freqs; logz; pooled genotype likelihoods) to the
# file named by $DINDEL_B200_DUMP_INFER, doubles as %.17g.  Golden vectors of the reference's own arithmetic for K2 / K3
# (tests/golden/make_infer_golden.py -> tests/golden/infer_golden.json.gz).
def DF(body):
    return ('if (getenv("DINDEL_B200_DUMP_INFER")) { FILE *df = fopen(getenv("DINDEL_B200_DUMP_INFER"), "a"); '
            + body + ' fclose(df); }\n')
DUMP_INF = [
    # diploidGLF: rl matrix + filtered (after the rl loop, DInDel.cpp:2962)
    ("\t// get all variants\n", "before", 1, 0,
     DF('fprintf(df, "DW %d %d\\n", int(nh), int(nr)); fprintf(df, "RL"); for (size_t i=0;i<nh*nr;i++) fprintf(df, " %.17g", rl[i]); '
        'fprintf(df, "\\nFILT"); for (size_t h=0;h<nh;h++) fprintf(df, " %d", filtered[h]); fprintf(df, "\\n");')),
    ("\t\tpairs_posterior[h1*nh+h2] = ll + prior[h1*nh+h2];\n", "after", 1, 0,
     DF('fprintf(df, "PP %d %d %.17g %.17g %d %d\\n", int(h1), int(h2), ll, prior[h1*nh+h2], hap_num_candidate_indels[h1], hap_num_candidate_indels[h2]);')),
    ("\tif (max_indel_pair[0]==-1 || max_indel_pair[1]==-1) throw string(\"Could not find indel allele\");\t\n", "before", 1, 0,
     DF('fprintf(df, "MAP %d %d %d %d %.17g %.17g %.17g\\n", max_indel_pair[0], max_indel_pair[1], max_noindel_pair[0], max_noindel_pair[1], max_ll_indel, max_ll_noindel, qual);')),
    ("\t\t\tgenoqual = - 10.0*( max_ll_altgeno - addLogs(max_ll_indel, max_ll_altgeno) )/log(10.0);\n", "after", 1, 0,
     DF('fprintf(df, "GQ %d %.17g %.17g\\n", it->first, max_ll_altgeno, genoqual);')),
    ("\t\t\tIGIt igit = genLiks.find(genotype);\n", "before", 1, 0,
     DF('fprintf(df, "SL %d %d %d %d %d %.17g %.17g\\n", pos, int(h1), int(h2), v1, v2, pr, ll);')),
    ("\t\t//cout << \"hx1: \" << hx1 << \" hx2: \" << hx2 << \" postprob: \" << maxll << endl;\n", "after", 1, 0,
     DF('fprintf(df, "SM %d %d %d %.17g", pos, hx1, hx2, maxll); for (IGIt g=genLiks.begin();g!=genLiks.end();g++) fprintf(df, " %d/%d:%.17g", *(g->first.begin()), *(g->first.rbegin()), g->second); fprintf(df, "\\n");')),
    # estimateHaplotypeFrequenciesBayesEM: rl matrix (filled at :2125-2139), then per set, then logz, then pooled genotype sums
    ("\tcout << \"ALWAYS CALLING ::filterHaplotypes\" << endl;\n", "before", 1, 0,
     DF('fprintf(df, "EW %d %d %.17g %.17g\\n", int(nh), int(nr), params.bayesa0, params.EMtol); fprintf(df, "RL"); for (size_t i=0;i<nh*nr;i++) fprintf(df, " %.17g", rl[i]); '
        'fprintf(df, "\\nPOOL"); for (size_t r=0;r<nr;r++) fprintf(df, " %d", reads[r].poolID); fprintf(df, "\\n");')),
    ("\t\tlogliks[th]=loglik;\n", "after", 1, 0,
     DF('fprintf(df, "SET %d %d %d %.17g %.17g C", th, numah, iter, loglik, logprior); for (size_t h=0;h<nh;h++) fprintf(df, " %d", compatible[h]); '
        'fprintf(df, " PI"); for (size_t h=0;h<nh;h++) fprintf(df, " %.17g", pi[h]); fprintf(df, " E %.17g\\n", eOld);')),
    ("\tfor (int a=0;a<nav;a++) {\n\t\tpost[a]=exp(logliks[a]+logpriors[a]-logz);\n", "before", 1, 0,
     DF('fprintf(df, "LOGZ %.17g FR", logz); for (size_t i=0;i<size_t(nav)*nh;i++) fprintf(df, " %.17g", freqs[i]); fprintf(df, "\\n");')),
    ("\t\tlogp-=logz;\n", "after", 1, 0,
     DF('fprintf(df, "VAR %d %.17g %.17g HF", idx, logp, freq); for (size_t h=0;h<nh;h++) fprintf(df, " %.17g", hapFreqs[h]); fprintf(df, " HV"); for (size_t h=0;h<nh;h++) fprintf(df, " %d", hapHasVar[h*nv+idx]); fprintf(df, "\\n");')),
    ("\t\t\t\t\tlik[genotype]=addLogs(lik[genotype],ll);\n", "before", 1, 0,
     DF('fprintf(df, "PG %d %d %d %d %.17g %.17g\\n", idx, int(b), int(h1), int(h2), pr, ll);')),
]
# --dump, fourth hook: the per-read aggregation loops behind a called site (DInDel.cpp:3126-3170 "dip.map" sites,
# :3396-3404 + :3504-3554 per variant position, :2697-2742 pooled per variant and pool member) append what they read
# (per read: mapQual / strand / unmapped / poolID; per (haplotype, read): the MLAlignment fields; per candidate
# haplotype: whether the coverage map has the site and startRead / endRead of its variant) and what they computed to
# the same file: golden vectors for K6 (tests/golden/make_stats_golden.py -> tests/golden/stats_golden.json.gz).
PEEK = ('struct dindel_b200_peek : AlignedVariant { static int end_read(const AlignedVariant & a) '
        '{ int AlignedVariant::*p = &dindel_b200_peek::endRead; return a.*p; } };\n')
RDLK = ('for (size_t r=0;r<reads.size();r++) fprintf(df, "RD %d %.17g %d %d %d\\n", int(r), reads[r].mapQual, int(reads[r].onReverseStrand), int(reads[r].isUnmapped()), reads[r].poolID); '
        'for (size_t h=0;h<haps.size();h++) for (size_t r=0;r<reads.size();r++) { const MLAlignment & m_ = liks[h][r]; '
        'fprintf(df, "LK %d %d %.17g %d %d %d %d %.17g %d %d %d %d\\n", int(h), int(r), m_.ll, int(m_.offHap), int(m_.indels.size()), m_.nBQT, m_.nmmBQT, m_.mLogBQ, m_.nMMLeft, m_.nMMRight, m_.firstBase, m_.lastBase); }')
CAND = ('{ int pres_=0, st_=0, en_=0; map<int,AlignedVariant>::const_iterator f_; '
        'if (KIND==1) { f_=haps[H_].indels.find(POS_); if (f_!=haps[H_].indels.end()) { pres_=1; st_=f_->second.getStartRead(); en_=dindel_b200_peek::end_read(f_->second); } } '
        'else if (KIND==2) { f_=haps[H_].snps.find(POS_); if (f_!=haps[H_].snps.end()) { pres_=1; st_=f_->second.getStartRead(); en_=dindel_b200_peek::end_read(f_->second); } } '
        'fprintf(df, " %d %d %d %d %d", int(H_), pres_, st_, en_, int(HV_)); }')
def cand(h, pos, kind, hv="0"):
    return CAND.replace("H_", h).replace("POS_", pos).replace("KIND", kind).replace("HV_", hv)
SITE_KIND = "(haps[H_].indels.find(pos)==haps[H_].indels.end() ? -1 : (haps[H_].indels.find(pos)->second.isIndel() ? 1 : (haps[H_].indels.find(pos)->second.isSNP() ? 2 : 0)))"
DUMP_STATS = [
    ("\t// get all variants\n", "before", 1, 0, DF(RDLK)),
    ("\tcout << \"ALWAYS CALLING ::filterHaplotypes\" << endl;\n", "before", 1, 0, DF(RDLK)),
    ("\n\t\t\tif (n!=0) msq=sqrt(msq/double(n)); else msq=0.0;\n", "before", 1, 0,
     DF('fprintf(df, "RSM %d %d", it->first, m); for (int i_=0;i_<2;i_++) { int h_=max_indel_pair[i_]; It q_=haps[h_].indels.find(it->first); '
        'int g_=(i_<m && q_!=haps[h_].indels.end() && q_->second.isIndel())?1:0; '
        'fprintf(df, " %d %d %d %d 0", h_, g_, g_?q_->second.getStartRead():0, g_?dindel_b200_peek::end_read(q_->second):0); } '
        'fprintf(df, " O %d %d %d %.17g %d\\n", numf, numr, n, msq, numUnmappedRealigned);')),
    ("\n\t\tif (n!=0) msq=sqrt(msq/double(n)); else msq=0.0;\n", "before", 1, 0,
     DF('fprintf(df, "RSS %d", pos); for (int i_=0;i_<2;i_++) { int h_=i_?hx2:hx1; ' + cand("h_", "pos", SITE_KIND.replace("H_", "h_")) + ' } '
        'fprintf(df, " O %d %d %d %.17g %.17g %d %d %d %d %.17g %d %d %d\\n", nf, nr, n, msq, allmsq, numOffBoth, numMappedIndels, nBQT, nmmBQT, mLogBQ, nMMLeft, nMMRight, numUnmappedRealigned);')),
    ("\n\t\t\t\tif (n!=0) msq=sqrt(msq/double(n)); else msq=0.0;\n", "before", 1, 0,
     DF('fprintf(df, "RSP %d %d %d", idx, int(b), pos); for (size_t h_=0;h_<nh;h_++) ' + cand("h_", "pos", "(pav.second.isIndel()?1:(pav.second.isSNP()?2:0))", "hapHasVar[h_*nv+idx]") +
        ' fprintf(df, " O %d %d %d %.17g\\n", nf, nr, n, msq);')),
]
if dump:
    assert text.count(inc) == 1
    text = text.replace(inc, inc + PEEK.replace("\\n", "\n"))
    for anchor, where, count, which, code in DUMP_INF + DUMP_STATS:
        assert text.count(anchor) == count, (anchor, text.count(anchor))
        i = -1
        for _ in range(which + 1):
            i = text.index(anchor, i + 1)
        if where == "before" and anchor.startswith("\n"):   # on a line of its own (the line before may end in a // comment)
            text = text[:i + 1] + code + text[i + 1:]
        else:
            text = text[:i] + (code + anchor if where == "before" else anchor + code) + text[i + len(anchor):]
    anchor = "\t\tvarCoverage[pav]=VariantCoverage(int(rf.size()), int(rr.size()));\n\t}\n"
    assert text.count(anchor) == 1
    text = text.replace(anchor, anchor + DUMP)
    a1 = "\t\t\t\t\t\tcigars[r]=getCIGAR(haps[hidx], reads[r], liks[hidx][r], refSeqPos);\n"
    a2 = "\t\t\t\t\t\tcigars[r]=getCIGAR(haps[hmax], reads[r], liks[hmax][r], refSeqPos);\n"
    assert text.count(a1) == 1 and text.count(a2) == 1
    text = text.replace(a1, a1 + DUMP_CIG.replace("HX", "hidx").replace("MODE", "B").replace("H1", "-1").replace("H2", "-1"))
    text = text.replace(a2, a2 + DUMP_CIG.replace("HX", "hmax").replace("MODE", "D").replace("H1", "hpl.h1").replace("H2", "hpl.h2"))
open(dst, "w", encoding="latin-1").write(text)
