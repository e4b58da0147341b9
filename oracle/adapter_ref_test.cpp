/* oracle/adapter_ref_test.cpp — TEST INFRASTRUCTURE ONLY (built into oracle/_ref/, see oracle/Makefile).
 *
 * Drop-in proof for include/dindel_b200.hpp: the adapter is instantiated with the reference's OWN
 * Haplotype / Read / MLAlignment / AlignedVariant types (headers included from /root/reference where
 * they lie) and its output is compared, field by field and map entry by map entry, with what the
 * reference's own loop body produces for the same window:
 *     ObservationModelFBMaxErr(hap, read, leftPos, obsParams).calcLikelihood()   (DInDel.cpp:1718-1719)
 * Needs a CUDA device (the adapter has no CPU path): run by tests/test_adapter.py on the GPU box.
 * Exit code 0 = every pair identical; prints the first difference otherwise.
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <cmath>
#include <iostream>
#include <sstream>
#include <string>
#include <vector>
#include "ObservationModelFB.hpp" /* the reference's: pulls Haplotype.hpp, Read.hpp, MLAlignment.hpp, bam.h */
#include "dindel_b200.hpp"
extern "C" {
#include "dindel_oracle.h" /* ora_choose_hap / ora_get_cigar: the restatement pinned on the reference's golden CIGARs */
}
struct RefCigar : public std::vector<std::pair<int, int> > { int refPos; }; /* DetInDel::CIGAR, DInDel.hpp:168-173 */
static bool same_cigar(const RefCigar &a, const RefCigar &b)
{
    return (const std::vector<std::pair<int, int> > &)a == (const std::vector<std::pair<int, int> > &)b && a.refPos == b.refPos;
}

static uint64_t rng_state = 0x9E3779B97F4A7C15ULL;
static uint32_t rnd()
{
    rng_state ^= rng_state << 13; rng_state ^= rng_state >> 7; rng_state ^= rng_state << 17;
    return (uint32_t)(rng_state >> 32);
}
static int rndint(int lo, int hi) { return lo + (int)(rnd() % (uint32_t)(hi - lo + 1)); }
static std::string rndseq(int n)
{
    std::string s(n, 'A');
    for (int i = 0; i < n; i++) s[i] = "ACGT"[rnd() & 3];
    return s;
}

/* Read.hpp:127-131,143-148 */
static double phred_to_prob(int phred)
{
    double q = (1.0 - pow(10.0, -(double)phred / 10.0));
    if (q < 1e-16) q = 1e-16;
    if (q > 1.0 - 1e-16) q = 1.0 - 1e-16;
    return q;
}

/* A Read as the BAM constructor leaves it (Read.hpp:123-180): sequence, probabilities converted from
 * the phred bytes, and a bam1_t whose data block holds those bytes where bam1_qual() looks. */
static void make_read(Read &r, const std::string &seq, const std::vector<uint8_t> &ph, int mapq, uint32_t start,
                      bool unmapped, bool reverse)
{
    const int R = (int)seq.size();
    r.seq.seq = seq;
    r.qual.resize(R);
    for (int b = 0; b < R; b++) r.qual[b] = phred_to_prob(ph[b]);
    r.mapQual = phred_to_prob(mapq);
    r.posStat.first = (double)start;
    r.posStat.second = 1.0;
    r.pos = (int32_t)start;
    r.onReverseStrand = reverse;
    r.bam = new bam1_t;
    memset(r.bam, 0, sizeof(bam1_t));
    const int nbytes = (R + 1) / 2 + R + 1;
    r.bam->data = new uint8_t[nbytes];
    memset(r.bam->data, 0, nbytes);
    r.bam->m_data = nbytes;
    r.bam->data_len = nbytes;
    r.bam->core.l_qseq = R;
    r.bam->core.qual = mapq;
    if (unmapped) r.bam->core.flag |= BAM_FUNMAP;
    if (reverse) r.bam->core.flag |= BAM_FREVERSE;
    memcpy(bam1_qual(r.bam), ph.empty() ? (const uint8_t *)"" : &ph[0], R);
    r.initBam = true;
}

static bool same_av(const AlignedVariant &a, const AlignedVariant &b)
{
    return a.getString() == b.getString() && a.getStartHap() == b.getStartHap() && a.getEndHap() == b.getEndHap() &&
           a.getLeftFlankRead() == b.getLeftFlankRead() && a.getRightFlankRead() == b.getRightFlankRead() &&
           a.getLeftFlankHap() == b.getLeftFlankHap() && a.getRightFlankHap() == b.getRightFlankHap() &&
           a.getType() == b.getType() && a.size() == b.size() && a.getSeq() == b.getSeq();
}
static bool same_avmap(const std::map<int, AlignedVariant> &a, const std::map<int, AlignedVariant> &b)
{
    if (a.size() != b.size()) return false;
    std::map<int, AlignedVariant>::const_iterator i = a.begin(), j = b.begin();
    for (; i != a.end(); ++i, ++j)
        if (i->first != j->first || !same_av(i->second, j->second)) return false;
    return true;
}
static bool same_bits(double a, double b) { return memcmp(&a, &b, sizeof(double)) == 0; }

/* stand-in for VariantCoverage (DInDel.hpp:53-66; DInDel.hpp itself drags in the whole driver) */
struct Cov {
    int nf, nr;
    Cov() : nf(0), nr(0) {}
    Cov(int f, int r) : nf(f), nr(r) {}
    bool operator==(const Cov &o) const { return nf == o.nf && nr == o.nr; }
};
typedef std::map<std::pair<int, AlignedVariant>, Cov> CovMap;
static bool same_cov(const CovMap &a, const CovMap &b)
{
    if (a.size() != b.size()) return false;
    CovMap::const_iterator i = a.begin(), j = b.begin();
    for (; i != a.end(); ++i, ++j)
        if (i->first.first != j->first.first || !same_av(i->first.second, j->first.second) || !(i->second == j->second)) return false;
    return true;
}
static bool same_mla(const MLAlignment &a, const MLAlignment &b)
{
    return same_bits(a.ll, b.ll) && same_bits(a.llOn, b.llOn) && same_bits(a.llOff, b.llOff) && same_bits(a.mLogBQ, b.mLogBQ) &&
           a.offHap == b.offHap && a.offHapHMQ == b.offHapHMQ && a.firstBase == b.firstBase && a.lastBase == b.lastBase &&
           a.numIndels == b.numIndels && a.numMismatch == b.numMismatch && a.nBQT == b.nBQT && a.nmmBQT == b.nmmBQT &&
           a.nMMLeft == b.nMMLeft && a.nMMRight == b.nMMRight && a.hpos == b.hpos && a.align == b.align &&
           same_avmap(a.indels, b.indels) && same_avmap(a.snps, b.snps) && a.hapIndelCovered == b.hapIndelCovered &&
           a.hapSNPCovered == b.hapSNPCovered;
}

#define CHECK(cond, what)                                                                              \
    do {                                                                                               \
        if (!(cond)) {                                                                                 \
            fprintf(stderr, "MISMATCH window %d hap %d read %d: %s\n", w, (int)h, (int)r, what);       \
            return 1;                                                                                  \
        }                                                                                              \
    } while (0)

int main(int argc, char **argv)
{
    const int n_windows = argc > 1 ? atoi(argv[1]) : 12;
    ObservationModelParameters op; /* struct defaults, then the CLI's values (DInDel.cpp:4122-4157, 3943-3948) */
    op.pError = 5e-4; op.pMut = 1e-5; op.maxLengthDel = 5; op.padCover = 2; op.mapQualThreshold = 100.0;
    dindel_b200::Engine *gpu = 0;
    try {
        gpu = new dindel_b200::Engine(op, 0);
    } catch (std::string &s) {
        fprintf(stderr, "no engine: %s\n", s.c_str());
        return 2;
    }
    long n_pairs = 0, n_ind = 0, n_snp = 0, n_cov = 0, n_filt = 0, n_covreads = 0;
    /* kept for the batch check below: every window's inputs and its one-window results */
    std::vector<std::vector<Haplotype> > allHaps;
    std::vector<std::vector<Read> > allReads;
    std::vector<uint32_t> allLeft;
    std::vector<std::vector<std::vector<MLAlignment> > > allLiks;
    std::vector<std::vector<int> > allOn, allFiltered;
    std::vector<CovMap> allCov;
    std::vector<std::vector<RefCigar> > allCigD, allCigP;
    std::vector<int> allH1, allH2, allRefPos;
    long n_cig = 0, n_cig_ops = 0;
    for (int w = 0; w < n_windows; w++) {
        const uint32_t leftPos = 100000u + 1000u * (uint32_t)w;
        const int H = rndint(100, 130), nh = rndint(2, 6), nr = rndint(8, 30);
        const std::string ref = rndseq(H);
        std::vector<Haplotype> haps(nh);
        haps[0].seq = ref;
        for (int b = 0; b < H; b++) haps[0].ml.hpos.push_back(b);
        for (int h = 1; h < nh; h++) {
            std::string s = ref;
            const int p = rndint(H / 3, 2 * H / 3), len = rndint(1, 5);
            std::string var;
            if (rnd() & 1) { var = std::string("-") + s.substr(p, len); s.erase(p, len); }
            else { std::string ins = rndseq(len); var = std::string("+") + ins; s.insert(p, ins); }
            haps[h].seq = s;
            /* what alignHaplotypes leaves in hap.ml.hpos (ObservationModelSeqAn.hpp:170-233): the haplotype's own
             * alignment to the reference segment, known here by construction */
            for (int b = 0; b < (int)s.size(); b++)
                haps[h].ml.hpos.push_back(var[0] == '-' ? (b < p ? b : b + len) : (b < p ? b : (b < p + len ? MLAlignment::INS : b - len)));
            /* what alignHaplotypes leaves in hap.indels / hap.snps (DInDel.cpp:1477-1482): variants with
             * their position on this haplotype; only isCovered() reads them on this path */
            haps[h].indels[p] = AlignedVariant(var, p, p + len - 1, p, p + (var[0] == '-' ? 0 : len - 1));
            if (rnd() & 1) {
                const int q = rndint(5, (int)s.size() - 6);
                std::string snp; snp += 'A'; snp += "=>"; snp += s[q];
                if (s[q] != 'A') haps[h].snps[q] = AlignedVariant(snp, q, q, q, q);
            }
        }
        std::vector<Read> reads(nr);
        for (int r = 0; r < nr; r++) {
            const std::string &src = haps[rnd() % nh].seq;
            const int R = rndint(30, 100);
            int off = rndint(-15, (int)src.size() - R + 15);
            std::string s;
            for (int b = 0; b < R; b++) {
                const int x = off + b;
                s += (x >= 0 && x < (int)src.size()) ? src[x] : "ACGT"[rnd() & 3];
            }
            std::vector<uint8_t> ph(R);
            for (int b = 0; b < R; b++) {
                ph[b] = (uint8_t)rndint(2, 41);
                if ((rnd() % 1000) < (uint32_t)(1000.0 * pow(10.0, -ph[b] / 10.0)) + 5) s[b] = "ACGT"[rnd() & 3];
            }
            if (rnd() % 7 == 0) { const int p = rndint(5, R - 10); s.erase(p, rndint(1, 3)); ph.resize(s.size()); }
            if (rnd() % 7 == 0) { const int p = rndint(5, (int)s.size() - 5); const std::string ins = rndseq(rndint(1, 3)); s.insert(p, ins); ph.resize(s.size(), 30); }
            if (rnd() % 23 == 0) s[rndint(0, (int)s.size() - 1)] = 'N';
            const int mq[4] = {60, 60, 40, 20};
            make_read(reads[r], s, ph, mq[rnd() & 3], (uint32_t)((int)leftPos + off + rndint(-2, 2)),
                      rnd() % 19 == 0, (rnd() & 1) != 0);
        }
        /* the adapter */
        std::vector<std::vector<MLAlignment> > liks;
        std::vector<int> onHap;
        gpu->computeLikelihoods(haps, reads, liks, leftPos, leftPos + (uint32_t)H, onHap);
        /* the reference's own loop body (DInDel.cpp:1715-1721), stdout silenced (BMIDERROR prints) */
        std::streambuf *old = std::cout.rdbuf();
        std::ostringstream sink;
        std::cout.rdbuf(sink.rdbuf());
        std::vector<int> refOn(nr, 0);
        for (size_t h = 0; h < haps.size(); h++)
            for (size_t r = 0; r < reads.size(); r++) {
                ObservationModelFBMaxErr oms(haps[h], reads[r], leftPos, op);
                const MLAlignment want = oms.calcLikelihood();
                const MLAlignment &got = liks[h][r];
                if (!want.offHapHMQ) refOn[r] = 1;
                std::cout.rdbuf(old);
                CHECK(same_bits(got.ll, want.ll), "ll");
                CHECK(same_bits(got.llOn, want.llOn), "llOn");
                CHECK(same_bits(got.llOff, want.llOff), "llOff");
                CHECK(same_bits(got.mLogBQ, want.mLogBQ), "mLogBQ");
                CHECK(got.offHap == want.offHap && got.offHapHMQ == want.offHapHMQ, "offHap / offHapHMQ");
                CHECK(got.firstBase == want.firstBase && got.lastBase == want.lastBase, "firstBase / lastBase");
                CHECK(got.numIndels == want.numIndels && got.numMismatch == want.numMismatch, "numIndels / numMismatch");
                CHECK(got.nBQT == want.nBQT && got.nmmBQT == want.nmmBQT, "nBQT / nmmBQT");
                CHECK(got.nMMLeft == want.nMMLeft && got.nMMRight == want.nMMRight, "nMMLeft / nMMRight");
                CHECK(got.relPos == want.relPos && got.hl == want.hl && got.hr == want.hr, "relPos / hl / hr");
                CHECK(got.hpos == want.hpos, "hpos");
                CHECK(got.align == want.align, "align");
                CHECK(same_avmap(got.indels, want.indels), "indels map");
                CHECK(same_avmap(got.snps, want.snps), "snps map");
                CHECK(got.hapIndelCovered == want.hapIndelCovered, "hapIndelCovered");
                CHECK(got.hapSNPCovered == want.hapSNPCovered, "hapSNPCovered");
                n_pairs++; n_ind += (long)want.indels.size(); n_snp += (long)want.snps.size();
                for (std::map<int, bool>::const_iterator it = want.hapIndelCovered.begin(); it != want.hapIndelCovered.end(); ++it)
                    n_cov += it->second ? 1 : 0;
                std::cout.rdbuf(sink.rdbuf());
            }
        std::cout.rdbuf(old);
        if (onHap != refOn) { fprintf(stderr, "MISMATCH window %d: onHap\n", w); return 1; }
        /* filterHaplotypes on the resident window (K4) */
        std::vector<int> filtered;
        CovMap cov;
        gpu->filterHaplotypes(haps, reads, liks, filtered, cov, true, op.maxMismatch);
        /* realigned-read CIGARs of the resident window (K5) against the restatement run on the REFERENCE's liks */
        std::vector<RefCigar> cigD, cigP;
        const int ch1 = (int)(rnd() % nh), ch2 = (int)(rnd() % nh);
        gpu->realignedCigars(haps, reads.size(), (int)leftPos, ch1, ch2, cigD);
        gpu->realignedCigars(haps, reads.size(), (int)leftPos, -1, -1, cigP);
        for (int mode = 0; mode < 2; mode++)
            for (size_t r = 0; r < reads.size(); r++) {
                const RefCigar &got = mode ? cigP[r] : cigD[r];
                if (mode && !onHap[r]) { if (!got.empty()) { fprintf(stderr, "MISMATCH window %d read %d: CIGAR for a read off all haplotypes\n", w, (int)r); return 1; } continue; }
                std::vector<double> ll(nh);
                std::vector<int32_t> ni(nh);
                for (int h = 0; h < nh; h++) { ll[h] = liks[h][r].ll; ni[h] = haps[h].countIndels(); }
                const int hx = ora_choose_hap(mode, nh, &ll[0], &ni[0], ch1, ch2);
                std::vector<int32_t> hp(liks[hx][r].hpos.begin(), liks[hx][r].hpos.end()), hm(haps[hx].ml.hpos.begin(), haps[hx].ml.hpos.end());
                int32_t refPos = 0, ops[64];
                const int n = ora_get_cigar((int)hp.size(), &hp[0], (int)hm.size(), &hm[0], (int)leftPos, &refPos, ops, 32);
                RefCigar want;
                for (int k = 0; k < n; k++) want.push_back(std::make_pair((int)ops[2 * k], (int)ops[2 * k + 1]));
                want.refPos = refPos;
                if (n < 1 || !same_cigar(got, want)) { fprintf(stderr, "MISMATCH window %d read %d mode %d: CIGAR (%d ops vs %d, refPos %d vs %d)\n", w, (int)r, mode, (int)got.size(), n, got.refPos, refPos); return 1; }
                n_cig++; n_cig_ops += n;
            }
        allCigD.push_back(cigD); allCigP.push_back(cigP); allH1.push_back(ch1); allH2.push_back(ch2); allRefPos.push_back((int)leftPos);
        allHaps.push_back(haps); allReads.push_back(reads); allLeft.push_back(leftPos);
        allLiks.push_back(liks); allOn.push_back(onHap); allFiltered.push_back(filtered); allCov.push_back(cov);
    }
    /* the throughput binding: all windows in ONE submit + ONE coverage call (dindel_b200::WindowBatch) must give
     * exactly what the window-at-a-time calls gave */
    for (int full = 1; full >= 0; full--) {
        dindel_b200::WindowBatch<Haplotype, Read> wb(*gpu, op.maxMismatch, true, full != 0);
        for (int w = 0; w < n_windows; w++)
            if (wb.add(allHaps[w], allReads[w], allLeft[w]) != w) { fprintf(stderr, "MISMATCH batch: window index\n"); return 1; }
        wb.flush();
        { /* every window's CIGARs in one K5 call == the window-at-a-time calls (also when the hpos rows never left the GPU) */
            std::vector<std::vector<RefCigar> > bD, bP;
            wb.cigars(allHaps, allRefPos, false, allH1, allH2, bD);
            wb.cigars(allHaps, allRefPos, true, allH1, allH2, bP);
            for (int w = 0; w < n_windows; w++)
                for (size_t r = 0; r < allReads[w].size(); r++)
                    if (!same_cigar(bD[w][r], allCigD[w][r]) || !same_cigar(bP[w][r], allCigP[w][r])) { fprintf(stderr, "MISMATCH batch (full %d) window %d read %d: CIGAR\n", full, w, (int)r); return 1; }
        }
        for (int w = 0; w < n_windows; w++) {
            std::vector<std::vector<MLAlignment> > liks;
            std::vector<int> onHap, filtered;
            CovMap cov;
            wb.result(w, allHaps[w], allReads[w], liks, onHap);
            wb.filterResult(w, filtered, cov);
            if (onHap != allOn[w] || filtered != allFiltered[w] || !same_cov(cov, allCov[w])) { fprintf(stderr, "MISMATCH batch window %d: onHap / filtered / coverage\n", w); return 1; }
            for (size_t h = 0; h < liks.size(); h++)
                for (size_t r = 0; r < liks[h].size(); r++) {
                    const MLAlignment &a = liks[h][r], &b = allLiks[w][h][r];
                    const bool ok = full ? same_mla(a, b)
                                         : (same_bits(a.ll, b.ll) && a.offHapHMQ == b.offHapHMQ && a.numIndels == b.numIndels && a.firstBase == b.firstBase &&
                                            a.lastBase == b.lastBase && a.hapIndelCovered == b.hapIndelCovered && a.hapSNPCovered == b.hapSNPCovered &&
                                            a.nBQT == b.nBQT && same_bits(a.mLogBQ, b.mLogBQ) && a.hpos.empty());
                    if (!ok) { fprintf(stderr, "MISMATCH batch (full %d) window %d hap %d read %d\n", full, w, (int)h, (int)r); return 1; }
                }
            if (full) {
                for (size_t h = 0; h < filtered.size(); h++) n_filt += filtered[h];
                for (CovMap::const_iterator it = cov.begin(); it != cov.end(); ++it) n_covreads += it->second.nf + it->second.nr;
            }
        }
    }
    if (n_covreads == 0) { fprintf(stderr, "MISMATCH: no covering reads counted at all\n"); return 1; }
    /* error behaviour: a haplotype shorter than maxLengthDel throws "hapSize error." like the reference
     * constructor (ObservationModelFB.cpp:47) */
    {
        std::vector<Haplotype> haps(2);
        haps[0].seq = rndseq(90); haps[1].seq = "ACG";
        std::vector<Read> reads(1);
        make_read(reads[0], haps[0].seq.substr(10, 50), std::vector<uint8_t>(50, 30), 60, 5010u, false, false);
        std::vector<std::vector<MLAlignment> > liks;
        std::vector<int> onHap;
        std::string gotErr, wantErr;
        try { gpu->computeLikelihoods(haps, reads, liks, 5000u, 5090u, onHap); } catch (std::string &s) { gotErr = s; }
        try { ObservationModelFBMaxErr oms(haps[1], reads[0], 5000u, op); oms.calcLikelihood(); } catch (std::string &s) { wantErr = s; }
        if (gotErr != wantErr || gotErr.empty()) {
            fprintf(stderr, "MISMATCH error behaviour: adapter threw '%s', reference threw '%s'\n", gotErr.c_str(), wantErr.c_str());
            return 1;
        }
    }
    delete gpu;
    printf("adapter_ref_test ok: %ld pairs identical to the reference class (%ld indel events, %ld snps, %ld covered hap indels); "
           "batched submit == window-at-a-time (%ld filtered haplotypes, %ld covering reads); %ld realigned-read CIGARs (%ld operations) "
           "identical to getCIGAR's restatement on the reference's alignments, batched == window-at-a-time\n",
           n_pairs, n_ind, n_snp, n_cov, n_filt, n_covreads, n_cig, n_cig_ops);
    return 0;
}
