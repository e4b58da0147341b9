// tests/cpp/adapter_mock_test.cpp — include/dindel_b200.hpp against test doubles that expose the same
// member names as the reference's Haplotype / Read / MLAlignment / AlignedVariant (no reference code
// needed).  Exit codes: 0 = all checks passed on a CUDA device, 2 = no CUDA device (the adapter has
// no CPU path and must say so), 1 = a check failed.
#include <stdio.h>
#include <string.h>
#include <map>
#include <string>
#include <vector>
#include "dindel_b200.hpp"

struct MockAV { // Variant.hpp:78-175: the members this path touches
    std::string str;
    int startHap, endHap, startRead, endRead;
    MockAV() : startHap(0), endHap(0), startRead(0), endRead(0) {}
    MockAV(const std::string &s, int sh, int eh, int sr, int er) : str(s), startHap(sh), endHap(eh), startRead(sr), endRead(er) {}
    bool isCovered(int pad, int firstBase, int lastBase) const { return firstBase + pad <= startRead && lastBase - pad >= endRead; }
};
// AlignedVariant as the K6 helpers see it (Variant.hpp:41-43, :148, :170): endRead is PROTECTED and has no getter
class MockAV2 {
public:
    enum { REF, INS, DEL, SNP };
    MockAV2() : kind(REF), startRead(-1), endRead(-1) {}
    MockAV2(int k, int sr, int er) : kind(k), startRead(sr), endRead(er) {}
    bool isIndel() const { return kind == INS || kind == DEL; }
    bool isSNP() const { return kind == SNP; }
    bool isRef() const { return kind == REF; }
    int getStartRead() const { return startRead; }
    bool isCovered(int pad, int firstBase, int lastBase) const { return firstBase + pad <= startRead && lastBase - pad >= endRead; }
protected:
    int kind, startRead, endRead;
};
struct MockHap2 { std::string seq; std::map<int, MockAV2> indels, snps; };
struct MockHapML { std::vector<int> hpos; };
struct MockHap { // Haplotype.hpp:40-70, countIndels :254-261, ml = its alignment to the reference segment
    std::string seq;
    std::map<int, MockAV> indels, snps;
    MockHapML ml;
    int countIndels() const { return (int)indels.size(); }
};
struct MockCigar : public std::vector<std::pair<int, int> > { int refPos; }; // DInDel.hpp:168-173
struct MockRead {
    std::string seq;
    std::vector<uint8_t> phred;
    uint8_t mapq;
    uint32_t start;
    bool unmapped, reverse;
};
struct MockMLA { // MLAlignment.hpp:28-76
    int relPos, firstBase, lastBase, hl, hr, numIndels, numMismatch, nBQT, nmmBQT, nMMLeft, nMMRight;
    std::map<int, MockAV> indels, snps;
    std::map<int, bool> hapIndelCovered, hapSNPCovered;
    double ll, llOn, llOff, mLogBQ;
    bool offHap, offHapHMQ;
    std::string align;
    std::vector<int> hpos;
    MockMLA() : relPos(-1), firstBase(0), lastBase(0), hl(-1), hr(-1), numIndels(0), numMismatch(0), nBQT(0), nmmBQT(0),
                nMMLeft(0), nMMRight(0), ll(0), llOn(0), llOff(0), mLogBQ(0), offHap(false), offHapHMQ(false) {}
};
namespace dindel_b200 {
template <> struct ReadTraits<MockRead> {
    static const std::string &seq(const MockRead &r) { return r.seq; }
    static const uint8_t *phred(const MockRead &r) { return r.phred.empty() ? (const uint8_t *)"" : &r.phred[0]; }
    static uint8_t mapq(const MockRead &r) { return r.mapq; }
    static bool unmapped(const MockRead &r) { return r.unmapped; }
    static uint32_t start(const MockRead &r) { return r.start; }
    static bool reverse(const MockRead &r) { return r.reverse; }
    static int pool(const MockRead &) { return 0; }
    static bool coverReverse(const MockRead &r) { return r.reverse; }
    static int insertMode(const MockRead &) { return DINDEL_INS_NONE; }
    static int32_t insertAnchor(const MockRead &, int) { return 0; }
    struct NoLib { int getMaxInsertSize() const { return 1; } double getProb(int) const { return 1.0; } double getNinetyFifthPctProb() const { return 1.0; } };
    static const NoLib *library(const MockRead &) { return 0; }
};
}
struct Params { // ObservationModel.hpp:28-99, CLI values (DInDel.cpp:4122-4157)
    double pError, pMut, mapQualThreshold, pFirstgLO, checkBaseQualThreshold;
    int maxLengthDel, padCover, bMid;
    bool forceReadOnHaplotype, mapUnmappedReads;
};

static MockRead mk(const std::string &s, uint32_t start)
{
    MockRead r;
    r.seq = s; r.phred.assign(s.size(), 30); r.mapq = 60; r.start = start; r.unmapped = false; r.reverse = false;
    return r;
}
#define CHECK(c) do { if (!(c)) { fprintf(stderr, "adapter_mock_test: check failed at line %d: %s\n", __LINE__, #c); return 1; } } while (0)

// a K3 provider that refuses the window the way the library does for more than 64 haplotypes / another error
struct RefusingEM {
    std::string why;
    void windowEM(const std::vector<uint8_t> &, size_t, double, double, std::vector<double> &, std::vector<int32_t> &,
                  std::vector<double> &, std::vector<double> &) { throw why; }
};

// a K6 provider that only counts its calls
struct FakeStats {
    bool resident; int calls;
    bool statsResident() const { return resident; }
    void windowStats(const std::vector<dindel_b200::StatsJob> &jobs, std::vector<dindel_read_stats> &out) { calls++; out.resize(jobs.size()); }
};

int main()
{
    { // host logic only: a window the EM kernel is not built for goes to the caller's own loop, other errors propagate
        dindel_b200::Inference inf;
        inf.enable(true);
        CHECK(inf.enter() && inf.gpu());
        RefusingEM r;
        r.why = "dindel_b200: batch shape not supported by the compiled kernels (more than 64 haplotypes in a window)";
        std::vector<uint8_t> comp(2 * 70, 1);
        inf.emBegin(r, comp, 2, 0.001, 1e-4);
        std::vector<int> c70(70, 1);
        std::vector<double> pi(70, 0.0);
        double ll = 0.0; int it = 0;
        CHECK(inf.risky() && inf.nEmUnsupported == 1 && !inf.em(0, c70, 1e-4, ll, pi, it) && inf.risky());
        inf.leave();
        CHECK(inf.enter());
        r.why = "dindel_b200: CUDA runtime error (an illegal memory access was encountered)";
        bool thrown = false;
        try { inf.emBegin(r, comp, 2, 0.001, 1e-4); } catch (std::string &) { thrown = true; }
        CHECK(thrown && inf.nEmUnsupported == 1);
        inf.leave();
    }
    { // host logic only: the K6 policy (by size unless $DINDEL_B200_DEVICE_STATS says otherwise; never when not resident)
        FakeStats fs = {true, 0};
        std::vector<dindel_b200::StatsJob> jobs(1, dindel_b200::StatsJob(DINDEL_STATS_SITE, -1));
        jobs[0].cands.resize(2);
        std::vector<dindel_read_stats> out;
        const int pol = dindel_b200::statsPolicy();
        const bool small = dindel_b200::deviceStats(fs, jobs, 50, out);            // 100 records
        const bool large = dindel_b200::deviceStats(fs, jobs, 5000, out);          // 10 000 records
        CHECK(small == (pol > 0) && large == (pol != 0) && fs.calls == (small ? 1 : 0) + (large ? 1 : 0));
        fs.resident = false;
        CHECK(!dindel_b200::deviceStats(fs, jobs, 5000, out));
        std::vector<dindel_b200::StatsJob> none;
        fs.resident = true;
        CHECK(!dindel_b200::deviceStats(fs, none, 5000, out));
    }
    Params p = {5e-4, 1e-5, 100.0, 0.01, 0.95, 5, 2, -1, false, false};
    dindel_b200::Engine *eng = 0;
    try {
        eng = new dindel_b200::Engine(p, 0);
    } catch (std::string &s) {
        printf("no engine: %s\n", s.c_str());
        return strstr(s.c_str(), "no usable CUDA device") ? 2 : 1;
    }
    const std::string ref = "ACGTTGCAAGGCTTAGCATCGGATACCGTTAGGCATCAAGTTCGGATCCATGCAAGTCGGCTAAGCTTGCAGGTCCATAGGCTAACGGTTACGATCGGA";
    const uint32_t leftPos = 7000;
    std::vector<MockHap> haps(2);
    haps[0].seq = ref;
    haps[1].seq = ref.substr(0, 50) + ref.substr(53); // 3-bp deletion after hap base 49
    haps[1].indels[50] = MockAV("-" + ref.substr(50, 3), 50, 52, 50, 50);
    std::vector<MockRead> reads;
    reads.push_back(mk(ref.substr(10, 60), leftPos + 10));                          // 0: plain reference read
    reads.push_back(mk(ref.substr(20, 30) + ref.substr(53, 30), leftPos + 20));     // 1: carries the deletion
    reads.push_back(mk(ref.substr(15, 25) + "TT" + ref.substr(40, 35), leftPos + 15)); // 2: 2-bp insertion before hap base 40
    std::string snp = ref.substr(30, 50);
    snp[20] = (snp[20] == 'A') ? 'C' : 'A';
    reads.push_back(mk(snp, leftPos + 30));                                         // 3: one mismatch at hap base 50
    std::vector<std::vector<MockMLA> > liks;
    std::vector<int> onHap;
    eng->computeLikelihoods(haps, reads, liks, leftPos, leftPos + 100, onHap);
    CHECK(liks.size() == 2 && liks[0].size() == 4 && onHap.size() == 4);
    for (int r = 0; r < 4; r++) CHECK(onHap[r] == 1);
    { // read 0 on the reference haplotype: ungapped, no variants
        const MockMLA &m = liks[0][0];
        CHECK(m.ll < 0 && m.ll > -1.0 && m.numIndels == 0 && m.indels.empty() && m.snps.empty());
        CHECK(m.firstBase == 10 && m.lastBase == 69 && m.hpos.size() == 60);
        for (int b = 0; b < 60; b++) CHECK(m.hpos[b] == 10 + b);
        CHECK(m.align == std::string(ref.size(), 'R') && !m.offHap && !m.offHapHMQ && m.nBQT == 60);
    }
    { // read 1: deletion against hap 0, clean against hap 1, and it covers hap 1's indel
        const MockMLA &a = liks[0][1], &b = liks[1][1];
        CHECK(a.numIndels == 1 && a.indels.size() == 1 && a.indels.begin()->second.str[0] == '-');
        CHECK(a.indels.begin()->second.str.size() == 4 && a.indels.begin()->second.endHap - a.indels.begin()->second.startHap == 2);
        CHECK(a.align.find("DDD") != std::string::npos);
        CHECK(b.numIndels == 0 && b.indels.empty() && b.ll > a.ll);
        CHECK(b.hapIndelCovered.size() == 1 && b.hapIndelCovered.begin()->first == 50 && b.hapIndelCovered.begin()->second);
        CHECK(liks[1][0].hapIndelCovered.size() == 1); // every read reports every hap indel, covered or not
    }
    { // read 2: insertion "+TT" recorded before hap base 40 with its read coordinates
        const MockMLA &m = liks[0][2];
        CHECK(m.numIndels == 1 && m.indels.size() == 1);
        const MockAV &v = m.indels.begin()->second;
        CHECK(v.str == "+TT" && v.startHap == v.endHap && v.endRead - v.startRead == 1);
        CHECK(m.hpos[v.startRead] == DINDEL_HPOS_INS && m.hpos[v.endRead] == DINDEL_HPOS_INS);
        CHECK(m.hpos[v.startRead - 1] + 1 == m.indels.begin()->first && m.hpos[v.endRead + 1] == m.indels.begin()->first);
    }
    { // read 3: one SNP at hap base 50, "X=>Y" with X the haplotype's base
        const MockMLA &m = liks[0][3];
        CHECK(m.snps.size() == 1 && m.snps.begin()->first == 50 && m.numMismatch == 1 && m.nmmBQT == 1);
        CHECK(m.snps.begin()->second.str[0] == ref[50] && m.snps.begin()->second.str.substr(1, 2) == "=>");
        CHECK(m.align[50] == snp[20]);
    }
    { // realigned-read CIGARs of the resident window (K5): diploid pair (0, 1), then the pooled choice
        std::vector<MockCigar> cig;
        std::string err;
        try { eng->realignedCigars(haps, reads.size(), 7000, 0, 1, cig); } catch (std::string &s) { err = s; }
        CHECK(err == "Haplotype has not been aligned!");
        for (int b = 0; b < (int)haps[0].seq.size(); b++) haps[0].ml.hpos.push_back(b);
        for (int b = 0; b < (int)haps[1].seq.size(); b++) haps[1].ml.hpos.push_back(b < 50 ? b : b + 3);
        for (int pooled = 0; pooled < 2; pooled++) {
            eng->realignedCigars(haps, reads.size(), 7000, pooled ? -1 : 0, pooled ? -1 : 1, cig);
            CHECK(cig.size() == 4);
            CHECK(cig[0].size() == 1 && cig[0][0] == std::make_pair(0, 60) && cig[0].refPos == 7010);
            CHECK(cig[1].size() == 3 && cig[1][0] == std::make_pair(0, 30) && cig[1][1] == std::make_pair(2, 3) &&
                  cig[1][2] == std::make_pair(0, 30) && cig[1].refPos == 7020);
            CHECK(cig[2].size() == 3 && cig[2][0].first == 0 && cig[2][1] == std::make_pair(1, 2) && cig[2][2].first == 0 &&
                  cig[2][0].second + cig[2][2].second == 60 && cig[2].refPos == 7015);
            CHECK(cig[3].size() == 1 && cig[3][0] == std::make_pair(0, 50) && cig[3].refPos == 7030);
        }
    }
    { // K6 through the adapter: candidates from the variant maps (protected endRead through VariantPeek), the site's
      // read statistics from the GPU against the reference's loop (DInDel.cpp:3504-3554) run here on the mock records
        eng->computeLikelihoods(haps, reads, liks, leftPos, leftPos + 100, onHap);
        std::vector<MockHap2> h2(2);
        h2[0].indels[50] = MockAV2(MockAV2::REF, -1, -1);
        h2[1].indels[50] = MockAV2(MockAV2::DEL, 50, 50);
        h2[0].snps[20] = MockAV2(MockAV2::SNP, 20, 20);
        const dindel_b200::StatsCand c0 = dindel_b200::siteCand(h2[0], 0, 50), c1 = dindel_b200::siteCand(h2[1], 1, 50);
        CHECK(c0.flags == 0 && c1.flags == DINDEL_CAND_PRESENT && c1.start == 50 && c1.end == 50 && c1.hap == 1);
        CHECK(dindel_b200::siteCand(h2[0], 0, 77).flags == 0);                       // no entry at the site
        const dindel_b200::StatsCand m1 = dindel_b200::mapCand(h2[1], 1, 50, true);
        CHECK(m1.flags == DINDEL_CAND_PRESENT && dindel_b200::mapCand(h2[1], 1, 50, false).flags == 0 && dindel_b200::mapCand(h2[0], 0, 50, true).flags == 0);
        const dindel_b200::StatsCand p0 = dindel_b200::poolCand(h2[0], 0, 20, MockAV2(MockAV2::SNP, 0, 0), true);
        CHECK(p0.flags == (DINDEL_CAND_PRESENT | DINDEL_CAND_HAS_VAR) && p0.start == 20 && p0.end == 20);
        CHECK(dindel_b200::poolCand(h2[1], 1, 20, MockAV2(MockAV2::SNP, 0, 0), false).flags == 0);
        std::vector<dindel_b200::StatsJob> jobs(1, dindel_b200::StatsJob(DINDEL_STATS_SITE, -1));
        jobs[0].cands.push_back(c0); jobs[0].cands.push_back(c1);
        std::vector<dindel_read_stats> st;
        eng->windowStats(jobs, st);
        CHECK(st.size() == 1);
        int nf = 0, n = 0, nBQT = 0, nInd = 0;
        double mLogBQ = 0.0;
        for (size_t r = 0; r < reads.size(); r++) {
            const int h = (liks[0][r].ll >= liks[1][r].ll) ? 0 : 1;
            const MockMLA &m = liks[h][r];
            nBQT += m.nBQT; nInd += m.numIndels; mLogBQ += m.mLogBQ;
            if (h == 1 && m.firstBase + p.padCover <= 50 && m.lastBase - p.padCover >= 50) { nf++; n++; }
        }
        CHECK(st[0].n_reads == 4 && st[0].nBQT == nBQT && st[0].numIndels == nInd && st[0].mLogBQ == mLogBQ);
        CHECK(st[0].n == n && st[0].nf == nf && st[0].nr == 0 && n >= 1 && st[0].numUnmappedRealigned == 0);
        CHECK(dindel_b200::deviceStats(*eng, jobs, reads.size(), st) == (dindel_b200::statsPolicy() > 0)); // 4 reads: below the size threshold
    }
    { // a haplotype shorter than maxLengthDel: "hapSize error." (ObservationModelFB.cpp:47)
        std::vector<MockHap> bad(2);
        bad[0].seq = ref; bad[1].seq = "ACG";
        std::string err;
        try { eng->computeLikelihoods(bad, reads, liks, leftPos, leftPos + 100, onHap); } catch (std::string &s) { err = s; }
        CHECK(err == "hapSize error.");
    }
    { // empty window
        std::vector<MockHap> none;
        eng->computeLikelihoods(none, reads, liks, leftPos, leftPos + 100, onHap);
        CHECK(liks.empty() && onHap.size() == 4);
    }
    delete eng;
    printf("adapter_mock_test ok\n");
    return 0;
}
