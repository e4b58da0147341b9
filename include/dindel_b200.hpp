// include/dindel_b200.hpp — C++ host adapter over the C ABI (include/dindel_b200.h).
//
// Source-level drop-in for the one loop of the reference that drives the likelihood engine,
//   DetInDel::computeLikelihoods(haps, reads, liks, leftPos, rightPos, onHap)   (DInDel.cpp:1707-1739)
// i.e. for every (hap, read): ObservationModelFBMaxErr(hap, read, leftPos, obsParams).calcLikelihood()
// (ObservationModelFB.hpp:160,133).  Header only, duck typed: it is written against the *names* the
// reference's value types expose, so it compiles unchanged against the reference's own
// Haplotype / Read / MLAlignment / AlignedVariant (Haplotype.hpp:40, Read.hpp:31,
// MLAlignment.hpp:28, Variant.hpp:78) and against any test double with the same members:
//   HapT : std::string seq;  std::map<int, AV> indels, snps;  (AV::isCovered(pad, first, last))
//   MLA  : ll llOn llOff offHap offHapHMQ firstBase lastBase numIndels numMismatch nBQT nmmBQT
//          mLogBQ nMMLeft nMMRight align hpos indels snps hapIndelCovered hapSNPCovered
//          (mapped_type of MLA::indels constructible from (string, hapStart, hapEnd, readStart, readEnd))
//   ReadT: read through ReadTraits<ReadT> (specialise it for other read types); the default reads the
//          reference's Read: seq.seq, getBam() -> phred bytes / mapping quality / BAM_FUNMAP,
//          posStat.first, onReverseStrand.
// The GPU computes the Viterbi path and the per-pair record; this header only re-expands that record
// into the reference's containers (reportVariants, ObservationModelFB.cpp:1351-1463) and reproduces
// computeLikelihoods' guards.  There is no CPU fallback: without a CUDA device Engine() throws.
#ifndef DINDEL_B200_HPP
#define DINDEL_B200_HPP
#include <stdint.h>
#include <stdlib.h>
#include <iostream>
#include <sstream>
#include <math.h>
#include <map>
#include <string>
#include <time.h>
#include <cstdlib>
#include <cstdio>
#include <vector>
#include "dindel_b200.h"

namespace dindel_b200 {

// How the adapter reads one read.  Default = the reference's Read (Read.hpp); needs samtools' bam.h
// (bam1_qual, BAM_FUNMAP) to be included before this header, as DInDel.cpp does.
template <class ReadT>
struct ReadTraits {
    static const std::string &seq(const ReadT &r) { return r.seq.seq; }
#ifdef BAM_BAM_H
    static const uint8_t *phred(const ReadT &r) { return (const uint8_t *)bam1_qual(r.getBam()); } // Read.hpp:143
    static uint8_t mapq(const ReadT &r) { return (uint8_t)r.getBam()->core.qual; }                   // Read.hpp:127
    static bool unmapped(const ReadT &r) { return (r.getBam()->core.flag & BAM_FUNMAP) != 0; }       // Read.hpp:200
#endif
    static uint32_t start(const ReadT &r) { return uint32_t(r.posStat.first); }                      // ObservationModelFB.cpp:52
    static bool reverse(const ReadT &r) { return r.onReverseStrand; }
    static int pool(const ReadT &r) { return r.poolID; }                                             // Read.hpp:73, DInDel.cpp:2580
#ifdef BAM_BAM_H
    // insert-size prior of computeBMidPrior (ObservationModelFB.cpp:279-293): which formula applies to this read
    // (the caller ANDs it with params.mapUnmappedReads), its anchor coordinate, and its Library (Read.hpp:258)
    static int insertMode(const ReadT &r)
    {
        const bam1_t *b = r.getBam();
        if (!(b->core.flag & BAM_FPAIRED) || (b->core.flag & BAM_FMUNMAP) || r.mateLen == -1 || b->core.tid != b->core.mtid)
            return DINDEL_INS_NONE;
        return (b->core.flag & BAM_FMREVERSE) ? DINDEL_INS_MATE_REVERSE : DINDEL_INS_MATE_FORWARD;
    }
    static int32_t insertAnchor(const ReadT &r, int mode)
    {
        return mode == DINDEL_INS_MATE_REVERSE ? int(r.matePos + r.mateLen) : int(r.matePos);
    }
    static __typeof__(&((const ReadT *)0)->getLibrary()) library(const ReadT &r) { return &r.getLibrary(); }
    // strand a covering read is counted on (DInDel.cpp:1987-1992): unmapped ? !mateIsReverse() : isReverse()
    static bool coverReverse(const ReadT &r)
    {
        const unsigned f = r.getBam()->core.flag;
        return (f & BAM_FUNMAP) ? !(f & BAM_FMREVERSE) : (f & BAM_FREVERSE) != 0;
    }
#endif
};

namespace detail {
// dindel_b200_cigars with 6 operation slots per read (36 bytes per read back over PCIe; nearly every realigned read is
// M, S M, M S, M I M or M D M), run again with the exact capacity in the rare batch where a read needs more.
inline int run_cigars(dindel_ctx *ctx, dindel_cigar_in &in, size_t nr, std::vector<int32_t> &rh, std::vector<int32_t> &rp,
                      std::vector<int32_t> &rn, std::vector<uint32_t> &ops)
{
    int cap = 6;
    if (getenv("DINDEL_B200_CIGAR_CAP") && atoi(getenv("DINDEL_B200_CIGAR_CAP")) > 0) cap = atoi(getenv("DINDEL_B200_CIGAR_CAP")); // tests: force the second pass
    rh.resize(nr); rp.resize(nr); rn.resize(nr);
    for (int pass = 0; pass < 2; pass++) {
        ops.assign(nr * (size_t)cap, 0u);
        in.cigar_cap = cap;
        dindel_cigar_out out;
        out.read_hap = &rh[0]; out.read_ref_pos = &rp[0]; out.read_n_ops = &rn[0]; out.ops = &ops[0];
        int rc = dindel_b200_cigars(ctx, &in, &out);
        if (rc != DINDEL_OK)
            throw std::string("dindel_b200: ") + dindel_b200_strerror(rc) + " (" + dindel_b200_last_error(ctx) + ")";
        int need = 0;
        for (size_t r = 0; r < nr; r++) if (rn[r] > need) need = rn[r];
        if (need <= cap) break;
        cap = need;
    }
    return cap;
}
} // namespace detail

// ---------------------------------------------------------------------------------------------------------------
// K6 (dindel_b200_read_stats): the per-read aggregation loops behind a called site (DInDel.cpp:3126-3170, :3396-3404,
// :3504-3554, :2697-2742) on the GPU, over the records K1 left there.  Bit-identical to the loops (integer counts and
// read-order FP64 sums), so no guard band is involved.  A JOB is one execution of one of the loops; its candidates
// carry what the loop's map lookups would find at the site for each haplotype it looks at.
// ---------------------------------------------------------------------------------------------------------------
struct StatsCand {
    int32_t hap, start, end; uint8_t flags;
    StatsCand() : hap(0), start(0), end(0), flags(0) {}
    StatsCand(int h, int s, int e, unsigned f) : hap(h), start(s), end(e), flags((uint8_t)f) {}
};
struct StatsJob {
    int mode, group; std::vector<StatsCand> cands;
    StatsJob(int m, int g) : mode(m), group(g) {}
};
// AlignedVariant::endRead is protected and has no getter (Variant.hpp:148, :170): read through a pointer to member
// formed in a derived class (the one access the language grants), not through a cast.
template <class AV> struct VariantPeek : AV {
    static int endOf(const AV &a) { int AV::*p = &VariantPeek::endRead; return a.*p; }
};
template <class HapT> struct HapT_variant; // mapped type of HapT::indels / snps (defined below)
// what `m.find(pos)` + isCovered() of the reference reads: present, startRead, endRead of the map's variant at pos
template <class Map>
inline StatsCand statsLookup(const Map &m, int pos, int hap, unsigned extraFlags)
{
    typename Map::const_iterator it = m.find(pos);
    if (it == m.end()) return StatsCand(hap, 0, 0, extraFlags);
    return StatsCand(hap, it->second.getStartRead(), VariantPeek<typename Map::mapped_type>::endOf(it->second), DINDEL_CAND_PRESENT | extraFlags);
}
// DInDel.cpp:3536-3543: the map consulted for haplotype h at a variant position depends on what h carries there
template <class HapT>
inline StatsCand siteCand(const HapT &hap, int h, int pos)
{
    typename std::map<int, typename HapT_variant<HapT>::type>::const_iterator hit = hap.indels.find(pos);
    if (hit == hap.indels.end()) return StatsCand(h, 0, 0, 0);            // (the reference dereferences end() here)
    if (hit->second.isIndel()) return statsLookup(hap.indels, pos, h, 0); // hapIndelCovered[pos], ObservationModelFB.cpp:1465-1468
    if (hit->second.isSNP()) return statsLookup(hap.snps, pos, h, 0);     // hapSNPCovered[pos], :1469-1472
    return StatsCand(h, 0, 0, 0);
}
// DInDel.cpp:3152-3160: haplotype h is looped over (i < m) and has an indel at the site
template <class HapT>
inline StatsCand mapCand(const HapT &hap, int h, int pos, bool looped)
{
    typename std::map<int, typename HapT_variant<HapT>::type>::const_iterator it = hap.indels.find(pos);
    if (!looped || it == hap.indels.end() || !it->second.isIndel()) return StatsCand(h, 0, 0, 0);
    return statsLookup(hap.indels, pos, h, 0);
}
// DInDel.cpp:2716-2724: the variant's own kind picks the map; hapHasVar gates the strand flags (:2726)
template <class HapT, class AV>
inline StatsCand poolCand(const HapT &hap, int h, int pos, const AV &var, bool hasVar)
{
    const unsigned f = hasVar ? DINDEL_CAND_HAS_VAR : 0u;
    if (var.isIndel()) return statsLookup(hap.indels, pos, h, f);
    if (var.isSNP()) return statsLookup(hap.snps, pos, h, f);
    return StatsCand(h, 0, 0, f);
}
// $DINDEL_B200_DEVICE_STATS: 0 = the reference's host loops always, 1 = K6 always, unset = K6 when the loops would
// touch at least DINDEL_B200_STATS_MIN_WORK (read, candidate) records (a K6 call costs ~50 us: a loss on 50-read
// diploid windows whose records are on the host anyway, a gain on deep pooled windows)
#ifndef DINDEL_B200_STATS_MIN_WORK
#define DINDEL_B200_STATS_MIN_WORK 8192
#endif
inline int statsPolicy()
{
    static int p = -2;
    if (p == -2) { const char *e = getenv("DINDEL_B200_DEVICE_STATS"); p = e ? atoi(e) : -1; }
    return p;
}
namespace detail {
inline void runStats(dindel_ctx *ctx, int w, const std::vector<StatsJob> &jobs, const int32_t *readGroup, std::vector<dindel_read_stats> &out)
{
    const size_t nj = jobs.size();
    std::vector<int32_t> jw(nj, (int32_t)w), jm(nj), jg(nj), ch, cs, ce;
    std::vector<int64_t> jc(nj + 1, 0);
    std::vector<uint8_t> cf;
    for (size_t j = 0; j < nj; j++) {
        jm[j] = jobs[j].mode; jg[j] = readGroup ? jobs[j].group : -1;
        for (size_t c = 0; c < jobs[j].cands.size(); c++) {
            const StatsCand &k = jobs[j].cands[c];
            ch.push_back(k.hap); cs.push_back(k.start); ce.push_back(k.end); cf.push_back(k.flags);
        }
        jc[j + 1] = (int64_t)ch.size();
    }
    static const int32_t z32 = 0; static const uint8_t z8 = 0;
    dindel_stats_in in;
    in.n_jobs = (int32_t)nj; in.reserved = 0; in.job_window = &jw[0]; in.job_mode = &jm[0]; in.job_group = &jg[0]; in.job_cand = &jc[0];
    in.cand_hap = ch.empty() ? &z32 : &ch[0]; in.cand_start = cs.empty() ? &z32 : &cs[0]; in.cand_end = ce.empty() ? &z32 : &ce[0];
    in.cand_flags = cf.empty() ? &z8 : &cf[0]; in.read_group = readGroup;
    out.resize(nj);
    int rc = dindel_b200_read_stats(ctx, &in, &out[0]);
    if (rc != DINDEL_OK) throw std::string("dindel_b200: ") + dindel_b200_strerror(rc) + " (" + dindel_b200_last_error(ctx) + ")";
}
} // namespace detail
// Provider: Engine (the resident window), WindowBatch through the batched driver's replay slot (one window of the
// flushed batch).  false = the caller's own loops run.
template <class Provider>
inline bool deviceStats(Provider &p, const std::vector<StatsJob> &jobs, size_t nReads, std::vector<dindel_read_stats> &out)
{
    if (jobs.empty() || !p.statsResident()) return false;
    const int pol = statsPolicy();
    if (pol == 0) return false;
    if (pol < 0) {
        size_t work = 0;
        for (size_t j = 0; j < jobs.size(); j++) work += nReads * jobs[j].cands.size();
        if (work < (size_t)DINDEL_B200_STATS_MIN_WORK) return false;
    }
    p.windowStats(jobs, out);
    return true;
}

// RAII owner of one dindel_ctx (one device).  Construct from anything with the fields of
// ObservationModelParameters (ObservationModel.hpp:28-99) that the path reads.
class Engine {
public:
    template <class ObsParams>
    explicit Engine(const ObsParams &op, int device = 0) : nStatsCalls(0), nStatsJobs(0), ctx_(0), last_nh_(0), last_nr_(0), t_cl_(0.0), t_fh_(0.0), n_cl_(0), mapUnmapped_(false), libOff_(1, 0), wantSums_(false), nPools_(0), multi_(0)
    {
        dindel_b200_default_params(&p_);
        p_.pError = op.pError; p_.pMut = op.pMut; p_.mapQualThreshold = op.mapQualThreshold;
        p_.pFirstgLO = op.pFirstgLO; p_.checkBaseQualThreshold = op.checkBaseQualThreshold;
        p_.maxLengthDel = op.maxLengthDel; p_.padCover = op.padCover; p_.bMid = op.bMid;
        p_.forceReadOnHaplotype = op.forceReadOnHaplotype ? 1 : 0;
        mapUnmapped_ = op.mapUnmappedReads ? true : false; // set by --libFile (DInDel.cpp:4268-4272)
        int rc = dindel_b200_create(&p_, device, &ctx_);
        if (rc != DINDEL_OK) throw std::string("dindel_b200: ") + dindel_b200_strerror(rc);
    }
    ~Engine() { if (multi_) dindel_b200_multi_destroy(multi_); if (ctx_) dindel_b200_destroy(ctx_); }
    // The window scheduler over several GPUs (dindel_b200_multi_*): WindowBatch::flush then shards every batch over
    // these devices; the per-window entries (computeLikelihoods, windowEM, ...) stay on the engine's own device.
    void enableMulti(const std::vector<int> &devices)
    {
        if (multi_ || devices.size() < 2) return;
        int rc = dindel_b200_multi_create(&p_, &devices[0], (int)devices.size(), &multi_);
        if (rc != DINDEL_OK) { multi_ = 0; throw std::string("dindel_b200: ") + dindel_b200_strerror(rc); }
        if (!libP95_.empty()) dindel_b200_multi_set_libraries(multi_, (int)libP95_.size(), &libOff_[0], &libProbs_[0], &libP95_[0]);
    }
    dindel_multi *multi() const { return multi_; }
    dindel_ctx *ctx() const { return ctx_; }
    const dindel_params &params() const { return p_; }
    // Which read sums computeLikelihoods also brings back for the window (K2 inside the same dindel_b200_process call):
    // NONE; ALL = one sum over all reads per haplotype pair (diploidGLF, DInDel.cpp:3086-3089 / :3369-3372);
    // nPools >= 1 = one sum per pool member (estimateHaplotypeFrequenciesBayesEM, DInDel.cpp:2687-2690, myBams.size()).
    void setInference(bool sums, int nPools) { wantSums_ = sums; nPools_ = nPools; }
    bool wantSums() const { return wantSums_; }
    // the sums of the window computeLikelihoods was last called with: [group][q], q enumerating h1 <= h2 row-major
    const std::vector<double> &windowSums() const { return sums_; }
    const double *windowSumsPtr() const { return sums_.empty() ? 0 : &sums_[0]; }
    int windowPools() const { return nPools_ > 0 ? nPools_ : 0; }
    size_t windowHaps() const { return last_nh_; }
    size_t windowReads() const { return last_nr_; }
    const std::vector<int> &windowPoolSizes() const { return poolSize_; }
    // K3 on the resident window: one EM per row of `compatible` (nSets x nh).  Fills loglik[nSets], iters[nSets],
    // pi[nSets*nh] (log frequencies, DInDel.cpp:2473), margin[nSets] (distance of the stopping rule from flipping).
    void windowEM(const std::vector<uint8_t> &compatible, size_t nSets, double a0, double tol, std::vector<double> &loglik,
                  std::vector<int32_t> &iters, std::vector<double> &pi, std::vector<double> &margin)
    {
        const size_t nh = last_nh_;
        loglik.assign(nSets, 0.0); iters.assign(nSets, 0); pi.assign(nSets * nh, 0.0); margin.assign(nSets, 0.0);
        if (nSets == 0) return;
        std::vector<int32_t> sw(nSets, 0);
        std::vector<int64_t> off(nSets + 1, 0);
        for (size_t k = 0; k <= nSets; k++) off[k] = (int64_t)(k * nh);
        std::vector<double> freqs(nSets * nh, 0.0);
        dindel_em_in in; in.n_sets = (int32_t)nSets; in.reserved = 0; in.set_window = &sw[0]; in.set_off = &off[0];
        in.compatible = &compatible[0]; in.a0 = a0; in.tol = tol;
        dindel_em_out out; out.loglik = &loglik[0]; out.iters = &iters[0]; out.pi = &pi[0]; out.freqs = &freqs[0];
        out.margin = &margin[0]; out.energy = 0;
        int rc = dindel_b200_bayes_em(ctx_, &in, &out);
        if (rc != DINDEL_OK) throw std::string("dindel_b200: ") + dindel_b200_strerror(rc) + " (" + dindel_b200_last_error(ctx_) + ")";
    }
    // K6 on the resident window (the one computeLikelihoods was last called with)
    bool statsResident() const { return last_nh_ > 0 && last_nr_ > 0; }
    void windowStats(const std::vector<StatsJob> &jobs, std::vector<dindel_read_stats> &out)
    {
        detail::runStats(ctx_, 0, jobs, rgroup_.empty() ? 0 : &rgroup_[0], out);
        nStatsCalls++; nStatsJobs += (long)jobs.size();
    }
    long nStatsCalls, nStatsJobs;
    bool mapUnmappedReads() const { return mapUnmapped_; }
    void setMapUnmappedReads(bool v) { mapUnmapped_ = v; } // the reference flips it after construction (DInDel.cpp:4270)

    // Index of an insert-size library (Library.hpp:31-129) among those uploaded to the GPU; a library seen for the
    // first time is read out through its public interface and all of them are re-sent (dindel_b200_set_libraries).
    template <class LibT>
    int libraryIndex(const LibT *lib)
    {
        std::map<const void *, int>::const_iterator it = libIndex_.find((const void *)lib);
        if (it != libIndex_.end()) return it->second;
        const int idx = (int)libP95_.size(), maxins = lib->getMaxInsertSize();
        for (int d = 0; d < maxins; d++) libProbs_.push_back(lib->getProb(d));
        libOff_.push_back((int64_t)libProbs_.size());
        libP95_.push_back(lib->getNinetyFifthPctProb());
        libIndex_[(const void *)lib] = idx;
        int rc = dindel_b200_set_libraries(ctx_, (int)libP95_.size(), &libOff_[0], &libProbs_[0], &libP95_[0]);
        if (rc == DINDEL_OK && multi_) rc = dindel_b200_multi_set_libraries(multi_, (int)libP95_.size(), &libOff_[0], &libProbs_[0], &libP95_[0]);
        if (rc != DINDEL_OK) throw std::string("dindel_b200: ") + dindel_b200_strerror(rc) + " (" + dindel_b200_last_error(ctx_) + ")";
        return idx;
    }

    template <class HapT, class ReadT, class MLA>
    void computeLikelihoods(const std::vector<HapT> &haps, const std::vector<ReadT> &reads,
                            std::vector<std::vector<MLA> > &liks, uint32_t leftPos, uint32_t rightPos,
                            std::vector<int> &onHap);

    // Drop-in for DetInDel::filterHaplotypes(haps, reads, liks, filtered, varCoverage, doFilter)
    // (DInDel.cpp:1932-2100) on the window computeLikelihoods was last called with: the per-base hpos rows
    // are tested on the GPU (dindel_b200_coverage) where K1 left them; only filtered[] and the (nf, nr)
    // counts come back.  CovMap = map<pair<int, AlignedVariant>, VariantCoverage> (any mapped type
    // constructible from (nf, nr)).  Returns the number of filtered haplotypes (the reference prints it).
    template <class HapT, class ReadT, class MLA, class CovMap>
    int filterHaplotypes(const std::vector<HapT> &haps, const std::vector<ReadT> &reads,
                         const std::vector<std::vector<MLA> > &liks, std::vector<int> &filtered,
                         CovMap &varCoverage, bool doFilter, int maxMismatch);

    // Drop-in for the per-read loops around DetInDel::getCIGAR of --outputRealignedBAM (DInDel.cpp:500-513 pooled:
    // pass h1 < 0; :596-611 diploid: pass the MAP pair hpl.h1 / hpl.h2 of getMaxHap) on the window
    // computeLikelihoods was last called with: the haplotype choice and getCIGAR (DInDel.cpp:728-882) run on the
    // GPU over the resident ll / hpos rows (K5, dindel_b200_cigars); CigarT = DetInDel::CIGAR (a
    // vector<pair<int,int> > of (BAM op, length) with an int refPos, DInDel.hpp:168-173).  Reads the pooled loop
    // skips (!onHap[r]) get an empty CIGAR; refPos of a wholly clipped read (unset by the reference) is -1.
    // The reference's throws ("Error(n)!", "How is this possible? (1)") are thrown as the same strings.
    template <class HapT, class CigarT>
    void realignedCigars(const std::vector<HapT> &haps, size_t nReads, int refSeqPos, int h1, int h2,
                         std::vector<CigarT> &cigars);

private:
    Engine(const Engine &);
    Engine &operator=(const Engine &);
    dindel_params p_;
    dindel_ctx *ctx_;
    size_t last_nh_, last_nr_; // shape of the resident window (filterHaplotypes checks it is the same one)
    // DINDEL_B200_TRACE: wall time inside the adapter's two entries (packing + C call + re-expansion), stderr
    double t_cl_, t_fh_;
    long n_cl_;
    static double now() { timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec; }
    struct Tick {
        double *acc; double t0;
        explicit Tick(double *a) : acc(a), t0(now()) {}
        ~Tick() { *acc += now() - t0; }
    };
    bool mapUnmapped_;
    std::map<const void *, int> libIndex_;
    std::vector<int64_t> libOff_;
    std::vector<double> libProbs_, libP95_;
    bool wantSums_;
    int nPools_;
    std::vector<double> sums_;
    std::vector<int> poolSize_;
    std::vector<int32_t> rgroup_;  // Read::poolID of the resident window's reads
    dindel_multi *multi_;
};

// ---------------------------------------------------------------------------------------------------------------
// Inference: the numeric cores of DetInDel::diploidGLF (DInDel.cpp:3085-3120, :3240-3266, :3349-3390) and of
// DetInDel::estimateHaplotypeFrequenciesBayesEM (:2431-2545, :2683-2694) take their read sums / EM results from the
// GPU (K2 / K3) while the TEXT the reference writes stays byte-identical.  The GPU values differ from glibc's at the
// ulp level (CUDA libm, summation order), so every place where such a value is COMPARED or PRINTED goes through a
// guard: a comparison closer than the error bound, or a printed field whose 6-digit text (OutputData::Line::set
// streams doubles at the default precision, OutputData.hpp:82-84) could change within the bound, marks the window
// "undecided".  The patched function runs under DINDEL_B200_TWO_PASS: output is buffered; an undecided window is run
// again with the reference's own CPU loops (mode CPU: every hook is transparent), so the result is the reference's
// by construction.  Counters report how often that happened.
class Inference {
public:
    enum Mode { IDLE, GPU, CPU };
    Inference() : nWindows(0), nRecheck(0), nEmSets(0), nEmMismatch(0), nEmUnsupported(0), enabled_(false), mode_(IDLE), risky_(false), S_(0),
                  nh_(0), nr_(0), G_(0), gsize_(0), eps_(0.0), emErr_(0.0), emSets_(0), emUnavailable_(false) { for (int i = 0; i < 6; i++) nTrip[i] = 0; }
    void enable(bool on) { enabled_ = on; }
    bool enabled() const { return enabled_; }
    bool enter() { if (!enabled_ || mode_ != IDLE) return false; mode_ = GPU; risky_ = false; eps_ = 0.0; emErr_ = 0.0; nWindows++; return true; }
    void leave() { mode_ = IDLE; }
    bool gpu() const { return mode_ == GPU; }
    bool risky() const { return risky_; }
    void toCpu() { mode_ = CPU; risky_ = false; nRecheck++; }
    void markRisky() { if (mode_ == GPU) risky_ = true; }
    // the window's read sums (K2): S[g * npairs + q]; nGroups == 0: one block over all reads
    void setSums(const double *S, size_t nh, size_t nr, int nGroups, const int *groupSize)
    { S_ = S; nh_ = nh; nr_ = nr; G_ = nGroups; gsize_ = groupSize; }
    static size_t tri(size_t h1, size_t h2, size_t nh) { return h1 * nh - h1 * (h1 - 1) / 2 + (h2 - h1); }
    // start + sum over the window's reads of log(.5) + addLogs(rl[r,h1], rl[r,h2])      (DInDel.cpp:3086-3089, :3369-3372)
    double sum(size_t h1, size_t h2, size_t nh, double start)
    {
        if (!S_ || nh != nh_) { markRisky(); return start; }
        const double s = S_[tri(h1, h2, nh)];
        widen(dindel_b200_sum_error_bound((int64_t)nr_ + 1, fabs(start) + fabs(s) + 1.0));
        return start + s;
    }
    // the same over the reads of pool member b                                          (DInDel.cpp:2687-2690)
    double sumGroup(int b, size_t h1, size_t h2, size_t nh, double start)
    {
        if (!S_ || nh != nh_ || b < 0 || b >= G_) { markRisky(); return start; }
        const double s = S_[(size_t)b * (nh * (nh + 1) / 2) + tri(h1, h2, nh)];
        widen(dindel_b200_sum_error_bound((int64_t)(gsize_ ? gsize_[b] : (int)nr_) + 1, fabs(start) + fabs(s) + 1.0));
        return start + s;
    }
    // a > b as the reference evaluates it; undecided when the two are closer than the bound.  gt: log-domain values
    // (sums, log-likelihoods: absolute bound); gtLin: probabilities / frequencies derived from the EM (relative bound)
    bool gt(double a, double b)
    {
        if (mode_ == GPU && fabs(a - b) <= 2.0 * eps_ + 2.0 * emErr_) trip(0);
        return a > b;
    }
    bool gtLin(double a, double b)
    {
        if (mode_ == GPU && fabs(a - b) <= 4.0 * emErr_ * (fabs(a) > fabs(b) ? fabs(a) : fabs(b))) trip(1);
        return a > b;
    }
    // a value about to be printed: its text must not depend on where inside the bound the true value lies
    double printed(double v)
    {
        if (mode_ == GPU) {
            const double e = 2.0 * eps_ + 2.0 * emErr_;
            if (text(v - e) != text(v + e)) trip(2);
        }
        return v;
    }
    double printedLin(double v)
    {
        if (mode_ == GPU) {
            const double e = 4.0 * emErr_ * fabs(v);
            if (text(v - e) != text(v + e)) trip(3);
        }
        return v;
    }
    // -10 * (ref - addLogs(alt, ref)) / log(10): qual (DInDel.cpp:3119) and genoqual (:3266), whose inputs are two
    // maxima of sums; evaluated at the four corners of the bound (the formula cancels, so a plain interval on its
    // output would not do)
    void checkPhred(double ref, double alt)
    {
        if (mode_ != GPU) return;
        const double e = 2.0 * eps_;
        const std::string t = text(phred(ref, alt));
        for (int i = 0; i < 4; i++)
            if (text(phred(ref + ((i & 1) ? e : -e), alt + ((i & 2) ? e : -e))) != t) { trip(4); return; }
    }
    // ---- Bayes EM (K3) ----
    // All EM runs of the window in one launch; em() then hands them out set by set and checks that the set it was
    // computed for is the one the reference arrived at.
    // Provider: anything with windowEM(compatible, nSets, a0, tol, loglik, iters, pi, margin) for the current window
    // (Engine: the resident window; the batched driver: one window of the flushed batch)
    template <class Provider>
    void emBegin(Provider &e, const std::vector<uint8_t> &compatible, size_t nSets, double a0, double tol)
    {
        emCompat_ = compatible; emSets_ = nSets; emUnavailable_ = false;
        if (mode_ != GPU) return;
        // A window the kernel is not built for (more haplotypes than K3 takes: 256) is not an error of the window: its EM
        // runs in the reference's own loop, like an undecided one.
        try { e.windowEM(compatible, nSets, a0, tol, emLoglik_, emIters_, emPi_, emMargin_); }
        catch (std::string &why) {
            if (why.find("shape not supported") == std::string::npos) throw;
            emUnavailable_ = true; risky_ = true; nEmUnsupported++;
        }
    }
    // returns true (converged) and fills loglik / pi / iter when the GPU's EM of set th stands; else the reference's
    // own while-loop runs (returns false)
    template <class IntVec, class DblVec>
    bool em(int th, const IntVec &compatible, double tol, double &loglik, DblVec &pi, int &iter)
    {
        if (mode_ != GPU) return false;
        if (emUnavailable_) { risky_ = true; return false; }
        nEmSets++;
        const size_t nh = compatible.size();
        if ((size_t)th >= emSets_ || emCompat_.size() != emSets_ * nh) { nEmMismatch++; risky_ = true; return false; }
        for (size_t h = 0; h < nh; h++)
            if ((compatible[h] != 0) != (emCompat_[th * nh + h] != 0)) { nEmMismatch++; risky_ = true; return false; }
        // stopping rule |eOld - eNew| < tol (DInDel.cpp:2515): the GPU's energies differ from the CPU's by ~1e-13
        // relative (measured 1e-15, tools/k23_deviation.py -> profiles/r2_k23_deviation.json); a pass that came closer to
        // tol than 1e-10 * (1 + |loglik|) could have gone the other way on the CPU
        const double guard = emGuardRel() * (1.0 + fabs(emLoglik_[th]));
        if (emMargin_[th] < guard) { trip(5); return false; }
        loglik = emLoglik_[th];
        for (size_t h = 0; h < nh; h++) pi[h] = emPi_[th * nh + h];
        iter = emIters_[th];
        const double e = emValueRel() * (1.0 + fabs(loglik));
        if (e > emErr_) emErr_ = e;
        return true;
    }
    static double emGuardRel() { return 1e-10; }  // margin guard, relative to 1 + |loglik| (measured energy deviation: 1e-15)
    static double emValueRel() { return 1e-12; }  // bound on the GPU EM's error, relative (measured: loglik 7e-16, pi 5e-14, freqs 5e-15)
    static std::string text(double v) { std::ostringstream os; os << v; return os.str(); }
    static double addLogsHost(double l1, double l2) // Utils.hpp:29-38
    {
        if (l1 > l2) { double diff = l2 - l1; return l1 + log(1.0 + exp(diff)); }
        double diff = l1 - l2; return l2 + log(1.0 + exp(diff));
    }
    static double phred(double ref, double alt) { return -10.0 * (ref - addLogsHost(alt, ref)) / log(10.0); }
    long nWindows, nRecheck, nEmSets, nEmMismatch, nEmUnsupported;
    long nTrip[6]; // which guard marked a window undecided: 0 gt, 1 gtLin, 2 printed, 3 printedLin, 4 checkPhred, 5 EM stop-rule margin
private:
    void widen(double e) { if (e > eps_) eps_ = e; }
    void trip(int which) { if (!risky_) nTrip[which]++; risky_ = true; } // counts the FIRST guard that tripped in a window
    bool enabled_;
    Mode mode_;
    bool risky_;
    const double *S_;
    size_t nh_, nr_;
    int G_;
    const int *gsize_;
    double eps_, emErr_;
    std::vector<uint8_t> emCompat_;
    size_t emSets_;
    bool emUnavailable_;
    std::vector<double> emLoglik_, emPi_, emMargin_;
    std::vector<int32_t> emIters_;
};

// Runs CALL (the function's own body, re-entered) with the glf output buffered: first with the GPU's numbers, and
// when a guard tripped once more with the reference's CPU loops.  OUT = the OutputData whose `out` stream the body
// writes to; RESET = statement(s) undoing what a discarded pass appended to output arguments.
struct TwoPass {
    Inference &inf; std::ostream **slot; std::ostream *real; std::ostringstream buf;
    TwoPass(Inference &i, std::ostream **s) : inf(i), slot(s), real(*s) { *slot = &buf; }
    ~TwoPass() { *slot = real; inf.leave(); }
    bool retry() { if (inf.gpu() && inf.risky()) { buf.str(std::string()); buf.clear(); inf.toCpu(); return true; } return false; }
    void commit() { *real << buf.str(); }
};
#define DINDEL_B200_TWO_PASS(INF, OUT, RESET, CALL)                                              \
    if ((INF).enter()) {                                                                         \
        dindel_b200::TwoPass _tp((INF), &(OUT).out);                                             \
        for (;;) {                                                                               \
            std::string _ex; bool _threw = false;                                                \
            try { CALL; } catch (std::string &_s) { _threw = true; _ex = _s; }                   \
            if (_tp.retry()) { RESET; continue; }                                                \
            _tp.commit();                                                                        \
            if (_threw) throw _ex;                                                               \
            return;                                                                              \
        }                                                                                        \
    }

// mapped types of MLA::indels and HapT::indels (C++98-friendly, the reference builds as gnu++98)
template <class M> struct map_value;
template <class K, class V, class C, class A> struct map_value<std::map<K, V, C, A> > { typedef V type; };
template <class MLA> struct MLA_indel_type {
    static MLA &ref();
    typedef typename map_value<__typeof__(ref().indels)>::type type;
};
template <class HapT> struct HapT_variant {
    static HapT &ref();
    typedef typename map_value<__typeof__(ref().indels)>::type type;
};

// Re-expands one GPU pair record into the reference's MLAlignment exactly as reportVariants does
// (ObservationModelFB.cpp:1351-1463).  The Viterbi states are recovered from hpos + mapState0:
// an on-haplotype base at hap index p is state x = p+1, LO is x = 0, RO is x = H+1, and a base inside
// an insertion keeps the column of the base before it (plane 1 is entered from plane 0 of the same
// column, cpp:1754/:1823) or, when the read starts inside the insertion, the column in mapState0.
template <class HapT, class MLA>
inline void expandAlignment(const dindel_pair_out &rec, const int16_t *hpos, const HapT &hap,
                            const std::string &readSeq, int padCover, MLA &ml)
{
    typedef typename MLA_indel_type<MLA>::type AV;
    const std::string &hseq = hap.seq;
    const int H = (int)hseq.size(), S = H + 2, R = (int)readSeq.size();
    ml.ll = rec.ll; ml.llOn = rec.llOn; ml.llOff = rec.llOff;                    // cpp:1120-1135
    ml.offHap = rec.offHap != 0; ml.offHapHMQ = rec.offHapHMQ != 0;
    ml.align = std::string((size_t)H, 'R');                                      // cpp:1353-1372
    ml.indels.clear(); ml.snps.clear();
    ml.hapIndelCovered.clear(); ml.hapSNPCovered.clear();
    ml.firstBase = rec.firstBase; ml.lastBase = rec.lastBase;
    ml.nBQT = rec.nBQT; ml.nmmBQT = rec.nmmBQT; ml.mLogBQ = rec.mLogBQ;
    ml.nMMRight = rec.nMMRight; ml.nMMLeft = rec.nMMLeft;
    ml.numIndels = rec.numIndels; ml.numMismatch = rec.numMismatch;
    ml.hpos.assign((size_t)R, 0);
    for (int b = 0; b < R; b++) ml.hpos[b] = hpos[b];

    // plane-0 column of read base b, or -1 inside an insertion
    struct Col {
        static int of(int hp, int H) { return hp >= 0 ? hp + 1 : (hp == DINDEL_HPOS_LO ? 0 : (hp == DINDEL_HPOS_RO ? H + 1 : -1)); }
    };
    int b = 0;
    while (b < R) {
        const int hp = hpos[b];
        if (hp == DINDEL_HPOS_INS) {                                             // cpp:1377-1396
            int pos;
            if (b > 0) pos = Col::of(hpos[b - 1], H);
            else pos = (int)rec.mapState0 - S;
            const int rpos = b;
            int len = 0;
            while (b < R && hpos[b] == DINDEL_HPOS_INS) { b++; len++; }
            ml.indels[pos] = AV(std::string("+").append(readSeq.substr(rpos, len)), pos, pos, rpos, b - 1);
            continue; // b already sits on the base after the run (the reference does b--, b++)
        }
        if (hp >= 0) {
            const int s = hp + 1;
            if (readSeq[b] != hseq[s - 1]) {                                     // cpp:1410-1433
                std::string snp;
                snp += hseq[s - 1];
                snp.append("=>");
                snp += readSeq[b];
                ml.snps[s - 1] = AV(snp, s - 1, s - 1, b, b);
                ml.align[s - 1] = readSeq[b];
            }
            if (b < R - 1 && hpos[b + 1] != DINDEL_HPOS_INS) {                    // cpp:1436-1453
                const int ns = Col::of(hpos[b + 1], H);
                if (ns - s > 1) {
                    const int pos = s, dlen = ns - s - 1;
                    for (int y = pos; y < dlen + pos; y++) ml.align[y] = 'D';
                    ml.indels[pos] = AV(std::string("-").append(hseq.substr(pos, dlen)), pos, pos + dlen - 1, b, b + 1);
                }
            }
        }
        b++;
    }
    for (typename std::map<int, typename HapT_variant<HapT>::type>::const_iterator it = hap.indels.begin();
         it != hap.indels.end(); ++it)                                           // cpp:1465-1468
        ml.hapIndelCovered[it->first] = it->second.isCovered(padCover, ml.firstBase, ml.lastBase);
    for (typename std::map<int, typename HapT_variant<HapT>::type>::const_iterator it = hap.snps.begin();
         it != hap.snps.end(); ++it)                                             // cpp:1469-1472
        ml.hapSNPCovered[it->first] = it->second.isCovered(padCover, ml.firstBase, ml.lastBase);
}

// DetInDel::computeLikelihoods (DInDel.cpp:1707-1739) for one window: packs the window into the C
// ABI's batch layout, runs the GPU engine, re-expands every pair.  Guards as in the reference, in
// its h-outer / r-inner order: a haplotype shorter than maxLengthDel throws "hapSize error."
// (ObservationModelFB.cpp:47), ll > 0.1 prints "Likelihood>0" and exits (DInDel.cpp:1722-1731),
// NaN / Inf throws "Nan detected" (DInDel.cpp:1732-1735).
template <class HapT, class ReadT, class MLA>
void Engine::computeLikelihoods(const std::vector<HapT> &haps, const std::vector<ReadT> &reads,
                                std::vector<std::vector<MLA> > &liks, uint32_t leftPos, uint32_t /*rightPos*/,
                                std::vector<int> &onHap)
{
    typedef ReadTraits<ReadT> RT;
    Tick tick(&t_cl_);
    if (++n_cl_ % 50 == 0 && getenv("DINDEL_B200_TRACE"))
        fprintf(stderr, "dindel_b200 adapter trace: %ld windows, computeLikelihoods %.3f s, filterHaplotypes %.3f s\n", n_cl_, t_cl_, t_fh_);
    const size_t nh = haps.size(), nr = reads.size();
    onHap = std::vector<int>(nr, 0);
    liks = std::vector<std::vector<MLA> >(nh, std::vector<MLA>(nr));
    last_nh_ = last_nr_ = 0; // nothing of an earlier window may answer for this one
    sums_.clear();
    if (nh == 0 || nr == 0) return;
    std::vector<int64_t> hap_off(1, 0), read_off(1, 0);
    std::string hap_seq, read_seq;
    std::vector<uint8_t> qual, mapq, flags, insMode;
    std::vector<uint32_t> start;
    std::vector<int32_t> insAnchor, insLib;
    bool anyIns = false;
    size_t maxR = 0;
    for (size_t h = 0; h < nh; h++) { hap_seq += haps[h].seq; hap_off.push_back((int64_t)hap_seq.size()); }
    for (size_t r = 0; r < nr; r++) {
        const std::string &sq = RT::seq(reads[r]);
        const uint8_t *ph = RT::phred(reads[r]);
        read_seq += sq; read_off.push_back((int64_t)read_seq.size());
        qual.insert(qual.end(), ph, ph + sq.size());
        mapq.push_back(RT::mapq(reads[r]));
        start.push_back(RT::start(reads[r]));
        flags.push_back((uint8_t)((RT::unmapped(reads[r]) ? DINDEL_READ_UNMAPPED : 0) |
                                  (RT::reverse(reads[r]) ? DINDEL_READ_REVERSE : 0)));
        if (sq.size() > maxR) maxR = sq.size();
        const int mode = mapUnmapped_ ? RT::insertMode(reads[r]) : DINDEL_INS_NONE;                   // cpp:279,283
        insMode.push_back((uint8_t)mode);
        insAnchor.push_back(mode != DINDEL_INS_NONE ? RT::insertAnchor(reads[r], mode) : 0);
        insLib.push_back(mode != DINDEL_INS_NONE ? libraryIndex(RT::library(reads[r])) : 0);
        anyIns = anyIns || mode != DINDEL_INS_NONE;
    }
    int32_t win_hap[2] = {0, (int32_t)nh}, win_read[2] = {0, (int32_t)nr};
    static const uint8_t none = 0;
    dindel_batch_in in;
    in.n_windows = 1; in.n_haps = (int32_t)nh; in.n_reads = (int32_t)nr; in.reserved = 0;
    in.hap_start = &leftPos; in.win_hap = win_hap; in.win_read = win_read;
    in.hap_off = &hap_off[0]; in.hap_seq = hap_seq.data();
    in.read_off = &read_off[0]; in.read_seq = read_seq.data();
    in.read_qual = qual.empty() ? &none : &qual[0];
    in.read_mapq = &mapq[0]; in.read_start = &start[0]; in.read_flags = &flags[0];
    in.read_ins_mode = anyIns ? &insMode[0] : 0; in.read_ins_anchor = anyIns ? &insAnchor[0] : 0; in.read_lib = anyIns ? &insLib[0] : 0;

    const size_t np = nh * nr;
    const int stride = (int)((maxR + 7) & ~(size_t)7) > 0 ? (int)((maxR + 7) & ~(size_t)7) : 8;
    std::vector<dindel_pair_out> pairs(np);
    std::vector<int16_t> hpos(np * (size_t)stride);
    std::vector<uint8_t> on(nr);
    int64_t win_pair[2];
    int32_t status = 0;
    dindel_batch_out out;
    out.n_pairs = (int64_t)np; out.pairs = &pairs[0]; out.hpos = &hpos[0]; out.hpos_stride = stride;
    out.flags = DINDEL_OUT_KEEP_HPOS; out.win_pair = win_pair; out.win_status = &status; out.read_on_hap = &on[0];
    last_nh_ = last_nr_ = 0;
    // one pass: K1 and, when the caller's inference runs on the GPU too, the read sums of the window (K2) while the
    // likelihood matrix is still in HBM
    dindel_job job;
    job.in = &in; job.out = &out; job.cover = 0; job.cover_out = 0; job.pair_sums = 0; job.win_sum_off = 0;
    job.n_groups = 0; job.reserved = 0; job.read_group = 0; job.em = 0; job.em_out = 0;
    std::vector<int32_t> &rgroup = rgroup_;
    rgroup.resize(nr);
    for (size_t r = 0; r < nr; r++) rgroup[r] = RT::pool(reads[r]);
    if (wantSums_) {
        const int G = nPools_ > 0 ? nPools_ : 1;
        sums_.assign(nh * (nh + 1) / 2 * (size_t)G, 0.0);
        job.pair_sums = &sums_[0];
        if (nPools_ > 0) {
            poolSize_.assign((size_t)nPools_, 0);
            for (size_t r = 0; r < nr; r++) {
                if (rgroup[r] >= 0 && rgroup[r] < nPools_) poolSize_[rgroup[r]]++;
            }
            job.n_groups = nPools_; job.read_group = &rgroup[0];
        }
    }
    int rc = dindel_b200_process(ctx_, &job);
    if (rc != DINDEL_OK)
        throw std::string("dindel_b200: ") + dindel_b200_strerror(rc) + " (" + dindel_b200_last_error(ctx_) + ")";
    last_nh_ = nh; last_nr_ = nr;

    for (size_t h = 0; h < nh; h++) {
        for (size_t r = 0; r < nr; r++) {
            const size_t i = r * nh + h; // pairs are r-major inside a window
            const dindel_pair_out &rec = pairs[i];
            if (rec.status == DINDEL_WIN_HAPSIZE) throw std::string("hapSize error.");
            expandAlignment(rec, &hpos[i * (size_t)stride], haps[h], RT::seq(reads[r]), p_.padCover, liks[h][r]);
            if (!liks[h][r].offHapHMQ) onHap[r] = 1;
            if (rec.status == DINDEL_WIN_LL_POSITIVE) {
                std::cout << std::string(25, ' ') << haps[h].seq << std::endl;
                std::cout << "hidx: " << h << " r: " << r << std::endl;
                std::cerr << "Likelihood>0" << std::endl;
                exit(1);
            }
            if (rec.status == DINDEL_WIN_NAN) {
                std::cout << "NAN/Inf error" << std::endl;
                throw std::string("Nan detected");
            }
        }
    }
}


// The throughput binding: the caller's window loop split into prepare -> add() ->
// one flush() per batch -> result() / filterResult() per window.  One dindel_b200_submit (pipelined H2D /
// kernels / D2H) and one dindel_b200_coverage serve every window added since the last clear(); results are
// what Engine::computeLikelihoods / Engine::filterHaplotypes return window by window (same code expands them).
template <class HapT, class ReadT>
class WindowBatch {
    typedef ReadTraits<ReadT> RT;
    typedef typename HapT_variant<HapT>::type AV;
    typedef std::pair<int, AV> PAV;
public:
    // fullAlignments = false: the per-base hpos rows stay on the GPU (DINDEL_OUT_KEEP_HPOS); result() then fills the
    // scalar MLAlignment fields and the hapIndelCovered / hapSNPCovered maps but leaves hpos / align / indels / snps
    // empty (nothing in the calling pipeline reads them once filterResult() replaces filterHaplotypes).
    explicit WindowBatch(Engine &e, int maxMismatch = 2, bool doFilter = false, bool fullAlignments = true)
        : eng_(e), maxMismatch_(maxMismatch), doFilter_(doFilter), flushed_(false), full_(fullAlignments), resident_(false) { clear(); }
    void clear()
    {
        hap_start_.clear(); win_hap_.assign(1, 0); win_read_.assign(1, 0); hap_off_.assign(1, 0); read_off_.assign(1, 0);
        hap_seq_.clear(); read_seq_.clear(); qual_.clear(); mapq_.clear(); flags_.clear(); strand_.clear(); start_.clear();
        win_entry_.assign(1, 0); win_key_.assign(1, 0); e_hap_.clear(); e_key_.clear(); e_left_.clear(); e_right_.clear();
        e_type_.clear(); keys_.clear(); maxR_ = 0; flushed_ = false;
        insMode_.clear(); insAnchor_.clear(); insLib_.clear(); anyIns_ = false; pool_.clear();
    }
    size_t size() const { return hap_start_.size(); }
    // copies what the engine needs of one window (DetInDel::computeLikelihoods' arguments + the variant entries of
    // DetInDel::filterHaplotypes, DInDel.cpp:1962-1974); returns the window's index in the batch
    int add(const std::vector<HapT> &haps, const std::vector<ReadT> &reads, uint32_t leftPos)
    {
        flushed_ = false;
        const int pad = eng_.params().padCover;
        std::map<PAV, int> keyOf;
        std::vector<PAV> keys;
        for (size_t h = 0; h < haps.size(); h++) {
            hap_seq_ += haps[h].seq; hap_off_.push_back((int64_t)hap_seq_.size());
            for (typename std::map<int, AV>::const_iterator it = haps[h].indels.begin(); it != haps[h].indels.end(); ++it) {
                const AV &av = it->second;
                const PAV pav(it->first, av);
                typename std::map<PAV, int>::iterator k = keyOf.find(pav);
                if (k == keyOf.end()) { k = keyOf.insert(std::make_pair(pav, (int)keys.size())).first; keys.push_back(pav); }
                e_hap_.push_back((int32_t)h); e_key_.push_back(k->second);
                e_type_.push_back(av.getType() == AV::DEL ? DINDEL_VAR_DEL : (av.getType() == AV::INS ? DINDEL_VAR_INS : DINDEL_VAR_OTHER));
                e_left_.push_back(av.getLeftFlankRead() - pad); e_right_.push_back(av.getRightFlankRead() + pad);
            }
        }
        for (size_t r = 0; r < reads.size(); r++) {
            const std::string &sq = RT::seq(reads[r]);
            const uint8_t *ph = RT::phred(reads[r]);
            read_seq_ += sq; read_off_.push_back((int64_t)read_seq_.size());
            qual_.insert(qual_.end(), ph, ph + sq.size());
            mapq_.push_back(RT::mapq(reads[r])); start_.push_back(RT::start(reads[r]));
            flags_.push_back((uint8_t)((RT::unmapped(reads[r]) ? DINDEL_READ_UNMAPPED : 0) | (RT::reverse(reads[r]) ? DINDEL_READ_REVERSE : 0)));
            strand_.push_back(RT::coverReverse(reads[r]) ? 1 : 0);
            pool_.push_back(RT::pool(reads[r]));
            if (sq.size() > maxR_) maxR_ = sq.size();
            const int mode = eng_.mapUnmappedReads() ? RT::insertMode(reads[r]) : DINDEL_INS_NONE;
            insMode_.push_back((uint8_t)mode);
            insAnchor_.push_back(mode != DINDEL_INS_NONE ? RT::insertAnchor(reads[r], mode) : 0);
            insLib_.push_back(mode != DINDEL_INS_NONE ? eng_.libraryIndex(RT::library(reads[r])) : 0);
            anyIns_ = anyIns_ || mode != DINDEL_INS_NONE;
        }
        hap_start_.push_back(leftPos);
        win_hap_.push_back((int32_t)(hap_off_.size() - 1)); win_read_.push_back((int32_t)(read_off_.size() - 1));
        win_entry_.push_back((int64_t)e_hap_.size()); win_key_.push_back(win_key_.back() + (int64_t)keys.size());
        keys_.push_back(keys);
        return (int)hap_start_.size() - 1;
    }
    void flush();
    template <class MLA>
    void result(int w, const std::vector<HapT> &haps, const std::vector<ReadT> &reads,
                std::vector<std::vector<MLA> > &liks, std::vector<int> &onHap) const;
    template <class CovMap>
    int filterResult(int w, std::vector<int> &filtered, CovMap &varCoverage) const;
    // K2 results of window w (flush() computes them when the engine's setInference asked for sums): hands them to the
    // window's Inference guard
    void bindSums(int w, Inference &inf) const
    {
        if (!flushed_ || sums_.empty() || w < 0 || (size_t)w >= size()) { inf.setSums(0, 0, 0, 0, 0); return; }
        const int G = G_ > 0 ? G_ : 1;
        inf.setSums(&sums_[(size_t)sum_off_[w] * G], (size_t)(win_hap_[w + 1] - win_hap_[w]), (size_t)(win_read_[w + 1] - win_read_[w]),
                    G_, G_ > 0 ? &poolSize_[(size_t)w * G_] : 0);
    }
    // K3 for window w of the flushed (still resident) batch: one EM per row of `compatible` (nSets x nh)
    void windowEM(int w, const std::vector<uint8_t> &compatible, size_t nSets, double a0, double tol, std::vector<double> &loglik,
                  std::vector<int32_t> &iters, std::vector<double> &pi, std::vector<double> &margin) const
    {
        const size_t nh = (size_t)(win_hap_[w + 1] - win_hap_[w]);
        loglik.assign(nSets, 0.0); iters.assign(nSets, 0); pi.assign(nSets * nh, 0.0); margin.assign(nSets, 0.0);
        if (nSets == 0) return;
        std::vector<int32_t> sw(nSets, (int32_t)w);
        std::vector<int64_t> off(nSets + 1, 0);
        for (size_t k = 0; k <= nSets; k++) off[k] = (int64_t)(k * nh);
        std::vector<double> freqs(nSets * nh, 0.0);
        dindel_em_in in; in.n_sets = (int32_t)nSets; in.reserved = 0; in.set_window = &sw[0]; in.set_off = &off[0];
        in.compatible = &compatible[0]; in.a0 = a0; in.tol = tol;
        dindel_em_out out; out.loglik = &loglik[0]; out.iters = &iters[0]; out.pi = &pi[0]; out.freqs = &freqs[0];
        out.margin = &margin[0]; out.energy = 0;
        int rc = dindel_b200_bayes_em(eng_.ctx(), &in, &out);
        if (rc != DINDEL_OK) throw std::string("dindel_b200: ") + dindel_b200_strerror(rc) + " (" + dindel_b200_last_error(eng_.ctx()) + ")";
    }
    // K6 for window w of the flushed batch, while it is still resident on the engine's own device (not when the batch
    // was sharded over several GPUs or streamed in more than one range)
    bool statsResident() const { return flushed_ && resident_; }
    void windowStats(int w, const std::vector<StatsJob> &jobs, std::vector<dindel_read_stats> &out) const
    {
        detail::runStats(eng_.ctx(), w, jobs, pool_.empty() ? 0 : &pool_[0], out);
        eng_.nStatsCalls++; eng_.nStatsJobs += (long)jobs.size(); // callers serialise GPU calls (the driver's gpuLock)
    }
    // Realigned-read CIGARs of EVERY window of the flushed batch in one K5 call (see Engine::realignedCigars).
    // allHaps[w] = the haplotypes window w was added with (their ml.hpos / countIndels() are read here: the
    // haplotype-to-reference alignment is not part of add()); refSeqPos[w]; h1[w] / h2[w] = the MAP pair of window w
    // (getMaxHap), h1[w] < 0 skips the window; pooled = the choice of DInDel.cpp:503-509 instead (h1 / h2 unused).
    template <class CigarT>
    void cigars(const std::vector<std::vector<HapT> > &allHaps, const std::vector<int> &refSeqPos, bool pooled,
                const std::vector<int> &h1, const std::vector<int> &h2, std::vector<std::vector<CigarT> > &out) const;
private:
    Engine &eng_;
    int maxMismatch_;
    bool doFilter_, flushed_, full_, anyIns_, resident_;
    size_t maxR_;
    std::vector<uint8_t> insMode_;
    std::vector<int32_t> insAnchor_, insLib_;
    int stride_;
    std::vector<uint32_t> hap_start_, start_;
    std::vector<int32_t> win_hap_, win_read_, e_hap_, e_key_, e_left_, e_right_, status_, nf_, nr_;
    std::vector<int64_t> hap_off_, read_off_, win_entry_, win_key_, win_pair_;
    std::string hap_seq_, read_seq_;
    std::vector<uint8_t> qual_, mapq_, flags_, strand_, e_type_, on_, filt_;
    std::vector<std::vector<PAV> > keys_;
    std::vector<dindel_pair_out> pairs_;
    std::vector<int16_t> hpos_;
    std::vector<int32_t> pool_;
    std::vector<double> sums_;          // K2: [sum_off_[w] * G + g * npairs + q]
    std::vector<int64_t> sum_off_;
    std::vector<int> poolSize_;         // [w * G + g] reads of window w in pool g
    int G_;
};


template <class HapT, class ReadT>
void WindowBatch<HapT, ReadT>::flush()
{
    const size_t nw = size();
    resident_ = false;
    if (nw == 0) { flushed_ = true; return; }
    static const uint8_t none8 = 0;
    static const uint32_t none32 = 0;
    static const int32_t zero32 = 0;
    dindel_batch_in in;
    in.n_windows = (int32_t)nw; in.n_haps = win_hap_.back(); in.n_reads = win_read_.back(); in.reserved = 0;
    in.hap_start = &hap_start_[0]; in.win_hap = &win_hap_[0]; in.win_read = &win_read_[0];
    in.hap_off = &hap_off_[0]; in.hap_seq = hap_seq_.data(); in.read_off = &read_off_[0]; in.read_seq = read_seq_.data();
    in.read_qual = qual_.empty() ? &none8 : &qual_[0]; in.read_mapq = mapq_.empty() ? &none8 : &mapq_[0];
    in.read_start = start_.empty() ? &none32 : &start_[0]; in.read_flags = flags_.empty() ? &none8 : &flags_[0];
    in.read_ins_mode = anyIns_ ? &insMode_[0] : 0; in.read_ins_anchor = anyIns_ ? &insAnchor_[0] : 0; in.read_lib = anyIns_ ? &insLib_[0] : 0;
    const int64_t np = dindel_b200_count_pairs(&in);
    stride_ = (int)((maxR_ + 7) & ~(size_t)7); if (stride_ < 8) stride_ = 8;
    pairs_.resize((size_t)(np ? np : 1));
    if (full_) hpos_.resize((size_t)(np ? np : 1) * (size_t)stride_);
    win_pair_.assign(nw + 1, 0); status_.assign(nw, 0); on_.assign(in.n_reads ? in.n_reads : 1, 0);
    dindel_batch_out out;
    out.n_pairs = np; out.pairs = &pairs_[0]; out.hpos = full_ ? &hpos_[0] : 0; out.hpos_stride = full_ ? stride_ : 0;
    out.flags = DINDEL_OUT_KEEP_HPOS; out.win_pair = &win_pair_[0]; out.win_status = &status_[0]; out.read_on_hap = &on_[0];
    // K4 over the rows K1 leaves in HBM, and the read sums of the inference (K2), in the same pass over the batch
    dindel_cover_in ci;
    ci.n_entries = (int32_t)e_hap_.size(); ci.n_keys = (int32_t)win_key_.back(); ci.maxMismatch = maxMismatch_; ci.doFilter = doFilter_ ? 1 : 0;
    ci.win_entry = &win_entry_[0]; ci.win_key = &win_key_[0];
    ci.entry_hap = e_hap_.empty() ? &zero32 : &e_hap_[0]; ci.entry_key = e_key_.empty() ? &zero32 : &e_key_[0];
    ci.entry_type = e_type_.empty() ? &none8 : &e_type_[0];
    ci.entry_left = e_left_.empty() ? &zero32 : &e_left_[0]; ci.entry_right = e_right_.empty() ? &zero32 : &e_right_[0];
    ci.read_strand = strand_.empty() ? &none8 : &strand_[0];
    filt_.assign(in.n_haps ? in.n_haps : 1, 0);
    nf_.assign(ci.n_keys ? ci.n_keys : 1, 0); nr_.assign(ci.n_keys ? ci.n_keys : 1, 0);
    dindel_cover_out co;
    co.hap_filtered = &filt_[0]; co.hap_uncovered = 0; co.key_nf = &nf_[0]; co.key_nr = &nr_[0];
    dindel_job job;
    job.in = &in; job.out = &out; job.cover = &ci; job.cover_out = &co; job.pair_sums = 0; job.win_sum_off = 0;
    job.n_groups = 0; job.reserved = 0; job.read_group = 0; job.em = 0; job.em_out = 0;
    sums_.clear(); G_ = 0;
    if (eng_.wantSums()) {
        G_ = eng_.windowPools();
        const int G = G_ > 0 ? G_ : 1;
        sum_off_.assign(nw + 1, 0);
        for (size_t w = 0; w < nw; w++) {
            const int64_t nh = win_hap_[w + 1] - win_hap_[w];
            sum_off_[w + 1] = sum_off_[w] + nh * (nh + 1) / 2;
        }
        sums_.assign((size_t)(sum_off_[nw] ? sum_off_[nw] : 1) * (size_t)G, 0.0);
        job.pair_sums = &sums_[0];
        if (G_ > 0) {
            job.n_groups = G_; job.read_group = pool_.empty() ? &zero32 : &pool_[0];
            poolSize_.assign(nw * (size_t)G_, 0);
            for (size_t w = 0; w < nw; w++)
                for (int32_t r = win_read_[w]; r < win_read_[w + 1]; r++)
                    if (pool_[r] >= 0 && pool_[r] < G_) poolSize_[w * (size_t)G_ + (size_t)pool_[r]]++;
        }
    }
    int rc;
    if (eng_.multi()) { // the batch sharded over the GPUs of the box; results land in window order
        rc = dindel_b200_multi_process(eng_.multi(), &job);
        if (rc != DINDEL_OK)
            throw std::string("dindel_b200: ") + dindel_b200_strerror(rc) + " (" + dindel_b200_multi_last_error(eng_.multi()) + ")";
    } else {
        rc = dindel_b200_process(eng_.ctx(), &job);
        if (rc != DINDEL_OK)
            throw std::string("dindel_b200: ") + dindel_b200_strerror(rc) + " (" + dindel_b200_last_error(eng_.ctx()) + ")";
        dindel_timing t;
        resident_ = dindel_b200_get_timing(eng_.ctx(), &t) == DINDEL_OK && t.n_ranges == 1;
    }
    flushed_ = true;
}

// DetInDel::computeLikelihoods' results for window w, with its guards (see Engine::computeLikelihoods).
template <class HapT, class ReadT>
template <class MLA>
void WindowBatch<HapT, ReadT>::result(int w, const std::vector<HapT> &haps, const std::vector<ReadT> &reads,
                                      std::vector<std::vector<MLA> > &liks, std::vector<int> &onHap) const
{
    if (!flushed_ || w < 0 || (size_t)w >= size()) throw std::string("dindel_b200: result() before flush() / bad window");
    const size_t nh = haps.size(), nr = reads.size();
    if ((size_t)(win_hap_[w + 1] - win_hap_[w]) != nh || (size_t)(win_read_[w + 1] - win_read_[w]) != nr)
        throw std::string("dindel_b200: result() called with another window's haplotypes / reads");
    onHap = std::vector<int>(nr, 0);
    liks = std::vector<std::vector<MLA> >(nh, std::vector<MLA>(nr));
    const int pad = eng_.params().padCover;
    for (size_t h = 0; h < nh; h++)
        for (size_t r = 0; r < nr; r++) {
            const size_t i = (size_t)win_pair_[w] + r * nh + h;
            const dindel_pair_out &rec = pairs_[i];
            MLA &ml = liks[h][r];
            if (rec.status == DINDEL_WIN_HAPSIZE) throw std::string("hapSize error.");
            if (full_) {
                expandAlignment(rec, &hpos_[i * (size_t)stride_], haps[h], RT::seq(reads[r]), pad, ml);
            } else {
                ml.ll = rec.ll; ml.llOn = rec.llOn; ml.llOff = rec.llOff; ml.offHap = rec.offHap != 0; ml.offHapHMQ = rec.offHapHMQ != 0;
                ml.firstBase = rec.firstBase; ml.lastBase = rec.lastBase; ml.nBQT = rec.nBQT; ml.nmmBQT = rec.nmmBQT;
                ml.mLogBQ = rec.mLogBQ; ml.nMMRight = rec.nMMRight; ml.nMMLeft = rec.nMMLeft;
                ml.numIndels = rec.numIndels; ml.numMismatch = rec.numMismatch;
                for (typename std::map<int, AV>::const_iterator it = haps[h].indels.begin(); it != haps[h].indels.end(); ++it)
                    ml.hapIndelCovered[it->first] = it->second.isCovered(pad, ml.firstBase, ml.lastBase);
                for (typename std::map<int, AV>::const_iterator it = haps[h].snps.begin(); it != haps[h].snps.end(); ++it)
                    ml.hapSNPCovered[it->first] = it->second.isCovered(pad, ml.firstBase, ml.lastBase);
            }
            if (!ml.offHapHMQ) onHap[r] = 1;
            if (rec.status == DINDEL_WIN_LL_POSITIVE) { std::cerr << "Likelihood>0" << std::endl; exit(1); }
            if (rec.status == DINDEL_WIN_NAN) { std::cout << "NAN/Inf error" << std::endl; throw std::string("Nan detected"); }
        }
}

// DetInDel::filterHaplotypes' results for window w (computed by K4 during flush()).
template <class HapT, class ReadT>
template <class CovMap>
int WindowBatch<HapT, ReadT>::filterResult(int w, std::vector<int> &filtered, CovMap &varCoverage) const
{
    typedef typename CovMap::mapped_type Coverage;
    if (!flushed_ || w < 0 || (size_t)w >= size()) throw std::string("dindel_b200: filterResult() before flush() / bad window");
    const size_t nh = (size_t)(win_hap_[w + 1] - win_hap_[w]);
    filtered = std::vector<int>(nh, 0);
    varCoverage.clear();
    int n = 0;
    for (size_t h = 0; h < nh; h++) { filtered[h] = filt_[win_hap_[w] + h]; n += filtered[h]; }
    for (size_t k = 0; k < keys_[w].size(); k++)
        varCoverage[keys_[w][k]] = Coverage(nf_[win_key_[w] + k], nr_[win_key_[w] + k]);
    return n;
}

template <class HapT, class ReadT, class MLA, class CovMap>
int Engine::filterHaplotypes(const std::vector<HapT> &haps, const std::vector<ReadT> &reads,
                             const std::vector<std::vector<MLA> > & /*liks: resident on the GPU*/,
                             std::vector<int> &filtered, CovMap &varCoverage, bool doFilter, int maxMismatch)
{
    typedef ReadTraits<ReadT> RT;
    typedef typename HapT_variant<HapT>::type AV;
    typedef typename CovMap::key_type PAV;           // pair<int, AlignedVariant> (DInDel.cpp:1943)
    typedef typename CovMap::mapped_type Coverage;   // VariantCoverage(nf, nr) (DInDel.hpp:53-66)
    Tick tick(&t_fh_);
    const size_t nh = haps.size(), nr = reads.size();
    filtered = std::vector<int>(nh, 0);
    varCoverage.clear();
    if (nh == 0) return 0;
    if (nh != last_nh_ || nr != last_nr_)
        throw std::string("dindel_b200: filterHaplotypes called on a window other than the resident one");
    std::map<PAV, int> keyOf; // PAV -> key index, in the map's own order (DInDel.cpp:1944)
    std::vector<int32_t> e_hap, e_key, e_left, e_right;
    std::vector<uint8_t> e_type, strand(nr ? nr : 1);
    for (size_t h = 0; h < nh; h++)
        for (typename std::map<int, AV>::const_iterator it = haps[h].indels.begin(); it != haps[h].indels.end(); ++it) {
            const AV &av = it->second;
            const PAV pav(it->first, av);
            typename std::map<PAV, int>::iterator k = keyOf.find(pav);
            if (k == keyOf.end()) k = keyOf.insert(std::make_pair(pav, (int)keyOf.size())).first;
            e_hap.push_back((int32_t)h);
            e_key.push_back(k->second);
            e_type.push_back(av.getType() == AV::DEL ? DINDEL_VAR_DEL : (av.getType() == AV::INS ? DINDEL_VAR_INS : DINDEL_VAR_OTHER));
            e_left.push_back(av.getLeftFlankRead() - p_.padCover);    // DInDel.cpp:1973
            e_right.push_back(av.getRightFlankRead() + p_.padCover);  // DInDel.cpp:1974
        }
    for (size_t r = 0; r < nr; r++) strand[r] = RT::coverReverse(reads[r]) ? 1 : 0;
    const int64_t win_entry[2] = {0, (int64_t)e_hap.size()}, win_key[2] = {0, (int64_t)keyOf.size()};
    static const int32_t zero32 = 0;
    static const uint8_t zero8 = 0;
    dindel_cover_in in;
    in.n_entries = (int32_t)e_hap.size(); in.n_keys = (int32_t)keyOf.size();
    in.maxMismatch = maxMismatch; in.doFilter = doFilter ? 1 : 0;
    in.win_entry = win_entry; in.win_key = win_key;
    in.entry_hap = e_hap.empty() ? &zero32 : &e_hap[0]; in.entry_key = e_key.empty() ? &zero32 : &e_key[0];
    in.entry_type = e_type.empty() ? &zero8 : &e_type[0];
    in.entry_left = e_left.empty() ? &zero32 : &e_left[0]; in.entry_right = e_right.empty() ? &zero32 : &e_right[0];
    in.read_strand = &strand[0];
    std::vector<uint8_t> filt(nh), unc(nh);
    std::vector<int32_t> nf(keyOf.size() ? keyOf.size() : 1), nrv(keyOf.size() ? keyOf.size() : 1);
    dindel_cover_out out;
    out.hap_filtered = &filt[0]; out.hap_uncovered = &unc[0]; out.key_nf = &nf[0]; out.key_nr = &nrv[0];
    int rc = dindel_b200_coverage(ctx_, &in, &out);
    if (rc != DINDEL_OK)
        throw std::string("dindel_b200: ") + dindel_b200_strerror(rc) + " (" + dindel_b200_last_error(ctx_) + ")";
    int numFiltered = 0;
    for (size_t h = 0; h < nh; h++) { filtered[h] = filt[h]; numFiltered += filt[h]; }
    // The reference creates a key when a haplotype's walk reaches the variant (DInDel.cpp:1966-1969) and
    // consumers read varCoverage[pav] with operator[] (absent == (0,0)), so every key is materialised here.
    for (typename std::map<PAV, int>::const_iterator k = keyOf.begin(); k != keyOf.end(); ++k)
        varCoverage[k->first] = Coverage(nf[k->second], nrv[k->second]);
    return numFiltered;
}

template <class HapT, class ReadT>
template <class CigarT>
void WindowBatch<HapT, ReadT>::cigars(const std::vector<std::vector<HapT> > &allHaps, const std::vector<int> &refSeqPos,
                                      bool pooled, const std::vector<int> &h1, const std::vector<int> &h2,
                                      std::vector<std::vector<CigarT> > &out) const
{
    const size_t nw = size();
    if (!flushed_) throw std::string("dindel_b200: cigars() before flush()");
    if (allHaps.size() != nw || refSeqPos.size() != nw || (!pooled && (h1.size() != nw || h2.size() != nw)))
        throw std::string("dindel_b200: cigars(): one entry per window expected");
    out.assign(nw, std::vector<CigarT>());
    const size_t nr = nw ? (size_t)win_read_.back() : 0;
    if (nr == 0) return;
    std::vector<int32_t> refpos, nind, wh1(nw, -1), wh2(nw, -1), wstart(nw);
    for (size_t w = 0; w < nw; w++) {
        if (allHaps[w].size() != (size_t)(win_hap_[w + 1] - win_hap_[w])) throw std::string("dindel_b200: cigars(): haplotypes of another window");
        for (size_t h = 0; h < allHaps[w].size(); h++) {
            const HapT &hp = allHaps[w][h];
            if (hp.ml.hpos.size() != hp.seq.size()) throw std::string("Haplotype has not been aligned!"); // DInDel.cpp:730
            refpos.insert(refpos.end(), hp.ml.hpos.begin(), hp.ml.hpos.end());
            nind.push_back(hp.countIndels());
        }
        wstart[w] = refSeqPos[w];
        if (!pooled) { wh1[w] = h1[w]; wh2[w] = h2[w]; }
    }
    static const int32_t zero32 = 0;
    dindel_cigar_in in;
    in.mode = pooled ? DINDEL_CIGAR_BEST : DINDEL_CIGAR_PAIR;
    in.win_h1 = &wh1[0]; in.win_h2 = &wh2[0]; in.win_ref_start = &wstart[0];
    in.hap_n_indels = nind.empty() ? &zero32 : &nind[0]; in.hap_ref_pos = refpos.empty() ? &zero32 : &refpos[0];
    std::vector<int32_t> rh, rp, rn;
    std::vector<uint32_t> ops;
    const int cap = detail::run_cigars(eng_.ctx(), in, nr, rh, rp, rn, ops);
    for (size_t w = 0; w < nw; w++) {
        out[w].assign((size_t)(win_read_[w + 1] - win_read_[w]), CigarT());
        for (size_t r = (size_t)win_read_[w]; r < (size_t)win_read_[w + 1]; r++) {
            if (rn[r] == -5) throw std::string("How is this possible? (1)");
            if (rn[r] < 0) { char msg[32]; snprintf(msg, sizeof msg, "Error(%d)!", -rn[r]); throw std::string(msg); }
                CigarT &c = out[w][r - (size_t)win_read_[w]];
            for (int k = 0; k < rn[r]; k++) {
                const uint32_t x = ops[r * cap + k];
                c.push_back(typename CigarT::value_type((int)(x & 15u), (int)(x >> 4)));
            }
            c.refPos = rp[r];
        }
    }
}

template <class HapT, class CigarT>
void Engine::realignedCigars(const std::vector<HapT> &haps, size_t nReads, int refSeqPos, int h1, int h2,
                             std::vector<CigarT> &cigars)
{
    const size_t nh = haps.size(), nr = nReads;
    cigars.assign(nr, CigarT());
    if (nh == 0 || nr == 0) return;
    if (nh != last_nh_ || nr != last_nr_)
        throw std::string("dindel_b200: realignedCigars called on a window other than the resident one");
    std::vector<int32_t> refpos, nind(nh);
    for (size_t h = 0; h < nh; h++) {
        if (haps[h].ml.hpos.size() != haps[h].seq.size()) throw std::string("Haplotype has not been aligned!"); // DInDel.cpp:730
        refpos.insert(refpos.end(), haps[h].ml.hpos.begin(), haps[h].ml.hpos.end());
        nind[h] = haps[h].countIndels();
    }
    static const int32_t zero32 = 0;
    const int32_t wh1 = h1, wh2 = h2, wstart = refSeqPos;
    dindel_cigar_in in;
    in.mode = h1 < 0 ? DINDEL_CIGAR_BEST : DINDEL_CIGAR_PAIR;
    in.win_h1 = &wh1; in.win_h2 = &wh2; in.win_ref_start = &wstart;
    in.hap_n_indels = &nind[0]; in.hap_ref_pos = refpos.empty() ? &zero32 : &refpos[0];
    std::vector<int32_t> rh, rp, rn;
    std::vector<uint32_t> ops;
    const int cap = detail::run_cigars(ctx_, in, nr, rh, rp, rn, ops);
    for (size_t r = 0; r < nr; r++) {
        if (rn[r] == -5) throw std::string("How is this possible? (1)");                 // DInDel.cpp:827
        if (rn[r] < 0) { char msg[32]; snprintf(msg, sizeof msg, "Error(%d)!", -rn[r]); throw std::string(msg); }
        CigarT &c = cigars[r];
        for (int k = 0; k < rn[r]; k++) {
            const uint32_t x = ops[r * cap + k];
            c.push_back(typename CigarT::value_type((int)(x & 15u), (int)(x >> 4)));
        }
        c.refPos = rp[r];
    }
}

} // namespace dindel_b200
#endif
