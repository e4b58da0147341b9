/* oracle/ref_harness.cpp — TEST INFRASTRUCTURE ONLY.
 *
 * Thin extern "C" harness around the reference's OWN, unmodified ObservationModelFBMaxErr
 * (/root/reference/ObservationModelFB.{hpp,cpp}); it is compiled together with the reference
 * sources where they lie (see oracle/Makefile) into oracle/_ref/libdindel_ref.so.  Nothing here
 * restates the algorithm: it only builds the reference's Haplotype / Read value types from flat
 * buffers, calls the reference class, and flattens MLAlignment into a POD.
 *
 * Used by: tests/ (to pin the C restatement in oracle/dindel_oracle.c), __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference arm.  Never linked into the product library.
 *
 * Mirrors the loop of DetInDel::computeLikelihoods (DInDel.cpp:1707-1739) in ref_window().
 */
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <string>
#include <vector>
#include <iostream>
#include <sstream>
#include <cmath>
#include "ObservationModelFB.hpp"

extern "C" {

/* Must match oracle/dindel_oracle.h: ora_params / ora_pair_out (checked by tests via sizeof exports). */
struct ref_params {
    double pError, pMut, mapQualThreshold, pFirstgLO, checkBaseQualThreshold;
    int32_t maxLengthDel, padCover, bMid, forceReadOnHaplotype;
};

struct ref_pair_out {
    double ll, llOn, llOff, mLogBQ;
    int32_t firstBase, lastBase;
    int32_t numIndels, numMismatch, nBQT, nmmBQT, nMMLeft, nMMRight;
    int32_t offHap, offHapHMQ;
    int32_t bMid, status; /* status: 0 ok, 1 "hapSize error." thrown */
};

int ref_sizeof_params(void) { return (int)sizeof(ref_params); }
int ref_sizeof_pair_out(void) { return (int)sizeof(ref_pair_out); }

static void fill_params(const ref_params *p, ObservationModelParameters &o)
{
    o.pError = p->pError;
    o.pMut = p->pMut;
    o.mapQualThreshold = p->mapQualThreshold;
    o.pFirstgLO = p->pFirstgLO;
    o.checkBaseQualThreshold = p->checkBaseQualThreshold;
    o.maxLengthDel = p->maxLengthDel;
    o.padCover = p->padCover;
    o.bMid = p->bMid;
    o.forceReadOnHaplotype = p->forceReadOnHaplotype != 0;
}

/* The reference's Read() default ctor leaves `bam` dangling while Init() calls read.isUnmapped();
 * attach a zeroed bam1_t (flag carries BAM_FUNMAP when asked) exactly as a BAM-backed Read has. */
static void attach_bam(Read &r, int unmapped)
{
    r.bam = new bam1_t;
    memset(r.bam, 0, sizeof(bam1_t));
    r.bam->data = new uint8_t[1];
    r.bam->m_data = 1;
    r.bam->data[0] = 0;
    if (unmapped) r.bam->core.flag |= BAM_FUNMAP;
    r.initBam = true;
}

static void make_read(Read &r, const char *seq, int R, const double *qual, double mapQual,
                      double posStatFirst, int unmapped)
{
    r.seq.seq.assign(seq, R);
    r.qual.assign(qual, qual + R);
    r.mapQual = mapQual;
    r.posStat.first = posStatFirst;
    r.pos = (int32_t)posStatFirst;
    attach_bam(r, unmapped);
}

static void flatten(const MLAlignment &ml, int bMid, ref_pair_out *out, int32_t *hpos, int R)
{
    out->ll = ml.ll; out->llOn = ml.llOn; out->llOff = ml.llOff; out->mLogBQ = ml.mLogBQ;
    out->firstBase = ml.firstBase; out->lastBase = ml.lastBase;
    out->numIndels = ml.numIndels; out->numMismatch = ml.numMismatch;
    out->nBQT = ml.nBQT; out->nmmBQT = ml.nmmBQT;
    out->nMMLeft = ml.nMMLeft; out->nMMRight = ml.nMMRight;
    out->offHap = ml.offHap ? 1 : 0; out->offHapHMQ = ml.offHapHMQ ? 1 : 0;
    out->bMid = bMid; out->status = 0;
    if (hpos) for (int b = 0; b < R; b++) hpos[b] = ml.hpos[b];
}

/* Access to the protected bMid / mapState for parity checks. */
class Probe : public ObservationModelFBMaxErr {
public:
    Probe(const Haplotype &h, const Read &r, uint32_t hs, const ObservationModelParameters &p)
        : ObservationModelFBMaxErr(h, r, hs, p) {}
    int getBMid() const { return bMid; }
};

/* One (haplotype, read) pair through the reference class.  qual[] and mapQual are probabilities
 * (as Read::qual / Read::mapQual hold them).  mapState (optional) receives the Viterbi path. */
int ref_pair(const char *hap, int H, const char *read, int R, const double *qual, double mapQual,
             double posStatFirst, int unmapped, uint32_t hapStart, const ref_params *p,
             ref_pair_out *out, int32_t *hpos, int32_t *mapState)
{
    ObservationModelParameters op;
    fill_params(p, op);
    Haplotype h;
    h.seq.assign(hap, H);
    Read r;
    make_read(r, read, R, qual, mapQual, posStatFirst, unmapped);
    /* the reference prints BMIDERROR on stdout; keep the harness quiet */
    std::streambuf *old = std::cout.rdbuf();
    std::ostringstream sink;
    std::cout.rdbuf(sink.rdbuf());
    int rc = 0;
    try {
        Probe om(h, r, hapStart, op);
        MLAlignment ml = om.calcLikelihood();
        flatten(ml, om.getBMid(), out, hpos, R);
        if (mapState) {
            std::vector<int> ms = om.getMapState();
            for (int b = 0; b < R; b++) mapState[b] = ms[b];
        }
    } catch (std::string &s) {
        memset(out, 0, sizeof(*out));
        out->status = 1;
        rc = 1;
    }
    std::cout.rdbuf(old);
    return rc;
}

/* Library(counts) as LibraryCollection::addFromFile builds it (Library.hpp:31-129): exports what the product is
 * given through dindel_b200_set_libraries: probs[0..maxins), and the 95th-percentile probability.  Returns maxins. */
int ref_library(const double *counts, int n, double *probs_out, double *p95_out)
{
    std::streambuf *old = std::cout.rdbuf();
    std::ostringstream sink;
    std::cout.rdbuf(sink.rdbuf());
    Library lib(std::vector<double>(counts, counts + n));
    std::cout.rdbuf(old);
    const int maxins = lib.getMaxInsertSize();
    for (int d = 0; d < maxins; d++) probs_out[d] = lib.getProb(d);
    *p95_out = lib.getNinetyFifthPctProb();
    return maxins;
}

/* One pair with the insert-size prior active (ObservationModelFB.cpp:279-293): a paired read whose mate is mapped
 * on the same target, params.mapUnmappedReads set as `--libFile` does (DInDel.cpp:4268-4272).  ins_mode 1 = mate on
 * the reverse strand (anchor = matePos + mateLen), 2 = forward (anchor = matePos), 0 = paired but mate unmapped. */
int ref_pair_ins(const char *hap, int H, const char *read, int R, const double *qual, double mapQual,
                 double posStatFirst, int unmapped, uint32_t hapStart, const ref_params *p,
                 ref_pair_out *out, int32_t *hpos, int32_t *mapState, int ins_mode, int ins_anchor,
                 const double *counts, int n_counts)
{
    ObservationModelParameters op;
    fill_params(p, op);
    op.mapUnmappedReads = true;
    std::streambuf *old = std::cout.rdbuf();
    std::ostringstream sink;
    std::cout.rdbuf(sink.rdbuf());
    Library lib(std::vector<double>(counts, counts + n_counts));
    Haplotype h;
    h.seq.assign(hap, H);
    Read r;
    make_read(r, read, R, qual, mapQual, posStatFirst, unmapped);
    r.bam->core.flag |= BAM_FPAIRED;
    r.bam->core.tid = 0; r.bam->core.mtid = 0;
    r.library = &lib;
    if (ins_mode == 0) {
        r.bam->core.flag |= BAM_FMUNMAP;
    } else if (ins_mode == 1) {
        r.bam->core.flag |= BAM_FMREVERSE;
        r.mateLen = 100; r.matePos = ins_anchor - 100;
    } else {
        r.mateLen = 100; r.matePos = ins_anchor;
    }
    int rc = 0;
    try {
        Probe om(h, r, hapStart, op);
        MLAlignment ml = om.calcLikelihood();
        flatten(ml, om.getBMid(), out, hpos, R);
        if (mapState) {
            std::vector<int> ms = om.getMapState();
            for (int b = 0; b < R; b++) mapState[b] = ms[b];
        }
    } catch (std::string &s) {
        memset(out, 0, sizeof(*out));
        out->status = 1;
        rc = 1;
    }
    std::cout.rdbuf(old);
    return rc;
}

/* Phred -> probability exactly as Read::Read(const bam1_t*, ...) does (Read.hpp:127-131,143-148). */
double ref_phred_to_prob(int phred)
{
    double ph = (double)phred;
    double q = (1.0 - pow(10.0, -ph / 10.0));
    if (q < 1e-16) q = 1e-16;
    if (q > 1.0 - 1e-16) q = 1.0 - 1e-16;
    return q;
}

/* A whole window: liks[h][r] for every hap x read, the double loop of
 * DetInDel::computeLikelihoods (DInDel.cpp:1715-1737).  Inputs are flat: haps concatenated with
 * hap_off[nh+1]; reads concatenated with read_off[nr+1]; phred bytes for base and mapping
 * quality; mReadStart (= uint32_t(posStat.first)); flags bit0 = unmapped.  Outputs are r-major
 * (out[r*nh+h]) like rl[] in DInDel.cpp:2129; hpos (optional) is [pair][max R] packed by
 * hpos_stride.  Returns number of pairs computed, or -1 when the reference threw. */
int ref_window(const char *hap_seq, const int32_t *hap_off, int nh, const char *read_seq,
               const uint8_t *read_qual, const int32_t *read_off, const uint8_t *mapq,
               const uint32_t *mReadStart, const uint8_t *flags, int nr, uint32_t hapStart,
               const ref_params *p, ref_pair_out *out, int32_t *hpos, int hpos_stride)
{
    ObservationModelParameters op;
    fill_params(p, op);
    std::vector<Haplotype> haps(nh);
    for (int h = 0; h < nh; h++) haps[h].seq.assign(hap_seq + hap_off[h], hap_off[h + 1] - hap_off[h]);
    std::vector<Read> reads(nr);
    std::vector<double> q;
    for (int r = 0; r < nr; r++) {
        int R = read_off[r + 1] - read_off[r];
        q.resize(R);
        for (int b = 0; b < R; b++) q[b] = ref_phred_to_prob(read_qual[read_off[r] + b]);
        make_read(reads[r], read_seq + read_off[r], R, R ? &q[0] : 0, ref_phred_to_prob(mapq[r]),
                  (double)mReadStart[r], flags[r] & 1);
    }
    std::streambuf *old = std::cout.rdbuf();
    std::ostringstream sink;
    std::cout.rdbuf(sink.rdbuf());
    int n = 0;
    try {
        for (int h = 0; h < nh; h++)
            for (int r = 0; r < nr; r++) {
                Probe om(haps[h], reads[r], hapStart, op);
                MLAlignment ml = om.calcLikelihood();
                int R = read_off[r + 1] - read_off[r];
                size_t pi = (size_t)r * nh + h;
                flatten(ml, om.getBMid(), &out[pi], hpos ? hpos + pi * hpos_stride : 0, R);
                n++;
            }
    } catch (std::string &s) {
        n = -1;
    }
    std::cout.rdbuf(old);
    return n;
}

} /* extern "C" */
