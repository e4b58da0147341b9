// k1_common.cuh — device-side data layout shared by the K1 (pair Viterbi) kernels.
//
// HBM layout of one resident batch (all arrays device pointers, SoA, CSR offsets; see
// include/dindel_b200.h for the host mirror):
//   windows : hap_start[nw], win_hap[nw+1], win_read[nw+1], win_pair[nw+1], win_unit[nw+1]
//   haps    : hap_off[nh+1] -> hap_seq bytes;  hap_hp: one homopolymer code per state column,
//             column base of hap h = hap_off[h] + 2*h, H+2 columns (x = 0 LO, 1..H, H+1 RO)
//   reads   : read_off[nr+1] -> read_seq / read_qual bytes; read_mapq, read_start, read_flags
//   results : pairs[n_pairs] (64 B each, r-major inside a window), optional hpos int16
// All transcendental values live in K1Luts, computed on the host with glibc (the reference's
// libm), so device arithmetic is FP64 add / compare / select only => bit-exact log-likelihoods.
#pragma once
#include <stdint.h>
#include "../../include/dindel_b200.h"

#define K1_HP_LUT 64     // homopolymer run-length codes 0 (default) .. 63 (>= 63, capped 0.99)
#ifndef K1_READ_CHUNK
#define K1_READ_CHUNK 4 // reads per work unit (window, hap, chunk): small units keep launch tails short
#endif
#define K1_TILED_SLOTS 16   // RO slots per shape class (columns per lane <= 16)
#define K1_TILED_BUCKETS 96 // 6 shape classes x 16 slots

struct K1Luts {
    // ObservationModelFB.cpp:232-234 by base phred (Read.hpp:143-148 conversion included)
    double eq[256], uq[256];
    // log10(1 - qual) (ObservationModelFB.cpp:1406) by base phred
    double log10q[256];
    // bit0: qual > checkBaseQualThreshold (cpp:1404,1420); bit1: qual > 0.95 (cpp:1429)
    uint8_t qflag[256];
    // computeBMidPrior (cpp:268-305) by mapping-quality phred: [plane][phred]
    double priorLO[2][256], priorOn[2][256];
    // the artificial high-mapping-quality prior (mapQual = 1-1e-10, cpp:1093): [plane]
    double priorLO_H[2], priorOn_H[2];
    // the same prior in parts, for reads with an insert-size term pinsert[x] (cpp:283-303):
    // prior[x] = (pinsert[x] + prior1m) + priorIns[plane]; prior[LO] = priorLO + pinsert[0]
    double prior1m[256], prior1m_H, priorIns[2];
    // setupTransitionProbs (cpp:1678-1703) by run-length code; code 0 = the 1e-5 default
    double hp_lpe[K1_HP_LUT], hp_lpn[K1_HP_LUT];
    // cpp:1643-1667
    double lLOgLO, lFirstgLO, lInsgIns, lNoInsgIns, lNoInsgNoIns;
    int32_t maxLengthDel, numT, bMidOverride, force;
};

struct K1Batch {
    int32_t n_windows, n_haps, n_reads, hpos_stride;
    long long n_units, n_pairs;
    // the window range this launch set works on (a whole batch, or one chunk of a pipelined submit)
    int32_t w_begin, w_end;
    long long unit_begin, unit_end; // generic kernel: units of that range (win_unit numbering)
    const uint32_t *hap_start;
    const int32_t *win_hap, *win_read;
    const long long *win_pair, *win_unit;
    const long long *hap_off;
    const uint8_t *hap_seq;
    const uint8_t *hap_hp;
    const long long *read_off;
    const uint8_t *read_seq, *read_qual, *read_mapq, *read_flags;
    const uint32_t *read_start;
    // insert-size prior (all null when no read has one): per read mode / anchor / library, per library the
    // log-probabilities by insert size (lib_off[l] .. lib_off[l+1]) and log(95th-percentile probability)
    const uint8_t *read_ins_mode;
    const int32_t *read_ins_anchor, *read_lib;
    const long long *lib_off;
    const double *lib_logp, *lib_log95;
    dindel_pair_out *pairs;
    int16_t *hpos;
    unsigned long long *counter; // work-queue heads (generic kernel: [0]; tiled kernel: one per class)
    // tiled kernel: work units (window, hap, read chunk) bucketed by shape class and RO slot
    // (k1_tiled_prepare); bucket b = class * K1_TILED_SLOTS + slot owns units[bucket_start[b] .. [b+1])
    uint2 *units;
    unsigned *bucket_start;      // [K1_TILED_BUCKETS + 1]
    unsigned *bucket_work;       // [2 * K1_TILED_BUCKETS] counts, then fill cursors
    int32_t class_c[6];          // tiled kernel: state columns per lane of each shape class (depends on J), 0 = unused
    const K1Luts *luts;
    uint8_t *bt_scratch;         // per-warp back-pointer slots in global memory (generic kernel when smem is
                                 // short; tiled kernel for J > 7, where a pair's rows are too large for smem)
    long long bt_slot_bytes;
    int32_t max_R, max_S;
};

// pinsert[x] of computeBMidPrior (cpp:283-290) for hap column x in [1, H]; plain int arithmetic like the reference
// (its hapStart member is an int, ObservationModelFB.hpp:112), Library::getProb's clamp (Library.hpp:61-66).
__device__ __forceinline__ double k1_pinsert(const double *logp, int maxins, int mode, int anchor, int hapStart,
                                             int x, int R, int bMid)
{
    int d = (mode == DINDEL_INS_MATE_REVERSE) ? (hapStart + x - bMid - anchor) : (hapStart + x + R - bMid - anchor);
    d = d < 0 ? -d : d;
    if (d < 0) d = -d;
    if (d >= maxins) d = maxins - 1;
    return logp[d];
}

// ObservationModelFB.cpp:35-102 (Init): anchor base of one pair.
__device__ __forceinline__ int k1_bmid(uint32_t hapStart, int H, int R, uint32_t mReadStart,
                                       bool unmapped, int bMidOverride)
{
    uint32_t hapEnd = hapStart + (uint32_t)H;
    uint32_t readEnd = mReadStart + (uint32_t)R - 1u;
    int bMid;
    if (unmapped || mReadStart > hapEnd || readEnd < hapStart) {
        bMid = R / 2;
    } else {
        uint32_t olStart = (hapStart > mReadStart) ? hapStart : mReadStart;
        uint32_t olEnd = (hapEnd > readEnd) ? readEnd : hapEnd;
        int mid = ((int)olEnd - (int)olStart) / 2 + (int)olStart;
        bMid = mid - (int)mReadStart;
    }
    if (bMidOverride != -1) bMid = bMidOverride;
    if (bMid < 0) bMid = 0;
    if (bMid >= R) bMid = R - 1;
    return bMid;
}

// Locate the window of a work unit: largest w with win_unit[w] <= unit.
__device__ __forceinline__ int k1_find_window(const long long *win_unit, int nw, long long unit)
{
    int lo = 0, hi = nw; // invariant: win_unit[lo] <= unit < win_unit[hi]
    while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (win_unit[mid] <= unit) lo = mid; else hi = mid;
    }
    return lo;
}

#define K1_NEG_INF (__longlong_as_double(0xfff0000000000000LL))
