/* include/dindel_b200.h — C ABI of the B200 read x candidate-haplotype likelihood engine.
 *
 * Drop-in boundary for the path DetInDel::empiricalDistributionMethod drives in the reference
 * (all citations relative to the reference tree):
 *   - DetInDel::computeLikelihoods (DInDel.cpp:1707-1739), i.e. for every (hap, read) of a window
 *     ObservationModelFBMaxErr(hap, read, leftPos, obsParams).calcLikelihood()
 *     (ObservationModelFB.hpp:160,133; ObservationModelFB.cpp:1068)         -> dindel_b200_submit
 *   - the pair-marginalisation core of DetInDel::diploidGLF (DInDel.cpp:3085-3113,3349-3390)
 *                                                                            -> dindel_b200_pair_sums
 *   - the variational-Bayes EM core of DetInDel::estimateHaplotypeFrequenciesBayesEM
 *     (DInDel.cpp:2431-2543)                                                 -> dindel_b200_bayes_em
 *   - DetInDel::filterHaplotypes (DInDel.cpp:1932-2100), the consumer of MLAlignment::hpos
 *                                                                            -> dindel_b200_coverage
 *   - the realigned-read CIGARs of --outputRealignedBAM: DetInDel::getCIGAR (DInDel.cpp:728-882) and the
 *     haplotype choice at its call sites (:503-509, :597-606)                -> dindel_b200_cigars
 * Plain C, POD only, caller-owned buffers, no exceptions: every entry returns 0 or a negative
 * DINDEL_E_* code (dindel_b200_strerror).  There is NO CPU fallback: without a CUDA device every
 * compute entry fails with DINDEL_E_NO_DEVICE.
 */
#ifndef DINDEL_B200_H
#define DINDEL_B200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* 4: dindel_b200_process / dindel_job, the multi-device calls, margins of the EM.  5 (additive: no existing struct or entry
 * changed): dindel_b200_read_stats (K6), dindel_b200_load_records, dindel_b200_run_static. */
#define DINDEL_B200_ABI_VERSION 5

/* error codes (batch-fatal) */
#define DINDEL_OK 0
#define DINDEL_E_INVALID (-1)    /* bad argument / inconsistent batch */
#define DINDEL_E_NO_DEVICE (-2)  /* no usable CUDA device */
#define DINDEL_E_CUDA (-3)       /* CUDA runtime error (see dindel_b200_last_error) */
#define DINDEL_E_NOMEM (-4)
#define DINDEL_E_UNSUPPORTED (-5) /* shape outside what the kernels are built for */

/* per-window status, mirrors the reference's per-window skip reasons (DInDel.cpp:1361-1408) */
#define DINDEL_WIN_OK 0
#define DINDEL_WIN_HAPSIZE 1     /* throw "hapSize error."  (ObservationModelFB.cpp:47) */
#define DINDEL_WIN_NAN 2         /* throw "Nan detected"    (DInDel.cpp:1732-1735) */
#define DINDEL_WIN_LL_POSITIVE 3 /* ll > 0.1 -> exit(1)     (DInDel.cpp:1722-1731) */

/* hpos codes (MLAlignment.hpp:31-34) */
#define DINDEL_HPOS_INS (-1)
#define DINDEL_HPOS_LO (-3)
#define DINDEL_HPOS_RO (-4)

/* read flag bits */
#define DINDEL_READ_UNMAPPED 1u /* Read::isUnmapped() (Read.hpp:200) */
#define DINDEL_READ_REVERSE 2u  /* Read::onReverseStrand */

/* Replaces ObservationModelParameters (ObservationModel.hpp:28-99): the fields the path reads.
 * dindel_b200_default_params() fills the struct defaults (ObservationModel.hpp:39-62);
 * dindel_b200_cli_params() fills the `dindel` CLI defaults (DInDel.cpp:4122-4157,3943-3948). */
typedef struct dindel_params {
    double pError, pMut, mapQualThreshold, pFirstgLO, checkBaseQualThreshold;
    int32_t maxLengthDel, padCover, bMid, forceReadOnHaplotype;
} dindel_params;

/* One batch of realignment windows, SoA with CSR offsets.  Replaces the (haps, reads, leftPos)
 * arguments of DetInDel::computeLikelihoods (DInDel.cpp:1707).
 * Sequences are the bytes Haplotype::seq / Read::seq hold (ASCII; emission test is plain char
 * equality or hap=='N', ObservationModelFB.cpp:246).  Qualities are the phred bytes of the BAM
 * record (Read.hpp:127,143): the library converts them exactly as Read's ctor does. */
typedef struct dindel_batch_in {
    int32_t n_windows;
    int32_t n_haps;             /* total over windows */
    int32_t n_reads;            /* total over windows */
    int32_t reserved;
    const uint32_t *hap_start;  /* [n_windows]   leftPos of the window (hapStart of Init, cpp:35) */
    const int32_t *win_hap;     /* [n_windows+1] window -> first haplotype */
    const int32_t *win_read;    /* [n_windows+1] window -> first read */
    const int64_t *hap_off;     /* [n_haps+1]    haplotype -> first base in hap_seq */
    const char *hap_seq;        /* [hap_off[n_haps]] */
    const int64_t *read_off;    /* [n_reads+1]   read -> first base in read_seq/read_qual */
    const char *read_seq;       /* [read_off[n_reads]] */
    const uint8_t *read_qual;   /* [read_off[n_reads]] phred */
    const uint8_t *read_mapq;   /* [n_reads] phred mapping quality (bam core.qual) */
    const uint32_t *read_start; /* [n_reads] uint32_t(read.posStat.first) (cpp:52) */
    const uint8_t *read_flags;  /* [n_reads] DINDEL_READ_* */
    /* Insert-size prior of computeBMidPrior (ObservationModelFB.cpp:279-293), used when the pipeline runs with
     * --libFile (params.mapUnmappedReads, DInDel.cpp:4268-4272).  All three NULL = no read has one.  Per read:
     *   read_ins_mode   DINDEL_INS_NONE unless mapUnmappedReads && read.isPaired() && !mateIsUnmapped() &&
     *                   mateLen != -1 && core.tid == core.mtid (cpp:279,283); then DINDEL_INS_MATE_REVERSE when
     *                   mateIsReverse() (cpp:284-285) else DINDEL_INS_MATE_FORWARD (cpp:286-287)
     *   read_ins_anchor int(matePos + mateLen) for MATE_REVERSE, int(matePos) for MATE_FORWARD
     *   read_lib        index of the read's Library among those given to dindel_b200_set_libraries */
    const uint8_t *read_ins_mode;
    const int32_t *read_ins_anchor;
    const int32_t *read_lib;
} dindel_batch_in;
#define DINDEL_INS_NONE 0
#define DINDEL_INS_MATE_REVERSE 1
#define DINDEL_INS_MATE_FORWARD 2

/* Flat image of one MLAlignment (MLAlignment.hpp:50-70), 64 bytes.  The AlignedVariant maps
 * (indels/snps/hapIndelCovered/hapSNPCovered) are pure functions of hpos / firstBase / lastBase
 * and the sequences; the host adapter rebuilds them on demand. */
typedef struct dindel_pair_out {
    double ll, llOn, llOff, mLogBQ;
    int32_t firstBase, lastBase;
    uint16_t numIndels, numMismatch, nBQT, nmmBQT, nMMLeft, nMMRight;
    uint8_t offHap, offHapHMQ;
    int16_t bMid;
    uint8_t status; /* DINDEL_WIN_* of this pair */
    uint8_t pad0;
    uint16_t mapState0; /* Viterbi state of read base 0 (x, or S+x in an insertion; S = H+2): with hpos it
                           determines every state, so the AlignedVariant maps can be rebuilt exactly */
    uint8_t pad[4];
} dindel_pair_out;

/* dindel_batch_out.flags: produce the hpos rows and keep them resident in HBM for
 * dindel_b200_coverage even when out->hpos == NULL (they then never cross PCIe). */
#define DINDEL_OUT_KEEP_HPOS 1

/* Results, caller-allocated.  Pair (w, h, r) lives at win_pair[w] + r*nh(w) + h  (r-major, like
 * rl[r*nh+h] at DInDel.cpp:2129).  win_pair is filled by the library ([n_windows+1]). */
typedef struct dindel_batch_out {
    int64_t n_pairs;            /* capacity of pairs[] (>= dindel_b200_count_pairs) */
    dindel_pair_out *pairs;     /* [n_pairs] */
    int16_t *hpos;              /* optional [n_pairs * hpos_stride]: MLAlignment::hpos */
    int32_t hpos_stride;        /* >= longest read when hpos != NULL; row entries past a read's length are unspecified */
    int32_t flags;              /* DINDEL_OUT_* */
    int64_t *win_pair;          /* [n_windows+1] */
    int32_t *win_status;        /* [n_windows] DINDEL_WIN_* (first failing pair's reason) */
    uint8_t *read_on_hap;       /* optional [n_reads]: onHap[r] (DInDel.cpp:1720) */
} dindel_batch_out;

/* timings of the last submit/run on the context's stream, CUDA events, milliseconds */
typedef struct dindel_timing {
    /* submit / process: h2d_ms / d2h_ms = busy time of the two copy streams (summed over chunks; they run under the
     * kernels), k1_ms = span from the first chunk's first kernel to the last chunk's last kernel, prep_ms = 0,
     * total_ms = first copy in to last copy out.  upload / run / download: the four stages one after the other. */
    float h2d_ms, prep_ms, k1_ms, d2h_ms, total_ms;
    int32_t k1_launches, prep_launches;
    int64_t n_pairs, n_cells; /* cells = sum over pairs R*(H+2) */
    float post_ms;            /* process: K4 / K2 / K3 stages incl. their copies, summed over ranges */
    float host_ms;            /* process: host wall time of the whole call (pack / index / scan + waiting) */
    int32_t n_ranges;         /* process: resident ranges the batch was streamed through (1 = whole batch resident) */
    int32_t n_chunks;         /* pipeline chunks over all ranges */
} dindel_timing;

typedef struct dindel_ctx dindel_ctx;

void dindel_b200_default_params(dindel_params *p);
void dindel_b200_cli_params(dindel_params *p);
int dindel_b200_abi_version(void);
const char *dindel_b200_strerror(int code);
const char *dindel_b200_last_error(const dindel_ctx *ctx);

/* one context = one device + one stream; use from one host thread at a time */
int dindel_b200_create(const dindel_params *p, int device, dindel_ctx **out);
int dindel_b200_destroy(dindel_ctx *ctx);

/* Insert-size libraries (Library.hpp:31-129): library l holds Library::probs (normalised, floored at 1e-10,
 * maxins = lib_off[l+1]-lib_off[l] entries; getProb clamps |x| to maxins-1, Library.hpp:61-66) and its
 * getNinetyFifthPctProb().  The library takes log() of them with the host libm, as cpp:285-289 does per call.
 * Replaces the set given before; needed only when some read has read_ins_mode != DINDEL_INS_NONE. */
int dindel_b200_set_libraries(dindel_ctx *ctx, int n_libs, const int64_t *lib_off, const double *probs,
                              const double *ninetyfifth_pct_prob);

int64_t dindel_b200_count_pairs(const dindel_batch_in *in);
int64_t dindel_b200_count_cells(const dindel_batch_in *in);

/* Window scheduler, host only (no device needed): split the batch into n_shards contiguous window
 * ranges of near-equal K1 cost (cells = sum R*(H+2) per window), one per GPU / process.  Windows are
 * independent (DInDel.cpp:1310-1411), so shards need no collective; this replaces the per-file
 * window split of makeWindows.py (python/makeWindows.py:48-55).  first_window[n_shards+1] receives
 * the range boundaries: shard k owns windows [first_window[k], first_window[k+1]). */
int dindel_b200_plan_shards(const dindel_batch_in *in, int n_shards, int32_t *first_window);

/* computeLikelihoods for a whole batch: host buffers in, host buffers out (H2D + kernels + D2H).
 * Equivalent to dindel_b200_process with only `in` / `out` set. */
int dindel_b200_submit(dindel_ctx *ctx, const dindel_batch_in *in, dindel_batch_out *out);

/* the same split in three for callers that keep a batch resident in HBM */
int dindel_b200_upload(dindel_ctx *ctx, const dindel_batch_in *in, int want_hpos);
int dindel_b200_run(dindel_ctx *ctx);                         /* kernels only, device resident */
int dindel_b200_download(dindel_ctx *ctx, dindel_batch_out *out);
int dindel_b200_sync(dindel_ctx *ctx);                        /* wait for run(); updates timing */
int dindel_b200_get_timing(const dindel_ctx *ctx, dindel_timing *t);
int dindel_b200_hpos_stride(const dindel_ctx *ctx);           /* stride of the resident hpos rows */
/* device time (CUDA events) of the kernel of the last pair_sums / bayes_em / coverage call, ms */
float dindel_b200_last_kernel_ms(const dindel_ctx *ctx);

/* Inference only: makes a batch of caller-computed read x haplotype log-likelihoods the resident batch, as if K1 had
 * produced them (rl[r*nh+h] of DInDel.cpp:2129 / :2953, r-major inside a window, windows back to back), so that
 * dindel_b200_pair_sums / dindel_b200_bayes_em run on likelihoods from elsewhere (the reference's --faster model, a
 * dump).  Every other field of the resident pair records is zero and there are no hpos rows. */
int dindel_b200_load_likelihoods(dindel_ctx *ctx, int32_t n_windows, const int32_t *win_hap, const int32_t *win_read,
                                 const double *ll);
/* The same with whole pair records (r-major per window, like dindel_batch_out.pairs) and the per-read bytes the stage
 * kernels read (mapping-quality phred, DINDEL_READ_* flags; NULL = zeros): K2 / K3 / K6 then run on records that were
 * computed elsewhere (tests load the reference's own MLAlignment fields this way). */
int dindel_b200_load_records(dindel_ctx *ctx, int32_t n_windows, const int32_t *win_hap, const int32_t *win_read,
                             const dindel_pair_out *records, const uint8_t *read_mapq, const uint8_t *read_flags);

/* ---- K2: haplotype-pair sums over the RESIDENT likelihood matrix (after run/submit) ----------
 * Replaces the inner loops of DetInDel::diploidGLF (DInDel.cpp:3085-3113: start = 0;
 * :3349-3390: start = prior - sitePrior) and the pooled genotype likelihoods (DInDel.cpp:2683-2694:
 * reads of one pool member).  Job j on window job_window[j]:
 *   out[job_off[j] + q] = start[job_off[j] + q] (+)= log(0.5) + addLogs(rl[r,h1], rl[r,h2])
 * summed over the window's reads r in order (only those with read_group[r] == job_group[j] when
 * job_group[j] >= 0); q enumerates h1 <= h2 row-major: (0,0),(0,1)..(0,nh-1),(1,1).. */
typedef struct dindel_pairsum_in {
    int32_t n_jobs;
    int32_t reserved;
    const int32_t *job_window; /* [n_jobs] */
    const int32_t *job_group;  /* [n_jobs] -1 = all reads */
    const int64_t *job_off;    /* [n_jobs+1]; job_off[j+1]-job_off[j] == nh*(nh+1)/2 of its window */
    const double *start;       /* [job_off[n_jobs]]; NULL = all zero */
    const int32_t *read_group; /* optional [n_reads] (Read::poolID) */
} dindel_pairsum_in;
int dindel_b200_pair_sums(dindel_ctx *ctx, const dindel_pairsum_in *in, double *out);
/* Bound on |GPU sum - the reference's CPU sum| for a sum of n_terms reads whose partial sums stay below `magnitude`
 * in absolute value (pass |start| + |sum| + 1): the terms differ by CUDA-libm-vs-glibc exp/log (<= 2 ulp of a term
 * plus 1e-15 for the log(1+exp()) part) and every partial sum rounds once, in either order of summation, so
 *   bound = 4 * n_terms * (2.22e-16 * magnitude + 1e-15)          (a factor 4 of head-room on the analysis).
 * The host adapter (include/dindel_b200.hpp) treats every comparison closer than this and every printed field whose
 * 6-digit text could change within it as undecided and re-evaluates that window with the reference's own CPU loop. */
double dindel_b200_sum_error_bound(int64_t n_terms, double magnitude);

/* ---- K3: variational-Bayes EM over haplotype frequencies (DInDel.cpp:2431-2545) --------------
 * One EM run per set of active variants: set s on window set_window[s] with compatible[] flags
 * (DInDel.cpp:2395-2411) for its nh haplotypes at set_off[s].  Outputs per set: loglik (logliks[th],
 * :2542), iterations, and per haplotype pi (log, :2473) and freqs = exp(pi)/sum (:2539). */
typedef struct dindel_em_in {
    int32_t n_sets;
    int32_t reserved;
    const int32_t *set_window;  /* [n_sets] */
    const int64_t *set_off;     /* [n_sets+1] */
    const uint8_t *compatible;  /* [set_off[n_sets]] */
    double a0;                  /* params.bayesa0 (CLI default 0.001, DInDel.cpp:4142) */
    double tol;                 /* params.EMtol  (1e-4, DInDel.hpp:244) */
} dindel_em_in;
typedef struct dindel_em_out {
    double *loglik;  /* [n_sets] */
    int32_t *iters;  /* [n_sets] */
    double *pi;      /* [set_off[n_sets]] */
    double *freqs;   /* [set_off[n_sets]] */
    /* optional [n_sets]: min over the EM passes of | |eOld - eNew| - tol |, the distance of the stopping rule
     * (DInDel.cpp:2515) from flipping.  The GPU's energy differs from the CPU's at the 1e-10 level (libm, reduction
     * order over reads); a set whose margin is below the adapter's guard is re-run with the reference's own loop. */
    double *margin;
    double *energy;  /* optional [n_sets]: eNew of the last pass */
} dindel_em_out;
int dindel_b200_bayes_em(dindel_ctx *ctx, const dindel_em_in *in, dindel_em_out *out);

/* ---- K4: haplotype filter + per-variant read coverage over the RESIDENT hpos rows -------------
 * Replaces DetInDel::filterHaplotypes (DInDel.cpp:1932-2100), the only consumer of MLAlignment::hpos
 * in the calling pipeline: with it the hpos rows (~200 B per pair, 85 % of the D2H bytes) stay in HBM.
 * The host lists, per haplotype and in the map order of Haplotype::indels, one ENTRY per variant
 * (DInDel.cpp:1962): type, the read-flank interval already widened by padCover
 * (left = getLeftFlankRead()-padCover, right = getRightFlankRead()+padCover, :1973-1974) and a KEY id
 * naming the (position, AlignedVariant) pair the entry stands for (`PAV`, :1943), shared by the
 * haplotypes of a window that carry the same variant.  Per entry the reads with !offHapHMQ and
 * numIndels == 0 (:1951) are tested: covered when every position of [left, right] is spanned by the read
 * with at most maxMismatch mismatches against the haplotype (hap 'N' exempt for deletions only,
 * :1997-2000, :2036-2041, :2047-2060).  A haplotype's entries after its first uncovered INS/DEL are
 * not evaluated (the `break` at :2071); filtered[h] = doFilter && an entry was uncovered (:2078-2083);
 * per key nf / nr = distinct covering reads on the forward / reverse strand over the unfiltered
 * haplotypes (:2089-2097).  The reference's stray read of hpos[b] at b == hpos.size() (loop bound `<=`
 * at :1995, :2034: heap slack) is "no hit" here, and a negative code (LO/RO/INS) never hits.
 * Needs a resident, computed batch with hpos rows (upload(want_hpos) or DINDEL_OUT_KEEP_HPOS). */
#define DINDEL_VAR_DEL 0
#define DINDEL_VAR_INS 1
#define DINDEL_VAR_OTHER 2 /* SNP / *REF entries of Haplotype::indels: key exists, no coverage test */
typedef struct dindel_cover_in {
    int32_t n_entries;          /* over the whole batch */
    int32_t n_keys;             /* over the whole batch */
    int32_t maxMismatch;        /* params.obsParams.maxMismatch (--flankMaxMismatch, DInDel.cpp:3949) */
    int32_t doFilter;           /* params.filterHaplotypes */
    const int64_t *win_entry;   /* [n_windows+1] window -> its entries, sorted by (haplotype, map order) */
    const int64_t *win_key;     /* [n_windows+1] window -> its keys */
    const int32_t *entry_hap;   /* [n_entries] haplotype index inside the window */
    const int32_t *entry_key;   /* [n_entries] key index inside the window */
    const uint8_t *entry_type;  /* [n_entries] DINDEL_VAR_* */
    const int32_t *entry_left;  /* [n_entries] */
    const int32_t *entry_right; /* [n_entries] */
    const uint8_t *read_strand; /* [n_reads] 1 = counted on the reverse strand: unmapped ? !mateIsReverse()
                                   : isReverse() (DInDel.cpp:1987-1992) */
} dindel_cover_in;
typedef struct dindel_cover_out {
    uint8_t *hap_filtered;      /* [n_haps] filtered[h] */
    uint8_t *hap_uncovered;     /* optional [n_haps] !allCovered, whatever doFilter says */
    int32_t *key_nf, *key_nr;   /* [n_keys] VariantCoverage (DInDel.hpp:53-66) */
} dindel_cover_out;
int dindel_b200_coverage(dindel_ctx *ctx, const dindel_cover_in *in, dindel_cover_out *out);

/* ---- K5: realigned-read CIGARs over the RESIDENT hpos rows (--outputRealignedBAM) --------------
 * Replaces the per-read loops at the two call sites of DetInDel::getCIGAR (DInDel.cpp:500-513 pooled,
 * :596-611 diploid) and getCIGAR itself (DInDel.cpp:728-882): with it the realigned-BAM option needs no hpos
 * row on the host either; 4 * (3 + cigar_cap) bytes per READ come back instead of 2 * R bytes per PAIR.
 * Per read r of window w:
 *   1. the haplotype it is realigned to.  DINDEL_CIGAR_PAIR (diploid, :597-606): the better of the window's MAP
 *      pair (win_h1[w], win_h2[w]) (getMaxHap, host); |ll1 - ll2| < 1e-8 -> h1 if hap_n_indels[h1] <
 *      hap_n_indels[h2] (Haplotype::countIndels, Haplotype.hpp:254-261) else h2.  DINDEL_CIGAR_BEST (pooled,
 *      :503-509): the first strict arg-max of ll over the window's haplotypes; reads with !onHap[r] are skipped
 *      (read_n_ops = 0, :502).  A window with win_h1[w] < 0 is skipped in PAIR mode (read_n_ops = 0).
 *   2. npos[b] = hap_ref_pos[hpos[b]] for hpos[b] >= 0, else hpos[b] (:745-747): hap_ref_pos is hap.ml.hpos, the
 *      haplotype's own alignment to the reference segment (ObservationModelSeqAn.hpp:170-233; >= 0 or a
 *      DINDEL_HPOS_* code), laid out like hap_seq (hap_off).
 *   3. the CIGAR state machine of :776-873, literally (soft clips either side of the first / last base with
 *      npos >= 0; M / I / D runs; the reference's silent no-op cases stay no-ops).
 * Results per read: read_hap (index inside the window), read_ref_pos (CIGAR::refPos = win_ref_start[w] + npos of
 * the first aligned base, :798; -1 for a wholly clipped read, where the reference leaves it unset), read_n_ops
 * (number of operations; the first min(n, cigar_cap) are stored; negative = the reference's throw:
 * -1..-4 "Error(n)!", -5 "How is this possible? (1)"), ops[r * cigar_cap + k] = len << 4 | op (BAM encoding,
 * what writeRealignedBAMFile stores at :695).  The reference's scan for the last aligned base (:779) runs off
 * the front of the vector when no base is aligned; that case is the whole-read soft clip of :782-786 here.
 * Needs a resident, computed batch with hpos rows (upload(want_hpos) or DINDEL_OUT_KEEP_HPOS). */
#define DINDEL_CIGAR_PAIR 0
#define DINDEL_CIGAR_BEST 1
typedef struct dindel_cigar_in {
    int32_t mode;                  /* DINDEL_CIGAR_* */
    int32_t cigar_cap;             /* operations per read the ops rows hold (>= 1) */
    const int32_t *win_h1, *win_h2;    /* [n_windows] PAIR mode only */
    const int32_t *win_ref_start;  /* [n_windows] refSeqPos (= leftPos, DInDel.cpp:401) */
    const int32_t *hap_n_indels;   /* [n_haps] PAIR mode only */
    const int32_t *hap_ref_pos;    /* [hap_off[n_haps]] */
} dindel_cigar_in;
typedef struct dindel_cigar_out {
    int32_t *read_hap;      /* [n_reads] */
    int32_t *read_ref_pos;  /* [n_reads] */
    int32_t *read_n_ops;    /* [n_reads] */
    uint32_t *ops;          /* [n_reads * cigar_cap] */
} dindel_cigar_out;
int dindel_b200_cigars(dindel_ctx *ctx, const dindel_cigar_in *in, dindel_cigar_out *out);

/* ---- K6: read statistics of a called site over the RESIDENT K1 records ---------------------------
 * Replaces the per-read aggregation loops that follow the arg-max in DetInDel::diploidGLF and in
 * DetInDel::estimateHaplotypeFrequenciesBayesEM: everything they read of liks[h][r] is a field of the 64-byte pair
 * record (hapIndelCovered / hapSNPCovered[pos] are AlignedVariant::isCovered(padCover, firstBase, lastBase),
 * ObservationModelFB.cpp:1465-1472 = firstBase + pad <= startRead && lastBase - pad >= endRead, Variant.hpp:125-128),
 * everything they read of reads[r] is resident too (mapping-quality phred, DINDEL_READ_* flags).
 * One JOB = one (window, site) with its CANDIDATE haplotypes; per candidate the host gives the window-relative
 * haplotype index, startRead / endRead of the haplotype's variant at the site and DINDEL_CAND_PRESENT when the
 * coverage map the reference consults has the site at all (`find(pos) != end()`), DINDEL_CAND_HAS_VAR = hapHasVar.
 * With mq(r) = -10*log10(1 - mapQual(r)) (host libm, by phred byte) and reads taken in window order, restricted to
 * read_group[r] == job_group when job_group >= 0:
 *   DINDEL_STATS_SITE (DInDel.cpp:3396-3404, :3504-3554; candidates = the site's arg-max pair hx1, hx2):
 *     numUnmappedRealigned = #unmapped reads with !offHap on hx1 or hx2; allmsq += mq^2; numOffBoth = #reads offHap on
 *     both; h = hx1 if ll[hx1][r] >= ll[hx2][r] else hx2; numIndels / nBQT / nmmBQT / mLogBQ summed over the chosen
 *     records, nMMLeft / nMMRight = #reads whose chosen record has >= 2; when the chosen record covers the site:
 *     nf or nr (Read::onReverseStrand) + 1, msq += mq^2, n + 1.
 *   DINDEL_STATS_MAP (:3126-3132, :3148-3170; candidates = max_indel_pair[0..1], PRESENT = it is one of the m
 *     haplotypes looped over and has an indel at the site): numUnmappedRealigned as above; for every PRESENT candidate
 *     in turn, for every read covering the site on it: nf or nr + 1, msq += mq^2, n + 1.
 *   DINDEL_STATS_POOL (:2697-2742; candidates = all haplotypes of the window in order): per read ml = max ll,
 *     the haplotypes with ll >= ml - 1e-7 that are PRESENT, cover the site and HAS_VAR set the read's strand flag;
 *     nf / nr = #reads with a forward / reverse flag; msq += mq^2 and n + 1 for every read.
 * The FP64 sums run in the reference's order (read order; MAP: candidate-major), so msq / allmsq / mLogBQ are
 * bit-identical to the CPU's; the caller applies sqrt(msq / n) etc.  n_reads = |readidx|.
 * Needs a resident, computed batch (the job windows inside the resident range). */
#define DINDEL_STATS_SITE 0
#define DINDEL_STATS_MAP 1
#define DINDEL_STATS_POOL 2
#define DINDEL_CAND_PRESENT 1u
#define DINDEL_CAND_HAS_VAR 2u
typedef struct dindel_stats_in {
    int32_t n_jobs;
    int32_t reserved;
    const int32_t *job_window;  /* [n_jobs] */
    const int32_t *job_mode;    /* [n_jobs] DINDEL_STATS_* */
    const int32_t *job_group;   /* [n_jobs] -1 = every read of the window */
    const int64_t *job_cand;    /* [n_jobs+1] job -> its candidates (SITE / MAP: exactly 2) */
    const int32_t *cand_hap;    /* haplotype index inside the window */
    const int32_t *cand_start;  /* AlignedVariant::startRead of the haplotype's variant at the site */
    const int32_t *cand_end;    /* AlignedVariant::endRead */
    const uint8_t *cand_flags;  /* DINDEL_CAND_* */
    const int32_t *read_group;  /* [n_reads] of the batch (Read::poolID); may be NULL when every job_group is -1 */
} dindel_stats_in;
typedef struct dindel_read_stats {
    double msq, allmsq, mLogBQ;   /* plain sums (no sqrt, no division) */
    int32_t n, nf, nr, numOffBoth, numIndels, nBQT, nmmBQT, nMMLeft, nMMRight, numUnmappedRealigned, n_reads, reserved;
} dindel_read_stats;
int dindel_b200_read_stats(dindel_ctx *ctx, const dindel_stats_in *in, dindel_read_stats *out /* [n_jobs] */);

/* ---- one pass over a batch of ANY size: K1 and, while the results are still in HBM, K4 / K2 / K3 --------------
 * The batch is streamed through the device in RANGES of windows that fit a fixed memory budget
 * ($DINDEL_B200_MEM_BUDGET bytes, default 48 GiB: inputs + 64 B per pair + the hpos rows); inside a range the
 * chunked H2D / kernel / D2H pipeline of submit runs, then the optional stages on the range's resident results, and
 * their (small) outputs come back before the next range reuses the buffers.  Device memory therefore does not grow
 * with the batch: one call can take millions of windows (BASELINE configs[4]).  What replaces what:
 *   out        computeLikelihoods (DInDel.cpp:1707-1739).  out->pairs may be NULL when only the stages below are
 *              wanted (then no per-pair record crosses PCIe); out->win_pair / win_status / read_on_hap as in submit.
 *   cover      filterHaplotypes (DInDel.cpp:1932-2100) for every window: same arrays as dindel_b200_coverage, over
 *              the whole batch.  Implies the hpos rows are produced (and kept in HBM unless out->hpos is set).
 *   pair_sums  the read sums of diploidGLF / the pooled genotype likelihoods (DInDel.cpp:3086-3089, :3369-3372,
 *              :2687-2690) for every window and every h1 <= h2, started at 0:
 *                n_groups == 0: pair_sums[sum_off[w] + q]                      = sum over all reads of the window
 *                n_groups >= 1: pair_sums[sum_off[w] * n_groups + g * npairs(w) + q] = sum over reads with read_group == g
 *              with sum_off[w] = sum over windows before w of nh(nh+1)/2 (returned in win_sum_off when non-NULL) and q
 *              enumerating h1 <= h2 row-major.  The host adds its own start value (prior - sitePrior) and keeps the
 *              printed text identical through dindel_b200_sum_error_bound (see there).
 *   em         estimateHaplotypeFrequenciesBayesEM's EM (DInDel.cpp:2431-2545) for sets known up front; set_window
 *              must be sorted ascending.
 * After the call the batch is resident (K2..K5 entries usable) only if it went through as ONE range
 * (dindel_timing.n_ranges == 1). */
typedef struct dindel_job {
    const dindel_batch_in *in;
    dindel_batch_out *out;
    const dindel_cover_in *cover;         /* optional */
    dindel_cover_out *cover_out;
    double *pair_sums;                     /* optional */
    int64_t *win_sum_off;                  /* optional [n_windows+1] */
    int32_t n_groups;
    int32_t reserved;
    const int32_t *read_group;             /* [n_reads] when n_groups >= 1 */
    const dindel_em_in *em;         /* optional */
    dindel_em_out *em_out;
} dindel_job;
int dindel_b200_process(dindel_ctx *ctx, const dindel_job *job);
/* number of doubles pair_sums must hold for this batch (per group when n_groups >= 1: multiply) */
int64_t dindel_b200_count_sums(const dindel_batch_in *in);

/* ---- the window scheduler over the GPUs of one box (north star (3); replaces the per-file process fan-out of
 * makeWindows.py, python/makeWindows.py:48-55) ----------------------------------------------------------------
 * One dindel_multi owns one context and one host worker thread per listed device.  multi_process plans contiguous
 * window shards of near-equal K1 cost (dindel_b200_plan_shards), runs dindel_b200_process on every shard
 * concurrently (each device streams its own shard through its own ranges) and returns when all are done; results
 * land in the caller's arrays in window order, byte-identical to a single-device run.  No collective: windows are
 * independent (DInDel.cpp:1310-1411).  Use from one host thread at a time. */
typedef struct dindel_multi dindel_multi;
int dindel_b200_multi_create(const dindel_params *p, const int *devices, int ndev, dindel_multi **out);
int dindel_b200_multi_destroy(dindel_multi *m);
int dindel_b200_multi_devices(const dindel_multi *m);
int dindel_b200_multi_set_libraries(dindel_multi *m, int n_libs, const int64_t *lib_off, const double *probs,
                                    const double *ninetyfifth_pct_prob);
int dindel_b200_multi_process(dindel_multi *m, const dindel_job *job);
/* per device of the last multi_process: its window shard [first, last) and its timing */
int dindel_b200_multi_get_shard(const dindel_multi *m, int k, int32_t *first_window, int32_t *last_window,
                                dindel_timing *t);
const char *dindel_b200_multi_last_error(const dindel_multi *m);

/* pinned host memory for the caller's buffers */
void *dindel_b200_alloc_pinned(size_t bytes);
void dindel_b200_free_pinned(void *p);

/* FP64 add-issue peak of the context's device (ops/s), register-resident DADD microbenchmark:
 * the denominator of the K1 roofline. */
int dindel_b200_measure_fp64_peak(dindel_ctx *ctx, double *dadd_per_s);

/* K1 over the resident batch WITHOUT ANY ALLOCATION.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Inputs: the batch made resident by dindel_b200_upload (which may allocate: set-up).
 * Outputs: `out` as for dindel_b200_download, or NULL to leave the results resident (dindel_b200_download later).
 * The call itself issues one memset, the generic K1 kernel with its back-pointer rows in global memory, the window
 * summary and (out != NULL) the device-to-host copies; it calls no cudaMalloc / cudaFree / allocating routine.  The
 * kernel's scratch (work-queue head + one back-pointer slot per resident warp) is a file-scope device array of
 * DINDEL_B200_STATIC_SCRATCH_BYTES that every GPU loading the library holds until the process exits.
 * Limit: align256((maxH + 2) * maxR) * 4 warps * SM count <= DINDEL_B200_STATIC_SCRATCH_BYTES - 256 (B200, 148 SMs:
 * (maxH + 2) * maxR <= 28 160, e.g. 100-base reads on haplotypes up to 279 bases); a larger batch is refused with
 * DINDEL_E_UNSUPPORTED and a message naming both sizes -- use dindel_b200_run.  Results are byte-identical to
 * dindel_b200_run's.  All calls on a device share the array, which keeps the last call's contents (the queue head is
 * cleared by the call; the slots need no clearing): calls that OVERLAP on the GPU (two contexts, streams or threads
 * on one device) race -- issue one at a time per device. */
#define DINDEL_B200_STATIC_SCRATCH_BYTES (16u << 20)
int dindel_b200_run_static(dindel_ctx *ctx, dindel_batch_out *out);
size_t dindel_b200_static_scratch_bytes(void);

#ifdef __cplusplus
}
#endif
#endif
