/* oracle/dindel_oracle.h — TEST INFRASTRUCTURE ONLY (see dindel_oracle.c). */
#ifndef DINDEL_ORACLE_H
#define DINDEL_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* The subset of ObservationModelParameters (ObservationModel.hpp:28-99) the path reads. */
typedef struct ora_params {
    double pError, pMut, mapQualThreshold, pFirstgLO, checkBaseQualThreshold;
    int32_t maxLengthDel, padCover, bMid, forceReadOnHaplotype;
} ora_params;

/* Flat image of the MLAlignment fields (MLAlignment.hpp:50-70) a pair produces. */
typedef struct ora_pair_out {
    double ll, llOn, llOff, mLogBQ;
    int32_t firstBase, lastBase;
    int32_t numIndels, numMismatch, nBQT, nmmBQT, nMMLeft, nMMRight;
    int32_t offHap, offHapHMQ;
    int32_t bMid, status; /* status: 0 ok, 1 "hapSize error." */
} ora_pair_out;

int ora_sizeof_params(void);
int ora_sizeof_pair_out(void);
double ora_phred_to_prob(int phred);

int ora_pair(const char *hap, int H, const char *read, int R, const double *qual, double mapQual,
             double posStatFirst, int unmapped, uint32_t hapStart, const ora_params *p,
             ora_pair_out *out, int32_t *hpos, int32_t *mapState);

int ora_pair_ins(const char *hap, int H, const char *read, int R, const double *qual, double mapQual,
                 double posStatFirst, int unmapped, uint32_t hapStart, const ora_params *p,
                 ora_pair_out *out, int32_t *hpos, int32_t *mapState, int ins_mode, int ins_anchor,
                 const double *lib_probs, int maxins, double p95);

int ora_window(const char *hap_seq, const int32_t *hap_off, int nh, const char *read_seq,
               const uint8_t *read_qual, const int32_t *read_off, const uint8_t *mapq,
               const uint32_t *mReadStart, const uint8_t *flags, int nr, uint32_t hapStart,
               const ora_params *p, ora_pair_out *out, int32_t *hpos, int hpos_stride);

void ora_pair_sums(const double *rl, int nr, int nh, const int32_t *read_group, int group,
                   const double *start, double *out);
int ora_bayes_em(const double *rl, int nr, int nh, const uint8_t *compatible, double a0, double tol,
                 double *loglik_out, double *pi_out, double *freqs_out);
void ora_filter_haplotypes(int nh, int nr, const char *hap_seq, const int32_t *hap_off, const char *read_seq,
                           const int32_t *read_off, const int32_t *hpos, int hpos_stride,
                           const uint8_t *offHapHMQ, const int32_t *numIndels, const uint8_t *read_strand,
                           int ne, const int32_t *entry_hap, const int32_t *entry_key, const uint8_t *entry_type,
                           const int32_t *entry_left, const int32_t *entry_right, int nk, int maxMismatch,
                           int doFilter, uint8_t *filtered, uint8_t *uncovered, int32_t *key_nf, int32_t *key_nr);
int ora_choose_hap(int mode, int nh, const double *ll, const int32_t *n_indels, int h1, int h2);
int ora_get_cigar(int R, const int32_t *hpos, int H, const int32_t *hml, int refSeqStart, int32_t *refPos,
                  int32_t *ops, int cap);
/* K6: read statistics of a called site (DInDel.cpp:3126-3170, :3396-3404, :3504-3554, :2697-2742) */
typedef struct ora_read_stats_t {
    double msq, allmsq, mLogBQ;
    int32_t n, nf, nr, numOffBoth, numIndels, nBQT, nmmBQT, nMMLeft, nMMRight, numUnmappedRealigned, n_reads, reserved;
} ora_read_stats_t;
void ora_read_stats(int mode, int nh, int nr, const ora_pair_out *recs, const double *mapQual, const uint8_t *reverse,
                    const uint8_t *unmapped, const int32_t *read_group, int group, int ncand, const int32_t *cand_hap,
                    const int32_t *cand_start, const int32_t *cand_end, const uint8_t *cand_flags, int pad,
                    ora_read_stats_t *o);
double ora_addlogs(double a, double b);
double ora_digamma(double x);

#ifdef __cplusplus
}
#endif
#endif
