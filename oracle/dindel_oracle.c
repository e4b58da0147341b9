/* oracle/dindel_oracle.c — TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C CPU restatement of Dindel's read x candidate-haplotype likelihood path
 * (ObservationModelFBMaxErr, a two-sided max-product HMM) and, further down, of the numeric cores
 * of the diploid marginalisation and the Bayesian EM that consume its output.  It is the checker
 * for the CUDA path: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may
 * load it; the product library never links or calls it.
 *
 * Parity status: PINNED for K1.  tests/test_oracle.py checks this file against
 *   (1) the four smain() known-answer fixtures (DInDel.cpp:3993-4066, values in SURVEY.md 8c), and
 *   (2) the reference's own ObservationModelFB.cpp compiled here into oracle/_ref (bit-equal ll /
 *       llOn / llOff / mLogBQ, identical Viterbi path, hpos and counters on randomised pairs), and
 *   (3) golden vectors produced by that compiled reference, committed under tests/golden/.
 * K2 (diploid) is a restatement of libm-only arithmetic; K3's digamma is a third-party function
 * (Boost.Math, version unpinned by the reference) => "parity unpinned" for that one function.
 *
 * Every function cites the reference lines (relative to /root/reference) it follows.
 */
#include "dindel_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

int ora_sizeof_params(void) { return (int)sizeof(ora_params); }
int ora_sizeof_pair_out(void) { return (int)sizeof(ora_pair_out); }

/* Read.hpp:127-131 (mapping quality) and :143-148 (base quality): phred byte -> probability. */
double ora_phred_to_prob(int phred)
{
    double ph = (double)phred;
    double q = (1.0 - pow(10.0, -ph / 10.0));
    if (q < 1e-16) q = 1e-16;
    if (q > 1.0 - 1e-16) q = 1.0 - 1e-16;
    return q;
}

/* ReadIndelErrorModel.hpp:28-50: homopolymer run length -> per-base indel error probability. */
static double hp_error(int hpLen)
{
    static const double base[10] = {2.9e-5, 2.9e-5, 2.9e-5, 2.9e-5, 4.3e-5,
                                    1.1e-4, 2.4e-4, 5.7e-4, 1.0e-3, 1.4e-3};
    int len = hpLen < 1 ? 1 : hpLen;
    double pbe = (len <= 10) ? base[len - 1] : base[9] + 4.3e-4 * (double)(len - 10);
    pbe *= (double)hpLen;
    if (pbe > 0.99) pbe = 0.99;
    return pbe;
}

/* ObservationModelFB.cpp:877-888 (EPS = 1e-10, ObservationModelFB.hpp:25). */
static inline void upd(double *dv, int *di, double nv, int ni)
{
    if (nv > *dv + 1e-10) {
        *dv = nv;
        *di = ni;
    } else if (nv >= *dv && nv <= *dv + 1e-5 && *di > ni) {
        *dv = nv;
        *di = ni;
    }
}

typedef struct {
    int H, R, S, RO, T; /* S = numS = H+2, T = numT = maxLengthDel+2 */
    double lLOgLO, lFirstgLO, lInsgIns, lNoInsgIns, lInsgNoIns, lNoInsgNoIns;
    double *lpe, *lpn; /* logProbError / logProbNoError, H+2 each */
} hmm_t;

/* ObservationModelFB.cpp:1641-1713: transition constants and the homopolymer tables. */
static void setup_transitions(hmm_t *m, const char *hap, const ora_params *p)
{
    int H = m->H, b, len;
    double perr;
    m->lLOgLO = log(1.0 - p->pFirstgLO);
    m->lFirstgLO = log(p->pFirstgLO);
    m->T = p->maxLengthDel + 2;
    m->lInsgIns = -.5;
    m->lNoInsgIns = log(1.0 - exp(m->lInsgIns));
    m->lInsgNoIns = log(p->pError);
    m->lNoInsgNoIns = log(1 - p->pError);
    for (b = 0; b < H + 2; b++) {
        m->lpe[b] = log(1e-5);
        m->lpn[b] = log(1 - 1e-5);
    }
    len = 1;
    perr = hp_error(1);
    m->lpe[1] = log(perr);
    m->lpn[1] = log(1.0 - perr);
    for (b = 1; b < H; b++) {
        if (hap[b] == hap[b - 1]) {
            len++;
        } else {
            perr = hp_error(len);
            m->lpe[b] = log(perr);
            m->lpn[b] = log(1.0 - perr);
            len = 1;
        }
    }
    perr = hp_error(len);
    m->lpe[H - 1] = log(perr);
    m->lpn[H - 1] = log(1.0 - perr);
}

/* ObservationModelFB.cpp:1775-1829: one left-half step, row b from row b-1. */
static void step_dec(const hmm_t *m, double *A, const double *P, const double *O, int *bt)
{
    int H = m->H, S = m->S, RO = m->RO, T = m->T, x, y;
    A[RO] = -HUGE_VAL;
    upd(&A[RO], &bt[RO], O[RO] + P[RO] + m->lLOgLO + m->lNoInsgNoIns, RO);
    upd(&A[RO], &bt[RO], O[H] + P[H] + m->lFirstgLO + m->lNoInsgNoIns, H);
    for (x = 1; x <= H; x++) {
        double lpt = m->lpe[x], lpn = m->lpn[x];
        A[x] = -HUGE_VAL;
        for (y = 1; y < T; y++) {
            int nx = x - y;
            double lp;
            if (nx < 0) nx = 0;
            lp = (y == 1) ? lpn : (lpt + (double)(y - 1) * m->lInsgIns);
            upd(&A[x], &bt[x], O[nx] + lp + P[nx] + lpn, nx);
        }
    }
    A[0] = O[0] + P[0] + m->lNoInsgNoIns;
    bt[0] = 0;
    upd(&A[RO], &bt[RO], O[S + RO] + P[S + RO] + m->lLOgLO + m->lpe[RO], S + RO);
    upd(&A[RO], &bt[RO], O[S + H] + P[S + H] + m->lFirstgLO + m->lpe[H], S + H);
    for (x = 1; x <= H; x++) {
        int nx = x - 1;
        upd(&A[x], &bt[x], O[S + nx] + P[S + nx] + m->lpe[x], S + nx);
    }
    for (x = 0; x <= H + 1; x++) {
        A[S + x] = O[S + x] + P[S + x] + m->lInsgIns;
        bt[S + x] = S + x;
    }
    for (x = 1; x <= H + 1; x++) upd(&A[S + x], &bt[S + x], O[x] + P[x] + m->lNoInsgIns, x);
}

/* ObservationModelFB.cpp:1715-1773: one right-half step, row b-1 from row b. */
static void step_inc(const hmm_t *m, double *B, const double *P, const double *O, int *bt)
{
    int H = m->H, S = m->S, RO = m->RO, T = m->T, x, y;
    B[0] = -HUGE_VAL;
    upd(&B[0], &bt[0], O[0] + P[0] + m->lLOgLO + m->lNoInsgNoIns, 0);
    upd(&B[0], &bt[0], O[1] + P[1] + m->lFirstgLO + m->lNoInsgNoIns, 1);
    for (x = 1; x <= H; x++) {
        B[x] = -HUGE_VAL;
        for (y = 1; y < T; y++) {
            int nx = x + y;
            double lpn, lpt, lp;
            if (nx > H) nx = RO;
            lpn = m->lpn[nx];
            lpt = m->lpe[nx];
            lp = (y == 1) ? lpn : (lpt + (double)(y - 1) * m->lInsgIns);
            upd(&B[x], &bt[x], lp + lpn + P[nx] + O[nx], nx);
        }
    }
    B[RO] = -HUGE_VAL;
    upd(&B[RO], &bt[RO], O[RO] + P[RO] + m->lpn[RO], RO);
    for (x = 0; x <= H; x++) upd(&B[x], &bt[x], O[S + x] + P[S + x] + m->lpe[x + 1], S + x);
    x = H + 1;
    upd(&B[x], &bt[x], O[S + x] + P[S + x], S + x);
    for (x = 0; x <= H + 1; x++) {
        B[S + x] = O[S + x] + P[S + x] + m->lInsgIns;
        bt[S + x] = S + x;
    }
    upd(&B[S], &bt[S], O[0] + P[0] + m->lNoInsgIns, 0);
    for (x = 1; x <= H + 1; x++) {
        int nx = x + 1;
        if (nx > RO) nx = RO;
        upd(&B[S + x], &bt[S + x], O[nx] + P[nx] + m->lNoInsgIns, nx);
    }
}

/* ObservationModelFB.cpp:268-305 with pinsert == 0 (no --libFile). */
/* Insert-size term of one read (ObservationModelFB.cpp:279-293): mode 0 = none; 1 = mate on the reverse strand,
 * anchor = int(matePos + mateLen); 2 = mate forward, anchor = int(matePos).  probs = Library::probs (maxins
 * entries), p95 = getNinetyFifthPctProb(). */
typedef struct {
    int mode, anchor, maxins;
    const double *probs;
    double p95;
} ins_t;

/* ObservationModelFB.cpp:268-305. */
static void bmid_prior(const hmm_t *m, double *prior, double mapQual, double mapQualThreshold, const ins_t *ins,
                       int hapStart, int bMid)
{
    int S = m->S, H = m->H, RO = m->RO, R = m->R, i, x;
    double mq = 1.0 - mapQual, pOff;
    double *pinsert = (double *)calloc((size_t)S, sizeof(double));
    if (-10.0 * log10(mq) > mapQualThreshold) mq = pow(10.0, -mapQualThreshold / 10.0);
    pOff = mq;
    if (ins && ins->mode != 0) { /* :283-290; Library::getProb clamps (Library.hpp:61-66) */
        for (x = 1; x < H + 1; x++) {
            int d = (ins->mode == 1) ? abs(hapStart + x - bMid - ins->anchor) : abs(hapStart + x + R - bMid - ins->anchor);
            if (d < 0) d = -d;
            if (d >= ins->maxins) d = ins->maxins - 1;
            pinsert[x] = log(ins->probs[d]);
        }
        pinsert[0] = log(ins->p95);
    }
    for (i = 0; i < 2; i++) {
        double lIns = (i == 1) ? m->lInsgNoIns : log(1.0 - exp(m->lInsgNoIns));
        prior[i * S + 0] = log(pOff) + lIns + pinsert[0];
        prior[i * S + RO] = -100.0;
        for (x = 1; x < H + 1; x++) prior[i * S + x] = pinsert[x] + log((1.0 - pOff)) + lIns;
    }
    free(pinsert);
}

/* ObservationModelFB.cpp:35-102: the anchor base. */
static int compute_bmid(int H, int R, double posStatFirst, int unmapped, uint32_t hapStart,
                        const ora_params *p)
{
    uint32_t hapEnd = hapStart + (uint32_t)H;
    uint32_t mReadStart = (uint32_t)posStatFirst;
    uint32_t readEnd = mReadStart + (uint32_t)R - 1;
    int bMid;
    if (unmapped) {
        bMid = R / 2;
    } else if (mReadStart > hapEnd) {
        bMid = R / 2;
    } else if (readEnd < hapStart) {
        bMid = R / 2;
    } else {
        uint32_t olStart = (hapStart > mReadStart) ? hapStart : mReadStart;
        uint32_t olEnd = (hapEnd > readEnd) ? readEnd : hapEnd;
        int mid = ((int)olEnd - (int)olStart) / 2 + (int)olStart;
        bMid = mid - (int)mReadStart;
    }
    if (p->bMid != -1) bMid = p->bMid;
    if (bMid < 0) bMid = 0;
    if (bMid >= R) bMid = R - 1;
    return bMid;
}

static int pair_core(const char *hap, int H, const char *read, int R, const double *qual, double mapQual,
                     double posStatFirst, int unmapped, uint32_t hapStart, const ora_params *p,
                     ora_pair_out *out, int32_t *hpos_out, int32_t *mapState_out, const ins_t *ins)
{
    hmm_t m;
    int S = H + 2, W = 2 * S, RO = H + 1, b, x, s, bMid;
    double *obs, *alpha, *beta, *priorR, *priorH;
    int *btf, *btb, *ms;
    memset(out, 0, sizeof(*out));
    /* ObservationModelFB.cpp:47 */
    if (p->maxLengthDel > H) {
        out->status = 1;
        return 1;
    }
    bMid = compute_bmid(H, R, posStatFirst, unmapped, hapStart, p);
    m.H = H; m.R = R; m.S = S; m.RO = RO;
    m.lpe = (double *)malloc(sizeof(double) * S);
    m.lpn = (double *)malloc(sizeof(double) * S);
    setup_transitions(&m, hap, p);
    obs = (double *)malloc(sizeof(double) * (size_t)R * W);
    alpha = (double *)malloc(sizeof(double) * (size_t)R * W);
    beta = (double *)malloc(sizeof(double) * (size_t)R * W);
    btf = (int *)calloc((size_t)R * W, sizeof(int));
    btb = (int *)calloc((size_t)R * W, sizeof(int));
    ms = (int *)calloc(R, sizeof(int));
    priorR = (double *)malloc(sizeof(double) * W);
    priorH = (double *)malloc(sizeof(double) * W);

    /* ObservationModelFB.cpp:335-338 */
    for (x = 0; x < W; x++) {
        alpha[x] = 0.0;
        beta[(size_t)(R - 1) * W + x] = 0.0;
    }
    /* ObservationModelFB.cpp:220-266: emissions */
    for (b = 0; b < R; b++) {
        double *o = obs + (size_t)b * W;
        double pr = qual[b] * (1.0 - p->pMut);
        double eq = log(.25 + .75 * pr);
        double uq = log(.75 + 1e-10 - .75 * pr);
        char nuc = read[b];
        o[S + 0] = eq; o[S + H + 1] = eq; o[0] = eq; o[H + 1] = eq;
        for (x = 0; x < H; x++) {
            o[S + x + 1] = eq;
            o[x + 1] = (hap[x] == 'N' || hap[x] == nuc) ? eq : uq;
        }
        if (p->forceReadOnHaplotype) { /* :307-316 */
            o[0] = -1000.0; o[RO] = -1000.0; o[S] = -1000.0; o[RO + S] = -1000.0;
        }
    }
    /* ObservationModelFB.cpp:1569-1581 */
    for (b = 1; b <= bMid; b++)
        step_dec(&m, alpha + (size_t)b * W, alpha + (size_t)(b - 1) * W, obs + (size_t)(b - 1) * W,
                 btf + (size_t)b * W);
    for (b = R - 1; b > bMid; b--)
        step_inc(&m, beta + (size_t)(b - 1) * W, beta + (size_t)b * W, obs + (size_t)b * W,
                 btb + (size_t)(b - 1) * W);

    /* ObservationModelFB.cpp:1075-1144 */
    {
        const double *a = alpha + (size_t)bMid * W, *be = beta + (size_t)bMid * W,
                     *o = obs + (size_t)bMid * W;
        double ll = -HUGE_VAL, llHMQ = -HUGE_VAL, off = -HUGE_VAL, on = -HUGE_VAL;
        int msR = 0;
        bmid_prior(&m, priorR, mapQual, p->mapQualThreshold, ins, (int)hapStart, bMid);
        bmid_prior(&m, priorH, 1.0 - 1e-10, p->mapQualThreshold, ins, (int)hapStart, bMid);
        for (s = 0; s < W; s++) {
            double v = a[s] + o[s] + be[s] + priorR[s];
            if (v > ll + 1e-10) { ll = v; msR = s; }
            if ((s % S) == 0) { if (v > off) off = v; }
            else if ((s % S) != RO) { if (v > on) on = v; }
            v = a[s] + o[s] + be[s] + priorH[s];
            if (v > llHMQ + 1e-10) { llHMQ = v; ms[bMid] = s; }
        }
        out->ll = ll; out->llOff = off; out->llOn = on;
        out->offHapHMQ = ((ms[bMid] % S) == 0 || (ms[bMid] % S) == RO);
        out->offHap = ((msR % S) == 0 || (msR % S) == RO);
    }
    /* ObservationModelFB.cpp:1148-1165 */
    for (b = bMid; b > 0; b--) ms[b - 1] = btf[(size_t)b * W + ms[b]];
    for (b = bMid; b < R - 1; b++) ms[b + 1] = btb[(size_t)b * W + ms[b]];

    /* ObservationModelFB.cpp:1351-1463 (the integer/FP summaries; the std::map / std::string
     * AlignedVariant records are rebuilt on the host from hpos when a caller needs them) */
    {
        int first = -1, last = -1;
        out->nBQT = 0; out->nmmBQT = 0; out->mLogBQ = 0.0; out->nMMRight = 0; out->nMMLeft = 0;
        out->numIndels = 0; out->numMismatch = 0;
        b = 0;
        while (b < R) {
            int hp;
            s = ms[b];
            if ((s % S) > 0 && (s % S) <= H) {
                if (s >= S) {
                    while (b < R && ms[b] >= S) {
                        if (hpos_out) hpos_out[b] = -1;
                        b++;
                    }
                    out->numIndels++;
                    b--;
                } else {
                    hp = s - 1;
                    if (hpos_out) hpos_out[b] = hp;
                    if (first == -1) first = hp; else if (hp < first) first = hp;
                    if (last == -1) last = hp; else if (hp > last) last = hp;
                    if (qual[b] > p->checkBaseQualThreshold) {
                        out->nBQT++;
                        out->mLogBQ += log10(1.0 - qual[b]);
                    }
                    if (read[b] != hap[s - 1]) {
                        if (qual[b] > p->checkBaseQualThreshold) out->nmmBQT++;
                        if (b < 6) out->nMMLeft++;
                        if (b > R - 6) out->nMMRight++;
                        if (qual[b] > 0.95) out->numMismatch++;
                    }
                    if (b < R - 1) {
                        int ns = ms[b + 1];
                        if (ns < S && ns - s > 1) out->numIndels++;
                    }
                }
            } else {
                if (hpos_out) hpos_out[b] = ((s % S) == 0) ? -3 : -4;
            }
            b++;
        }
        out->firstBase = first;
        out->lastBase = last;
    }
    out->bMid = bMid;
    out->status = 0;
    if (mapState_out) for (b = 0; b < R; b++) mapState_out[b] = ms[b];
    free(m.lpe); free(m.lpn); free(obs); free(alpha); free(beta); free(btf); free(btb); free(ms);
    free(priorR); free(priorH);
    return 0;
}

int ora_pair(const char *hap, int H, const char *read, int R, const double *qual, double mapQual,
             double posStatFirst, int unmapped, uint32_t hapStart, const ora_params *p,
             ora_pair_out *out, int32_t *hpos, int32_t *mapState)
{
    return pair_core(hap, H, read, R, qual, mapQual, posStatFirst, unmapped, hapStart, p, out, hpos, mapState, 0);
}

/* The same with the insert-size prior of a paired read under --libFile (params.mapUnmappedReads). */
int ora_pair_ins(const char *hap, int H, const char *read, int R, const double *qual, double mapQual,
                 double posStatFirst, int unmapped, uint32_t hapStart, const ora_params *p,
                 ora_pair_out *out, int32_t *hpos, int32_t *mapState, int ins_mode, int ins_anchor,
                 const double *lib_probs, int maxins, double p95)
{
    ins_t ins;
    ins.mode = ins_mode; ins.anchor = ins_anchor; ins.maxins = maxins; ins.probs = lib_probs; ins.p95 = p95;
    return pair_core(hap, H, read, R, qual, mapQual, posStatFirst, unmapped, hapStart, p, out, hpos, mapState, &ins);
}

/* The double loop of DetInDel::computeLikelihoods (DInDel.cpp:1715-1737) over flat buffers;
 * results r-major like rl[] (DInDel.cpp:2129). */
int ora_window(const char *hap_seq, const int32_t *hap_off, int nh, const char *read_seq,
               const uint8_t *read_qual, const int32_t *read_off, const uint8_t *mapq,
               const uint32_t *mReadStart, const uint8_t *flags, int nr, uint32_t hapStart,
               const ora_params *p, ora_pair_out *out, int32_t *hpos, int hpos_stride)
{
    int h, r, b, n = 0, maxR = 0;
    double *q;
    for (r = 0; r < nr; r++)
        if (read_off[r + 1] - read_off[r] > maxR) maxR = read_off[r + 1] - read_off[r];
    q = (double *)malloc(sizeof(double) * (maxR + 1));
    for (h = 0; h < nh; h++)
        for (r = 0; r < nr; r++) {
            int R = read_off[r + 1] - read_off[r];
            size_t pi = (size_t)r * nh + h;
            for (b = 0; b < R; b++) q[b] = ora_phred_to_prob(read_qual[read_off[r] + b]);
            if (ora_pair(hap_seq + hap_off[h], hap_off[h + 1] - hap_off[h], read_seq + read_off[r], R,
                         q, ora_phred_to_prob(mapq[r]), (double)mReadStart[r], flags[r] & 1, hapStart,
                         p, &out[pi], hpos ? hpos + pi * hpos_stride : 0, 0)) {
                free(q);
                return -1;
            }
            n++;
        }
    free(q);
    return n;
}

/* Utils.hpp:29-38 */
double ora_addlogs(double l1, double l2)
{
    if (l1 > l2) {
        double diff = l2 - l1;
        return l1 + log(1.0 + exp(diff));
    } else {
        double diff = l1 - l2;
        return l2 + log(1.0 + exp(diff));
    }
}

/* Digamma for x > 0.  The reference calls boost::math::digamma (DInDel.cpp:2466,2472), a
 * third-party function absent from /root/reference and unpinned in version ("parity unpinned").
 * Published algorithm restated: upward recurrence psi(x) = psi(x+1) - 1/x to x >= 10, then the
 * asymptotic series ln x - 1/(2x) - sum B_2k / (2k x^2k).  The CUDA K3 mirrors this exactly. */
double ora_digamma(double x)
{
    double r = 0.0, f, t;
    while (x < 10.0) {
        r -= 1.0 / x;
        x += 1.0;
    }
    f = 1.0 / (x * x);
    t = f * (-1.0 / 12.0 + f * (1.0 / 120.0 + f * (-1.0 / 252.0 + f * (1.0 / 240.0 + f * (-1.0 / 132.0
        + f * (691.0 / 32760.0 + f * (-1.0 / 12.0)))))));
    return r + log(x) - 0.5 / x + t;
}

/* ------------------------------------------------------------------------------------------
 * K2: haplotype-pair sums.  DInDel.cpp:3085-3093 (start 0.0, `ll+=log(0.5)+addLogs(...)`),
 * :3378-3381 (start `pr`), :2683-2690 (reads of one pool member).  rl is r-major [nr*nh].
 * q enumerates h1 <= h2 row-major.  group < 0: all reads.
 * ------------------------------------------------------------------------------------------ */
void ora_pair_sums(const double *rl, int nr, int nh, const int32_t *read_group, int group,
                   const double *start, double *out)
{
    int h1, h2, r, q = 0;
    double log5 = log(0.5);
    for (h1 = 0; h1 < nh; h1++)
        for (h2 = h1; h2 < nh; h2++, q++) {
            double ll = start[q];
            for (r = 0; r < nr; r++) {
                if (group >= 0 && read_group && read_group[r] != group) continue;
                ll += log5 + ora_addlogs(rl[r * nh + h1], rl[r * nh + h2]);
            }
            out[q] = ll;
        }
}

/* ------------------------------------------------------------------------------------------
 * K3: one variational-Bayes EM run, DInDel.cpp:2413-2542 statement by statement (z, nk, lpi, pi,
 * eNew, convergence `fabs(eOld-eNew)<tol || iter>25`), digamma = ora_digamma (Boost's is absent:
 * parity unpinned for that function).  Outputs loglik (:2542), pi (log), freqs (:2539), iters.
 * ------------------------------------------------------------------------------------------ */
int ora_bayes_em(const double *rl, int nr, int nh, const uint8_t *compatible, double a0, double tol,
                 double *loglik_out, double *pi_out, double *freqs_out)
{
    double *z = (double *)malloc(sizeof(double) * (size_t)nr * nh);
    double *pi = (double *)malloc(sizeof(double) * nh), *lpi = (double *)malloc(sizeof(double) * nh);
    double *nk = (double *)malloc(sizeof(double) * nh), *ak = (double *)malloc(sizeof(double) * nh);
    int h, r, idx, numah = 0, iter = 0, converged = 0;
    double eOld = -HUGE_VAL, eNew, loglik = 0.0, zs;
    for (h = 0; h < nh; h++) if (compatible[h]) numah++;
    for (h = 0; h < nh; h++) lpi[h] = compatible[h] ? log(1.0 / (double)numah) : -100;
    while (!converged) {
        double ahat = 0.0, dahat;
        for (h = 0; h < nh; h++) nk[h] = 0.0;
        loglik = 0.0;
        idx = 0;
        for (r = 0; r < nr; r++) {
            double lognorm = -HUGE_VAL;
            for (h = 0; h < nh; h++) {
                z[idx] = lpi[h] + (rl[idx]);
                lognorm = ora_addlogs(lognorm, z[idx]);
                idx++;
            }
            idx -= nh;
            for (h = 0; h < nh; h++) {
                z[idx] -= lognorm;
                z[idx] = exp(z[idx]);
                nk[h] += z[idx];
                idx++;
            }
            loglik += lognorm;
        }
        for (h = 0; h < nh; h++) if (compatible[h]) {
            ak[h] = nk[h] + a0;
            ahat += ak[h];
        }
        dahat = ora_digamma(ahat);
        for (h = 0; h < nh; h++) {
            if (compatible[h]) {
                lpi[h] = ora_digamma(ak[h]) - dahat;
                pi[h] = log((a0 + nk[h]) / ((double)numah * a0 + (double)nr));
            } else {
                lpi[h] = -100;
                pi[h] = -100;
            }
        }
        idx = 0;
        eNew = 0.0;
        for (r = 0; r < nr; r++)
            for (h = 0; h < nh; h++) {
                eNew += z[idx] * (pi[h] + rl[idx]);
                idx++;
            }
        converged = (fabs(eOld - eNew)) < tol || iter > 25;
        eOld = eNew;
        iter++;
    }
    zs = 0.0;
    for (h = 0; h < nh; h++) zs += exp(pi[h]);
    for (h = 0; h < nh; h++) {
        pi_out[h] = pi[h];
        freqs_out[h] = exp(pi[h]) / zs;
    }
    *loglik_out = loglik;
    free(z); free(pi); free(lpi); free(nk); free(ak);
    return iter;
}

/* ------------------------------------------------------------------------------------------
 * a13: DetInDel::filterHaplotypes (DInDel.cpp:1932-2100) for one window, on flat arrays.
 * Pair (h, r) lives at r*nh+h: offHapHMQ / numIndels / hpos row (int32, hpos_stride apart).
 * Entries = the (haplotype, variant) iterations of the loop at :1962, sorted by haplotype then map
 * order, each naming its PAV key (:1943) and its interval left = leftFlankRead - padCover,
 * right = rightFlankRead + padCover (:1973-1974).
 *   selReads (:1948-1956); per DEL entry :1984-2014; per INS entry :2018-2063 (both `lenins` branches
 *   end in "every position of [left,right] hit by this read && nmm <= maxMismatch", so lenins does not
 *   enter); break at the first uncovered entry :2066-2071; filtered :2078-2083; per key the union over
 *   unfiltered haplotypes, per strand :2089-2097.
 * Two deliberate deviations from undefined behaviour in the reference (SURVEY.md 7 "hard parts"): the
 * read of hpos[b] at b == hpos.size() (loop bound `<=`, :1995/:2034) is "no hit", and negative hpos codes
 * (LO/RO/INS) never hit (the reference would index hap.seq with them).
 * ------------------------------------------------------------------------------------------ */
void ora_filter_haplotypes(int nh, int nr, const char *hap_seq, const int32_t *hap_off, const char *read_seq,
                           const int32_t *read_off, const int32_t *hpos, int hpos_stride,
                           const uint8_t *offHapHMQ, const int32_t *numIndels, const uint8_t *read_strand,
                           int ne, const int32_t *entry_hap, const int32_t *entry_key, const uint8_t *entry_type,
                           const int32_t *entry_left, const int32_t *entry_right, int nk, int maxMismatch,
                           int doFilter, uint8_t *filtered, uint8_t *uncovered, int32_t *key_nf, int32_t *key_nr)
{
    /* hVarCoverage[key][h + strand*nh] as read bit maps */
    uint8_t *cov = (uint8_t *)calloc((size_t)(nk > 0 ? nk : 1) * 2 * nh * (nr > 0 ? nr : 1), 1);
    int h, r, e, b, k, x;
    for (h = 0; h < nh; h++) {
        int allCovered = 1;
        const char *hs = hap_seq + hap_off[h];
        const int H = hap_off[h + 1] - hap_off[h];
        for (e = 0; e < ne; e++) {
            int left, right, len, covered = 0;
            uint8_t *hit;
            if (entry_hap[e] != h) continue;
            if (entry_type[e] != 0 && entry_type[e] != 1) continue; /* key exists, nothing to test */
            left = entry_left[e]; right = entry_right[e]; len = right - left + 1;
            hit = (uint8_t *)malloc(len > 0 ? len : 1);
            for (r = 0; r < nr; r++) {
                const size_t pi = (size_t)r * nh + h;
                const int R = read_off[r + 1] - read_off[r];
                const char *rs = read_seq + read_off[r];
                int nmm = 0, csize = 0, all = 1;
                if (offHapHMQ[pi] || numIndels[pi] != 0) continue; /* selReads */
                for (x = 0; x < len; x++) hit[x] = 0;
                for (b = 0; b < R; b++) { /* the reference runs to b == R: see header */
                    const int hb = hpos[pi * hpos_stride + b];
                    if (hb >= 0 && hb < H && hb >= left && hb <= right) {
                        if (!hit[hb - left]) csize++;
                        hit[hb - left] = 1;
                        if (entry_type[e] == 0) { if (hs[hb] != 'N' && hs[hb] != rs[b]) nmm++; }
                        else { if (hs[hb] != rs[b]) nmm++; }
                    }
                }
                for (x = 0; x < len; x++) if (!hit[x]) all = 0;
                if (entry_type[e] == 0) {
                    if (csize >= len && nmm <= maxMismatch) { covered = 1; cov[((size_t)entry_key[e] * 2 * nh + h + (read_strand[r] ? nh : 0)) * nr + r] = 1; }
                } else {
                    if (nmm <= maxMismatch && all && csize >= len) { covered = 1; cov[((size_t)entry_key[e] * 2 * nh + h + (read_strand[r] ? nh : 0)) * nr + r] = 1; }
                }
            }
            free(hit);
            if (!covered) { allCovered = 0; break; }
        }
        if (uncovered) uncovered[h] = (uint8_t)!allCovered;
        filtered[h] = (uint8_t)(doFilter && !allCovered);
    }
    for (k = 0; k < nk; k++) {
        int nf = 0, nrv = 0;
        for (r = 0; r < nr; r++) {
            int f = 0, v = 0;
            for (h = 0; h < nh; h++) if (!filtered[h]) {
                f |= cov[((size_t)k * 2 * nh + h) * nr + r];
                v |= cov[((size_t)k * 2 * nh + nh + h) * nr + r];
            }
            nf += f; nrv += v;
        }
        key_nf[k] = nf; key_nr[k] = nrv;
    }
    free(cov);
}

/* ---- realigned-BAM CIGARs (--outputRealignedBAM) --------------------------------------------
 * ora_choose_hap: which haplotype a read is realigned to.
 *   mode 0 (diploid, DInDel.cpp:597-606): the better of the MAP pair (h1, h2); |ll1-ll2| < 1e-8 -> h1 when it has
 *           strictly fewer INS/DEL variants (Haplotype::countIndels, Haplotype.hpp:254-261), else h2.
 *   mode 1 (pooled, DInDel.cpp:503-509): first strict arg-max of ll over all haplotypes, hidx = 0 to start with.
 * ll[h] is the read's likelihood on haplotype h (liks[h][r].ll). */
int ora_choose_hap(int mode, int nh, const double *ll, const int32_t *n_indels, int h1, int h2)
{
    if (mode == 0) {
        if (fabs(ll[h1] - ll[h2]) < 1e-8) return n_indels[h1] < n_indels[h2] ? h1 : h2;
        return ll[h1] > ll[h2] ? h1 : h2;
    }
    double llmax = -HUGE_VAL;
    int hidx = 0;
    for (int h = 0; h < nh; h++)
        if (ll[h] > llmax) { llmax = ll[h]; hidx = h; }
    return hidx;
}

/* ora_get_cigar: DetInDel::getCIGAR (DInDel.cpp:728-882).  hpos[R] = the read's MLAlignment::hpos on the chosen
 * haplotype, hml[H] = that haplotype's own alignment to the reference segment (hap.ml.hpos,
 * ObservationModelSeqAn.hpp:170-233: >= 0, or INS / LO / RO).  ops[] receives (op, len) pairs in BAM numbering
 * (M 0, I 1, D 2, S 4), at most cap of them are stored; returns the number of operations, or -(n) for the
 * reference's `throw string("Error(n)!")` (n = 1..4) and -5 for "How is this possible? (1)".  *refPos = CIGAR::refPos
 * (left unset by the reference for a wholly clipped read: reported as -1 here).
 * The reference's scan for the last aligned base (`while (npos[b]<0) b--`, :779) runs off the front of the vector
 * when no base is aligned; that case is "lastbonh < 0" here. */
int ora_get_cigar(int R, const int32_t *hpos, int H, const int32_t *hml, int refSeqStart, int32_t *refPos,
                  int32_t *ops, int cap)
{
    enum { M = 0, I = 1, D = 2, S = 4, INS = -1 };
    int n = 0;
#define PUSH(op, len) do { if (n < cap) { ops[2 * n] = (op); ops[2 * n + 1] = (len); } n++; } while (0)
#define NPOS(b) (hpos[b] >= 0 ? (hpos[b] < H ? hml[hpos[b]] : -5) : hpos[b])   /* :745-747 */
    *refPos = -1;
    int b = R - 1;
    while (b >= 0 && NPOS(b) < 0) b--;
    const int lastbonh = b;
    if (lastbonh < 0) { PUSH(S, R); return n; }                                  /* :782-786 */
    b = 0;
    while (NPOS(b) < 0) b++;
    if (b > 0) PUSH(S, b);
    int prevponr = NPOS(b);
    *refPos = refSeqStart + prevponr;
    int curr = M, len = 1;
    while (b < lastbonh) {                                                       /* :806-862 */
        const int chp = NPOS(b), nhp = NPOS(b + 1);
        if (nhp == INS) {
            if (chp == INS) {
                if (curr != I) return -1;
                len++;
            } else if (chp >= 0) {
                if (curr != M) return -2;
                PUSH(M, len);
                len = 1; curr = I; prevponr = chp;
            } else return -5;
        } else if (chp >= 0 && nhp >= 0 && nhp - chp == 1) {
            if (curr != M) return -3;
            len++; prevponr = nhp;
        } else if (chp >= 0 && nhp >= 0 && nhp - chp > 1) {
            if (curr != M) return -4;
            PUSH(M, len);
            PUSH(D, nhp - chp - 1);
            curr = M; len = 1; prevponr = nhp;
        } else if (chp == INS && nhp - prevponr == 1) {
            PUSH(I, len);
            curr = M; len = 1; prevponr = nhp;
        } else if (chp == INS && nhp - prevponr > 1) {
            PUSH(I, len);
            PUSH(D, nhp - prevponr - 1);
            curr = M; len = 1; prevponr = nhp;
        }
        b++;
    }
    PUSH(curr, len);
    if (R - 1 - lastbonh > 0) PUSH(S, R - 1 - lastbonh);
#undef PUSH
#undef NPOS
    return n;
}

/* ------------------------------------------------------------------------------------------
 * K6: the per-read aggregation loops behind a called site.  recs is r-major [nr*nh] like rl; mapQual is
 * Read::mapQual; cand_* describe, per candidate haplotype, what the reference's map lookups find at the site
 * (present = `find(pos) != end()` of the coverage map it consults, start / end = startRead / endRead of that
 * haplotype's variant: hapIndelCovered / hapSNPCovered[pos] = isCovered(padCover, firstBase, lastBase),
 * ObservationModelFB.cpp:1465-1472, Variant.hpp:125-128).  Sums are left as sums (the caller's sqrt / divide
 * are at DInDel.cpp:2744, :3173, :3557, :3561, :3626).
 *   mode 0  DInDel.cpp:3396-3404 + :3504-3554 (candidates = hx1, hx2)
 *   mode 1  DInDel.cpp:3126-3132 + :3148-3170 (candidates = max_indel_pair; present = looped over and isIndel)
 *   mode 2  DInDel.cpp:2697-2742 (candidates = every haplotype; has_var = hapHasVar[h*nv+idx])
 * ------------------------------------------------------------------------------------------ */
static int covered(const ora_pair_out *p, int pad, int start, int end)
{
    return (p->firstBase + pad <= start && p->lastBase - pad >= end) ? 1 : 0;
}

void ora_read_stats(int mode, int nh, int nr, const ora_pair_out *recs, const double *mapQual, const uint8_t *reverse,
                    const uint8_t *unmapped, const int32_t *read_group, int group, int ncand, const int32_t *cand_hap,
                    const int32_t *cand_start, const int32_t *cand_end, const uint8_t *cand_flags, int pad,
                    ora_read_stats_t *o)
{
    int r, i, h;
    memset(o, 0, sizeof(*o));
#define IN_GROUP(r) (!(group >= 0 && read_group) || read_group[r] == group)
#define REC(h, r) (&recs[(size_t)(r) * nh + (h)])
    if (mode == 0) {
        const int hx1 = cand_hap[0], hx2 = cand_hap[1];
        for (r = 0; r < nr; r++)                                                        /* :3396-3404 */
            if (unmapped[r] && (!REC(hx1, r)->offHap || !REC(hx2, r)->offHap)) o->numUnmappedRealigned++;
        for (r = 0; r < nr; r++) {                                                      /* :3504 */
            double mq;
            int k, nrt = 0, nft = 0;
            const ora_pair_out *p;
            if (!IN_GROUP(r)) continue;
            o->n_reads++;
            mq = -10 * log10(1.0 - mapQual[r]);
            o->allmsq += (mq * mq);
            if (REC(hx1, r)->offHap && REC(hx2, r)->offHap) o->numOffBoth++;
            k = (REC(hx1, r)->ll >= REC(hx2, r)->ll) ? 0 : 1;                          /* :3514-3520 */
            p = REC(cand_hap[k], r);
            o->numIndels += p->numIndels;
            o->nBQT += p->nBQT;
            o->nmmBQT += p->nmmBQT;
            o->mLogBQ += p->mLogBQ;
            if (p->nMMLeft >= 2) o->nMMLeft++;
            if (p->nMMRight >= 2) o->nMMRight++;
            if ((cand_flags[k] & 1) && covered(p, pad, cand_start[k], cand_end[k])) {   /* :3537-3550 */
                double mq2 = -10 * log10(1.0 - mapQual[r]);
                if (reverse[r]) nrt = 1; else nft = 1;
                o->msq += mq2 * mq2;
                o->n++;
            }
            if (nft) o->nf++;
            if (nrt) o->nr++;
        }
    } else if (mode == 1) {
        for (r = 0; r < nr; r++)                                                        /* :3126-3132 */
            if (ncand >= 2 && unmapped[r] && (!REC(cand_hap[0], r)->offHap || !REC(cand_hap[1], r)->offHap))
                o->numUnmappedRealigned++;
        for (r = 0; r < nr; r++) if (IN_GROUP(r)) o->n_reads++;
        for (i = 0; i < ncand && i < 2; i++) {                                          /* :3152 */
            if (!(cand_flags[i] & 1)) continue;                                         /* :3155 */
            h = cand_hap[i];
            for (r = 0; r < nr; r++) {
                if (!IN_GROUP(r)) continue;
                if (covered(REC(h, r), pad, cand_start[i], cand_end[i])) {              /* :3160-3166 */
                    double mq = -10 * log10(1.0 - mapQual[r]);
                    if (reverse[r]) o->nr++; else o->nf++;
                    o->msq += mq * mq;
                    o->n++;
                }
            }
        }
    } else {
        for (r = 0; r < nr; r++) {                                                      /* :2697 */
            double ml = -HUGE_VAL, mq;
            int nrt = 0, nft = 0;
            if (!IN_GROUP(r)) continue;
            o->n_reads++;
            for (i = 0; i < ncand; i++)                                                 /* :2702-2707 */
                if (REC(cand_hap[i], r)->ll >= ml) ml = REC(cand_hap[i], r)->ll;
            for (i = 0; i < ncand; i++) {                                               /* :2708-2734 */
                const ora_pair_out *p = REC(cand_hap[i], r);
                if (!(p->ll >= ml - 1e-7)) continue;
                if ((cand_flags[i] & 1) && covered(p, pad, cand_start[i], cand_end[i]) && (cand_flags[i] & 2)) {
                    if (reverse[r]) nrt = 1; else nft = 1;
                }
            }
            mq = -10 * log10(1.0 - mapQual[r]);                                         /* :2736-2738 */
            o->msq += mq * mq;
            o->n++;
            if (nft) o->nf++;
            if (nrt) o->nr++;
        }
    }
#undef IN_GROUP
#undef REC
}
