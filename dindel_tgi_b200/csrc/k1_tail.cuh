// k1_tail.cuh — the part of one pair after the two Viterbi halves: back-trace
// (ObservationModelFB.cpp:1148-1165) and the alignment summary (reportVariants, cpp:1351-1463).
// Warp-cooperative; shared by the generic and the register-tiled K1 kernels, which differ only in
// how a back-pointer code is fetched (functor `Bt`).
//
// Back-pointer code of one state column x at one row: code0 = winner among the plane-0 ("not in
// insertion") candidates in reference source order, bit1 = 1 when the plane-1 state was reached
// from plane 0 instead of staying in the insertion.
//   left half  (passMessageTwoDec, cpp:1775): x in 1..H: code0 = y-1 (from x-y, clamped to 0),
//              J = from plane-1 x-1;  x == 0: 0;  x == RO: 0 RO, 1 H, 2 S+RO, 3 S+H
//              plane 1: bit1 -> from plane-0 x
//   right half (passMessageTwoInc, cpp:1715): x in 1..H: code0 = y-1 (to x+y, clamped to RO),
//              J = plane-1 x;  x == 0: 0 -> 0, 1 -> 1, J -> S;  x == RO: 0 RO, J S+RO
//              plane 1: bit1 -> plane-0 min(x+1, RO)  (x == 0: plane-0 0)
#pragma once
#include <type_traits>
#include "k1_common.cuh"

// Back-pointer rows that live in global memory (J = 11 classes): the walk reads ONE entry per row, each on its own
// 128-byte line that left the L1 long ago, so every step would wait a full L2 / DRAM round trip.  The walk follows
// the alignment's diagonal almost everywhere (x moves by one per row), so the entry K1_TAIL_PF_AHEAD rows further
// down the diagonal is prefetched; a wrong guess (indel in between) costs one useless line.  8 rows measured best
// (32 and 96 slightly worse: profiles/r2_k1_j11.md).
#ifndef K1_TAIL_PF_AHEAD
#define K1_TAIL_PF_AHEAD 8
#endif
// speculative parallel back-trace (k1_backtrace_parallel) for callers with their rows in shared / global memory; 0 = the
// sequential 2-lane walk.  Measured (profiles/r2_k1_regions.md): configs[1] 289.3 -> 306.9 GCUPS, configs[3] 141.2 -> 153.2
#ifndef K1_TAIL_PAR_SMEM
#define K1_TAIL_PAR_SMEM 1
#endif
#ifndef K1_TAIL_PAR_GLOBAL
#define K1_TAIL_PAR_GLOBAL 1
#endif
struct K1NoPrefetch {
    __device__ __forceinline__ void operator()(int, int) const {}
};

__device__ __forceinline__ int k1_prev_state_dec(int s, int code0, int bit1, int S, int H, int J)
{
    int RO = H + 1;
    if (s >= S) return bit1 ? (s - S) : s;
    if (s == 0) return 0;
    if (s == RO) return (code0 == 0) ? RO : (code0 == 1) ? H : (code0 == 2) ? S + RO : S + H;
    if (code0 >= J) return S + s - 1;
    int nx = s - code0 - 1;
    return nx < 0 ? 0 : nx;
}

__device__ __forceinline__ int k1_next_state_inc(int s, int code0, int bit1, int S, int H, int J)
{
    int RO = H + 1;
    if (s >= S) {
        int x = s - S;
        if (!bit1) return s;
        if (x == 0) return 0;
        return (x + 1 > RO) ? RO : x + 1;
    }
    if (s == 0) return (code0 == 0) ? 0 : (code0 == 1) ? 1 : S;
    if (s == RO) return (code0 == 0) ? RO : S + RO;
    if (code0 >= J) return S + s;
    int nx = s + code0 + 1;
    return nx > H ? RO : nx;
}

// ---- speculative parallel back-trace -------------------------------------------------------------------------
// The back-trace (cpp:1148-1165) is a chain of R-1 dependent steps, but an alignment follows a diagonal almost
// everywhere, so its states can be GUESSED: each of the 32 lanes takes K consecutive steps of one walk, starts them
// from the state the diagonal predicts, and the warp then checks that every lane's entry state is the previous lane's
// exit state.  Lane 0's entry is known; the first lane whose guess was wrong gets the true state (the exit of the lane
// before it, right by induction), the lanes after it a new guess along the diagonal from there, and the round is
// repeated.  Every round settles at least one more lane (at most 32 rounds; an alignment without indels: one), and the
// last round, in which all entries are right, writes the final states.  The guess only steers speed; what is stored
// is always produced by the same decode as the sequential walk.
template <class Bt>
__device__ __forceinline__ int k1_bt_step(const Bt &bt, bool left, int row, int s, int S, int H, int J)
{
    const int RO = H + 1;
    const bool plane1 = s >= S;
    const int x = plane1 ? s - S : s;
    const int c = bt(row, x);
    const int code = c & 127;
    if ((unsigned)(s - 1) < (unsigned)H) { // plane 0, interior column (k1_prev_state_dec / k1_next_state_inc, 1 <= s <= H)
        const int t = left ? x - code - 1 : x + code + 1;
        const int jump = left ? (t < 0 ? 0 : t) : (t > H ? RO : t);
        return (code >= J) ? (left ? S + x - 1 : S + x) : jump;
    }
    const bool bit1 = (c >> 7) != 0;
    const int p1d = bit1 ? x : s;
    const int xi = (x + 1 > RO) ? RO : x + 1;
    const int p1i = !bit1 ? s : (x == 0 ? 0 : xi);
    const int i0 = (code == 0) ? 0 : (code == 1) ? 1 : S;
    const int dR = (code == 0) ? RO : (code == 1) ? H : (code == 2) ? S + RO : S + H;
    const int iR = (code == 0) ? RO : S + RO;
    const int d = plane1 ? p1d : (x == 0 ? 0 : dR);
    const int u = plane1 ? p1i : (x == 0 ? i0 : iR);
    return left ? d : u;
}
// the state k steps further along the diagonal (a guess: LO / RO / insertion states are assumed to stay)
__device__ __forceinline__ int k1_bt_guess(bool left, int s, int k, int S, int H)
{
    if ((unsigned)(s - 1) >= (unsigned)H) return s;
    const int t = left ? s - k : s + k;
    return left ? (t < 0 ? 0 : t) : (t > H ? H + 1 : t);
}
template <class Bt>
__device__ __forceinline__ void k1_backtrace_parallel(const Bt &bt, int lane, int R, int H, int J, int bMid, int msMid, int16_t *sm_ms)
{
    const int S = H + 2;
    if (lane == 0) sm_ms[bMid] = (int16_t)msMid;
#pragma unroll 1
    for (int dir = 0; dir < 2; dir++) {
        const bool left = dir == 0;
        const int n = left ? bMid : (R - 1 - bMid);       // steps of this walk
        if (n <= 0) continue;
        const int K = (n + 31) >> 5;                       // steps per lane
        const int nl = (n + K - 1) / K;                    // lanes that have steps
        const int j0 = lane * K;                           // this lane's first step
        int e = k1_bt_guess(left, msMid, j0, S, H);        // entry state (lane 0: msMid itself)
        int settled = 1;                                   // lanes below this have right entries
        for (;;) {
            int s = e;
            for (int j = 0; j < K; j++) {
                const int st = j0 + j;
                if (st < n) {
                    // left: from read base bMid-st to bMid-st-1 through btf[bMid-st]; right: from bMid+st to bMid+st+1 through btb[bMid+st]
                    s = k1_bt_step(bt, left, left ? bMid - st : bMid + st + 1, s, S, H, J);
                    sm_ms[left ? bMid - st - 1 : bMid + st + 1] = (int16_t)s;
                }
            }
            const int prev = __shfl_up_sync(0xffffffffu, s, 1);
            const unsigned bad = __ballot_sync(0xffffffffu, lane >= settled && lane < nl && e != prev);
            if (!bad) break;
            const int f = __ffs(bad) - 1;                  // first lane with a wrong entry; the lanes before it are right
            const int tru = __shfl_sync(0xffffffffu, s, f - 1);
            if (lane >= f) e = k1_bt_guess(left, tru, (lane - f) * K, S, H);
            settled = f + 1;
        }
    }
}

// Everything after ms[bMid] is known.  `bt(row, x)` returns code0 | bit1 << 7 of state column x;
// rows 1..bMid hold the left half (btf[b]), rows bMid+1..R-1 hold btb[row-1].
// sm_ms: int16[R] scratch; sm_ch / sm_ph: the read's bases and phred bytes; sm_acc: double[R].
template <class Bt, class Pf>
__device__ __forceinline__ void k1_tail(const Bt &bt, const Pf &pf, int lane, int R, int H, int J, int bMid,
                                        int msMid, const uint8_t *hap, const uint8_t *sm_ch,
                                        const uint8_t *sm_ph, int16_t *sm_ms, double *sm_acc,
                                        const K1Luts *__restrict__ L, double ll, double llOn,
                                        double llOff, int offHap, int offHapHMQ,
                                        dindel_pair_out *out, int16_t *hpos_out)
{
    const int S = H + 2;
    // --- back-trace, cpp:1148-1165.  As a 2-lane sequential walk (lane 0 left through btf, lane 1 right through btb) it
    //     was 10 % of the J = 6 kernel's issued instructions (profiles/r2_k1_regions.md); the default is the speculative
    //     32-lane walk above, the sequential loops below stay as the measured alternative (K1_TAIL_PAR_* = 0) ---
    constexpr bool kParallel = std::is_same<Pf, K1NoPrefetch>::value ? (K1_TAIL_PAR_SMEM != 0) : (K1_TAIL_PAR_GLOBAL != 0);
    if constexpr (kParallel) {
        k1_backtrace_parallel(bt, lane, R, H, J, bMid, msMid, sm_ms);
    } else
    if (lane < 2) {
        const bool left = (lane == 0);
        const int RO = H + 1;
        const int n = left ? bMid : (R - 1 - bMid);
        int s = msMid, b = bMid;
        if (left) sm_ms[bMid] = (int16_t)s;
        if constexpr (!std::is_same<Pf, K1NoPrefetch>::value) { // the first rows of the diagonal
            const int x0 = s >= S ? s - S : s;
            for (int k = 1; k < K1_TAIL_PF_AHEAD && k < n; k++) {
                const int xp = left ? x0 - k : x0 + k;
                pf(left ? b - k : b + 1 + k, xp < 0 ? 0 : (xp > RO ? RO : xp));
            }
        }
        if constexpr (std::is_same<Pf, K1NoPrefetch>::value) {
            // rows in shared memory (J = 6 classes, generic kernel): nearly every step sits on an interior plane-0
            // column, which takes a short path of its own (+2.0 % on configs[1])
            for (int i = 0; i < n; i++) {
                const bool plane1 = s >= S;
                const int x = plane1 ? s - S : s;
                const int c = bt(left ? b : b + 1, x);
                const int code = c & 127;
                if ((unsigned)(s - 1) < (unsigned)H) {
                    // a jump of code+1 columns, or into / out of the insertion plane (k1_prev_state_dec /
                    // k1_next_state_inc, 1 <= s <= H)
                    const int t = left ? x - code - 1 : x + code + 1;
                    const int jump = left ? (t < 0 ? 0 : t) : (t > H ? RO : t);
                    s = (code >= J) ? (left ? S + x - 1 : S + x) : jump;
                } else {
                    const bool bit1 = (c >> 7) != 0;
                    // plane 1 (s >= S)
                    const int p1d = bit1 ? x : s;
                    const int xi = (x + 1 > RO) ? RO : x + 1;
                    const int p1i = !bit1 ? s : (x == 0 ? 0 : xi);
                    // plane 0, LO and RO
                    const int i0 = (code == 0) ? 0 : (code == 1) ? 1 : S;
                    const int dR = (code == 0) ? RO : (code == 1) ? H : (code == 2) ? S + RO : S + H;
                    const int iR = (code == 0) ? RO : S + RO;
                    const int d = plane1 ? p1d : (x == 0 ? 0 : dR);
                    const int u = plane1 ? p1i : (x == 0 ? i0 : iR);
                    s = left ? d : u;
                }
                b += left ? -1 : 1;
                sm_ms[b] = (int16_t)s;
            }
        } else {
            // rows in global memory (J = 11 classes): the walk waits for the load, the branch-free form (both decodes
            // evaluated, the lane's own one selected) measured 1.8 % faster there than the short path
            for (int i = 0; i < n; i++) {
                const bool plane1 = s >= S;
                const int x = plane1 ? s - S : s;
                if (i + K1_TAIL_PF_AHEAD < n) {
                    const int xp = left ? x - K1_TAIL_PF_AHEAD : x + K1_TAIL_PF_AHEAD;
                    pf(left ? b - K1_TAIL_PF_AHEAD : b + 1 + K1_TAIL_PF_AHEAD, xp < 0 ? 0 : (xp > RO ? RO : xp));
                }
                const int c = bt(left ? b : b + 1, x);
                const int code = c & 127;
                const bool bit1 = (c >> 7) != 0;
                // plane 1 (k1_prev_state_dec / k1_next_state_inc, s >= S)
                const int p1d = bit1 ? x : s;
                const int xi = (x + 1 > RO) ? RO : x + 1;
                const int p1i = !bit1 ? s : (x == 0 ? 0 : xi);
                // plane 0, interior column
                const int jd0 = x - code - 1, ji0 = x + code + 1;
                const int gd = (code >= J) ? S + x - 1 : (jd0 < 0 ? 0 : jd0);
                const int gi = (code >= J) ? S + x : (ji0 > H ? RO : ji0);
                // plane 0, LO and RO
                const int i0 = (code == 0) ? 0 : (code == 1) ? 1 : S;
                const int dR = (code == 0) ? RO : (code == 1) ? H : (code == 2) ? S + RO : S + H;
                const int iR = (code == 0) ? RO : S + RO;
                const int d = plane1 ? p1d : (x == 0 ? 0 : (x == RO ? dR : gd));
                const int u = plane1 ? p1i : (x == 0 ? i0 : (x == RO ? iR : gi));
                s = left ? d : u;
                b += left ? -1 : 1;
                sm_ms[b] = (int16_t)s;
            }
        }
    }
    __syncwarp();
    // --- summary, cpp:1351-1463, one read base per lane ---
    int first = 0x7fffffff, last = -1;
    int nInd = 0, nMis = 0, nBQT = 0, nmmBQT = 0, nL = 0, nR = 0;
    for (int b = lane; b < R; b += 32) {
        int s = sm_ms[b];
        int plane1 = s >= S;
        int x = plane1 ? s - S : s;
        bool on = (x > 0) && (x <= H);
        int hp;
        double acc = 0.0;
        if (!on) {
            hp = (x == 0) ? DINDEL_HPOS_LO : DINDEL_HPOS_RO;
        } else if (plane1) {
            hp = DINDEL_HPOS_INS;
            if (b == 0 || sm_ms[b - 1] < S) nInd++; // one event per insertion run (cpp:1379-1395)
        } else {
            hp = x - 1;
            first = min(first, hp);
            last = max(last, hp);
            int ph = sm_ph[b];
            int qf = L->qflag[ph];
            if (qf & 1) {
                nBQT++;
                acc = L->log10q[ph];
            }
            if (sm_ch[b] != hap[x - 1]) { // cpp:1410-1430
                if (qf & 1) nmmBQT++;
                if (b < 6) nL++;
                if (b > R - 6) nR++;
                if (qf & 2) nMis++;
            }
            if (b < R - 1) { // deletion, cpp:1437-1453
                int ns = sm_ms[b + 1];
                if (ns < S && ns - s > 1) nInd++;
            }
        }
        sm_acc[b] = acc;
        if (hpos_out) hpos_out[b] = (int16_t)hp;
    }
    first = __reduce_min_sync(0xffffffffu, first);
    last = __reduce_max_sync(0xffffffffu, last);
    nInd = __reduce_add_sync(0xffffffffu, nInd);
    nMis = __reduce_add_sync(0xffffffffu, nMis);
    nBQT = __reduce_add_sync(0xffffffffu, nBQT);
    nmmBQT = __reduce_add_sync(0xffffffffu, nmmBQT);
    nL = __reduce_add_sync(0xffffffffu, nL);
    nR = __reduce_add_sync(0xffffffffu, nR);
    __syncwarp();
    if (lane == 0) {
        // mLogBQ is a left-to-right FP64 sum in the reference (cpp:1406); adding +0.0 for the
        // bases that do not count leaves every partial sum bit-identical.
        double m = 0.0;
        for (int b = 0; b < R; b++) m += sm_acc[b];
        __align__(16) dindel_pair_out o;
        o.ll = ll; o.llOn = llOn; o.llOff = llOff; o.mLogBQ = m;
        o.firstBase = (last < 0) ? -1 : first;
        o.lastBase = last;
        o.numIndels = (uint16_t)nInd; o.numMismatch = (uint16_t)nMis;
        o.nBQT = (uint16_t)nBQT; o.nmmBQT = (uint16_t)nmmBQT;
        o.nMMLeft = (uint16_t)nL; o.nMMRight = (uint16_t)nR;
        o.offHap = (uint8_t)offHap; o.offHapHMQ = (uint8_t)offHapHMQ;
        o.bMid = (int16_t)bMid;
        // cpp DInDel.cpp:1722-1735: ll > 0.1 is fatal, NaN/Inf skips the window
        o.status = (ll > 0.1) ? DINDEL_WIN_LL_POSITIVE
                              : ((ll != ll || ll - ll != 0.0) ? DINDEL_WIN_NAN : DINDEL_WIN_OK);
        o.pad0 = 0;
        o.mapState0 = (uint16_t)sm_ms[0];
        for (int i = 0; i < 4; i++) o.pad[i] = 0;
        const int4 *src = reinterpret_cast<const int4 *>(&o);
        int4 *dst = reinterpret_cast<int4 *>(out);
        dst[0] = src[0]; dst[1] = src[1]; dst[2] = src[2]; dst[3] = src[3];
    }
    __syncwarp();
}

template <class Bt>
__device__ __forceinline__ void k1_tail(const Bt &bt, int lane, int R, int H, int J, int bMid, int msMid, const uint8_t *hap,
                                        const uint8_t *sm_ch, const uint8_t *sm_ph, int16_t *sm_ms, double *sm_acc,
                                        const K1Luts *__restrict__ L, double ll, double llOn, double llOff, int offHap,
                                        int offHapHMQ, dindel_pair_out *out, int16_t *hpos_out)
{
    k1_tail(bt, K1NoPrefetch(), lane, R, H, J, bMid, msMid, hap, sm_ch, sm_ph, sm_ms, sm_acc, L, ll, llOn, llOff, offHap,
            offHapHMQ, out, hpos_out);
}

// A pair that the reference refuses with "hapSize error." (cpp:47).
__device__ __forceinline__ void k1_write_hapsize_error(dindel_pair_out *out, int lane)
{
    if (lane == 0) {
        dindel_pair_out o;
        o.ll = 0.0; o.llOn = 0.0; o.llOff = 0.0; o.mLogBQ = 0.0;
        o.firstBase = -1; o.lastBase = -1;
        o.numIndels = 0; o.numMismatch = 0; o.nBQT = 0; o.nmmBQT = 0; o.nMMLeft = 0; o.nMMRight = 0;
        o.offHap = 0; o.offHapHMQ = 0; o.bMid = 0; o.status = DINDEL_WIN_HAPSIZE;
        o.pad0 = 0;
        o.mapState0 = 0;
        for (int i = 0; i < 4; i++) o.pad[i] = 0;
        *out = o;
    }
}
