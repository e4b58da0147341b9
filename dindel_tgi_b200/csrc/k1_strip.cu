// k1_strip.cu — K1 for wide shapes (J = maxLengthDel+1 = 11, haplotypes up to 382 bases: configs[3]): the register-
// tiled pair Viterbi of k1_tiled.cu walked in COLUMN STRIPS.
//
// A pair's 2*(H+2) states do not fit a warp at 4 columns per lane once H+2 > 128, and at 10-11 columns per lane the
// row loops of k1_tiled.cu become 20 KB of straight-line code holding 250 registers (profiles/r1_k1_tiled_j11.md:
// two warps per scheduler, instruction-fetch and latency bound, 0.46 of the FP64-issue roofline).  Here a warp still
// owns one (read, haplotype) pair but walks the haplotype in strips of 128 columns (4 per lane), KS_RB rows at a
// time: block of rows by block of rows, strip 0 runs its rows, then strip 1, then strip 2 (left half: rows 1..bMid,
// dependencies point left; the right half, rows R-1..bMid+1, runs the strips right to left).  Inside a strip the code
// is the 4-column row loop (~1/3 of the code, half the registers).  What a strip needs from its neighbour -- the
// previous row's values of the J columns across the boundary (+ one plane-1 value) -- the neighbour left, row by row,
// in a small per-warp EDGE buffer in shared memory (KS_RB + 1 rows x 13 doubles per boundary); between two blocks a
// strip's own row (8 doubles per lane) is parked in the warp's global scratch slot (L2: re-read ~40 us later).
// (A first version ran every strip over ALL rows with the edges in global memory: 250 rows later they had left the
// L2 and every strip-row waited for DRAM: 67 GCUPS against the tiled classes' 133.)
//
// Back-pointers: 4-bit winner code per column (16 bits per lane and row) + the plane-1 bit of the lane's 4 columns
// gathered by 4 ballots (16 bytes per row): 80 bytes per strip-row in a per-warp slot of global memory.
// Arithmetic per cell is that of k1_tiled.cu's shared-table path, operand for operand (same tables, same association,
// same specialised updateMax rules), so records and hpos rows are byte-identical to the generic kernel's.
#include <type_traits>
#include "k1_common.cuh"
#include "k1_tail.cuh"

#define K1_EPS 1e-10
#define FULLMASK 0xffffffffu
#define KS_C 4
#define KS_W 128          // columns per strip
#define KS_WPC 4          // warps per CTA
#define KS_MAX_STRIPS 3   // S <= 384
#define KS_EDGE 13        // doubles per edge-buffer row: [0..11] plane-0 values of the 12 boundary columns, [12] plane 1 of the last
#ifndef KS_RB
#define KS_RB 16          // rows per block
#endif
#ifndef K1_STRIP_MINB
#define K1_STRIP_MINB 3
#endif

__device__ __forceinline__ double ks_up(double v, int d) { return __shfl_up_sync(FULLMASK, v, d); }
__device__ __forceinline__ double ks_down(double v, int d) { return __shfl_down_sync(FULLMASK, v, d); }
__device__ __forceinline__ double ks_warp_max(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        double w = __shfl_xor_sync(FULLMASK, v, o);
        v = (w > v) ? w : v;
    }
    return v;
}
__device__ __forceinline__ uint32_t ks_accept(const uint8_t *hap, int x, int H)
{
    if (x < 1 || x > H) return 0xffffffffu;
    uint8_t c = hap[x - 1];
    return (c == 'N') ? 0xffffffffu : (1u << (c & 31));
}

__host__ __device__ inline size_t ks_smem_per_warp(int maxR)
{
    size_t b = (size_t)maxR * 16                       // eq / uq
               + (size_t)2 * KS_MAX_STRIPS * KS_W * 8  // alpha[bMid]+obs[bMid], then the meeting-row sums, by state column
               + (size_t)2 * (KS_RB + 1) * KS_EDGE * 8 // edge rows of the (at most two) strips that have a successor
               + (size_t)maxR * 2                      // map state
               + (size_t)maxR * 2;                     // read chars + phreds
    return (b + 15) & ~(size_t)15;
}
// per-warp global scratch: back-pointer rows of every strip + the strips' parked rows
__host__ __device__ inline size_t ks_bt_rows(int maxR) { return (size_t)maxR + 1; }
__host__ __device__ inline size_t ks_slot_bytes(int maxR)
{
    const size_t bt = (size_t)KS_MAX_STRIPS * ks_bt_rows(maxR) * 80;
    const size_t save = (size_t)KS_MAX_STRIPS * 2 * KS_C * 32 * 8; // a strip's row between two blocks
    return (bt + save + 255) & ~(size_t)255;
}

// prior of the meeting row for state column x (cpp:268-305 through the LUTs; insert-size term when INS)
struct KsPrior {
    double pLO0, pLO1, pOn0, pOn1, hLO0, hLO1, hOn0, hOn1, p1m, p1mH, pIns0, pIns1;
    int insMode, insAnchor, insMax, hapStart, R, bMid;
    const double *insLogp;
    double ins0;
};
template <bool INS>
__device__ __forceinline__ void ks_prior(const KsPrior &Q, int x, int RO, double &qR0, double &qR1, double &qH0, double &qH1)
{
    const bool lo = (x == 0), ro = (x == RO);
    qR0 = lo ? Q.pLO0 : Q.pOn0; qR1 = lo ? Q.pLO1 : Q.pOn1; qH0 = lo ? Q.hLO0 : Q.hOn0; qH1 = lo ? Q.hLO1 : Q.hOn1;
    if (INS && Q.insMode != DINDEL_INS_NONE && !ro) {
        if (lo) { // (log(pOff) + logpIns) + pinsert[0]  (cpp:298)
            qR0 = Q.pLO0 + Q.ins0; qR1 = Q.pLO1 + Q.ins0; qH0 = Q.hLO0 + Q.ins0; qH1 = Q.hLO1 + Q.ins0;
        } else {  // (pinsert[x] + log(1 - pOff)) + logpIns  (cpp:301)
            const double pin = k1_pinsert(Q.insLogp, Q.insMax, Q.insMode, Q.insAnchor, Q.hapStart, x, Q.R, Q.bMid);
            const double aR = pin + Q.p1m, aH = pin + Q.p1mH;
            qR0 = aR + Q.pIns0; qR1 = aR + Q.pIns1; qH0 = aH + Q.pIns0; qH1 = aH + Q.pIns1;
        }
    }
    if (ro) { qR0 = -100.0; qR1 = -100.0; qH0 = -100.0; qH1 = -100.0; }
}

// Sequential "v > best + EPS, first wins" scan over the 2S states (cpp:1096-1116) of one prior (HMQ selects the
// artificial high-mapping-quality one) over the meeting-row sums in shared memory: as scan_first_wins of k1_tiled.cu,
// the few states within 1e-9 of the maximum are replayed in state order (all of them if the guard band
// [M-2e-9, M-1e-9) is populated).
template <bool INS, bool HMQ>
__device__ __forceinline__ void ks_scan(const double *t0, const double *t1, const KsPrior &Q, int S, int RO, int lane,
                                        double &best, int &bestState, double *onMaxOut)
{
    auto val = [&](int i) -> double { // state i in [0, 2S)
        const int x = i < S ? i : i - S;
        double a, b, c, d;
        ks_prior<INS>(Q, x, RO, a, b, c, d);
        const double q = HMQ ? (i < S ? c : d) : (i < S ? a : b);
        return (i < S ? t0[x] : t1[x]) + q;
    };
    double m = K1_NEG_INF, on = K1_NEG_INF;
    for (int i = lane; i < 2 * S; i += 32) {
        const double v = val(i);
        m = (v > m) ? v : m;
        const int x = i < S ? i : i - S;
        if (x >= 1 && x < RO) on = (v > on) ? v : on;
    }
    m = ks_warp_max(m);
    if (onMaxOut) *onMaxOut = ks_warp_max(on);
    const double tNear = m - 1e-9, tGuard = m - 2e-9;
    bool guard = false;
    for (int i = lane; i < 2 * S; i += 32) {
        const double v = val(i);
        guard = guard || (v < tNear && v >= tGuard);
    }
    const bool all = __any_sync(FULLMASK, guard);
    best = K1_NEG_INF;
    bestState = 0;
    for (int i0 = 0; i0 < 2 * S; i0 += 32) {
        const int i = i0 + lane;
        const double v = (i < 2 * S) ? val(i) : K1_NEG_INF;
        unsigned near = __ballot_sync(FULLMASK, (i < 2 * S) && (all || v >= tNear));
        while (near) {
            const int src = __ffs(near) - 1;
            const double vv = __shfl_sync(FULLMASK, v, src);
            if (vv > best + K1_EPS) { best = vv; bestState = i0 + src; }
            near &= near - 1;
        }
    }
}

template <int J, int MINB, bool INS>
__global__ void __launch_bounds__(32 * KS_WPC, MINB) k1_strip_kernel(K1Batch B, int cls)
{
    constexpr int C = KS_C;
    extern __shared__ __align__(16) unsigned char ks_smem[];
    __shared__ double s_hp_lpe[K1_HP_LUT], s_hp_lpn[K1_HP_LUT];
    __shared__ double s_tl[K1_HP_LUT * (J - 1)]; // [code][y-2] = logProbError + (y-1)*lInsgIns (cpp:1791)
    __shared__ double s_tr[K1_HP_LUT * J];       // [code][y-1] = lp(y) + logProbNoError    (cpp:1732-1735)
    const int lane = threadIdx.x & 31;
    const int wib = threadIdx.x >> 5;
    const K1Luts *__restrict__ L = B.luts;
    const int maxR = B.max_R;
    for (int i = threadIdx.x; i < K1_HP_LUT; i += blockDim.x) {
        s_hp_lpe[i] = L->hp_lpe[i];
        s_hp_lpn[i] = L->hp_lpn[i];
    }
    for (int i = threadIdx.x; i < K1_HP_LUT * (J - 1); i += blockDim.x) {
        const int code = i / (J - 1), y = i % (J - 1) + 2;
        s_tl[i] = L->hp_lpe[code] + (double)(y - 1) * L->lInsgIns;
    }
    for (int i = threadIdx.x; i < K1_HP_LUT * J; i += blockDim.x) {
        const int code = i / J, y = i % J + 1;
        const double lpn_s = L->hp_lpn[code], lpe_s = L->hp_lpe[code];
        const double lp = (y == 1) ? lpn_s : (lpe_s + (double)(y - 1) * L->lInsgIns);
        s_tr[i] = lp + lpn_s;
    }
    __syncthreads();
    unsigned char *base = ks_smem + ks_smem_per_warp(maxR) * wib;
    double *sm_eq = reinterpret_cast<double *>(base);
    double *sm_uq = sm_eq + maxR;
    double *sm_ao = sm_uq + maxR; // [2][KS_MAX_STRIPS * KS_W]
    constexpr int AO1 = KS_MAX_STRIPS * KS_W;
    double *sm_edge = sm_ao + 2 * AO1; // [2][KS_RB + 1][KS_EDGE]
    int16_t *sm_ms = reinterpret_cast<int16_t *>(sm_edge + 2 * (KS_RB + 1) * KS_EDGE);
    uint8_t *sm_ch = reinterpret_cast<uint8_t *>(sm_ms + maxR);
    uint8_t *sm_ph = sm_ch + maxR;
    // per-warp global scratch
    unsigned char *slot = B.bt_scratch + ((long long)blockIdx.x * KS_WPC + wib) * B.bt_slot_bytes;
    const size_t btRows = ks_bt_rows(maxR);
    uint16_t *g_codes = reinterpret_cast<uint16_t *>(slot);                                                  // [strip][row][32]
    uint32_t *g_bits = reinterpret_cast<uint32_t *>(slot + (size_t)KS_MAX_STRIPS * btRows * 64);              // [strip][row][4]
    double *g_save = reinterpret_cast<double *>(slot + (size_t)KS_MAX_STRIPS * btRows * 80);                 // [strip][plane][c][32]

    const double lLOgLO = L->lLOgLO, lFirstgLO = L->lFirstgLO, lInsgIns = L->lInsgIns,
                 lNoInsgIns = L->lNoInsgIns, lNoInsgNoIns = L->lNoInsgNoIns;

    const unsigned u_begin = B.bucket_start[cls * K1_TILED_SLOTS];
    const unsigned u_end = B.bucket_start[(cls + 1) * K1_TILED_SLOTS];
    for (;;) {
        unsigned ui = 0;
        if (lane == 0) ui = u_begin + (unsigned)atomicAdd(B.counter + 1 + cls, 1ULL);
        ui = __shfl_sync(FULLMASK, ui, 0);
        if (ui >= u_end) break;
        const uint2 ud = B.units[ui];
        const int w = (int)ud.x, h = (int)(ud.y >> 16), chunk = (int)(ud.y & 0xffffu);
        const int h0 = B.win_hap[w], nh = B.win_hap[w + 1] - h0;
        const int r0 = B.win_read[w], nr = B.win_read[w + 1] - r0;
        const long long hoff = B.hap_off[h0 + h];
        const int H = (int)(B.hap_off[h0 + h + 1] - hoff);
        const int S = H + 2, RO = H + 1;
        const int NS = (S + KS_W - 1) / KS_W;
        const uint8_t *hap = B.hap_seq + hoff;
        const uint8_t *hp = B.hap_hp + hoff + 2LL * (h0 + h);
        const uint32_t hapStart = B.hap_start[w];
        const bool hapsize_err = L->maxLengthDel > H;
        const int rb = r0 + chunk * K1_READ_CHUNK;
        const int re = min(rb + K1_READ_CHUNK, r0 + nr);
        const int cRO = RO & (C - 1), lRO = (RO & (KS_W - 1)) >> 2; // RO sits in the last strip

        for (int r = rb; r < re; r++) {
            const long long pair = B.win_pair[w] + (long long)(r - r0) * nh + h;
            dindel_pair_out *out = B.pairs + pair;
            if (hapsize_err) {
                k1_write_hapsize_error(out, lane);
                continue;
            }
            const long long roff = B.read_off[r];
            const int R = (int)(B.read_off[r + 1] - roff);
            if (R <= 0) {
                if (lane == 0) {
                    k1_write_hapsize_error(out, 0);
                    out->status = DINDEL_WIN_NAN;
                }
                continue;
            }
            const int mapq = B.read_mapq[r];
            const bool unmapped = (B.read_flags[r] & DINDEL_READ_UNMAPPED) != 0;
            const int bMid = k1_bmid(hapStart, H, R, B.read_start[r], unmapped, L->bMidOverride);
            __syncwarp();
            for (int b = lane; b < R; b += 32) {
                int ph = B.read_qual[roff + b];
                sm_ch[b] = B.read_seq[roff + b];
                sm_ph[b] = (uint8_t)ph;
                sm_eq[b] = L->eq[ph];
                sm_uq[b] = L->uq[ph];
            }
            __syncwarp();

            // =================== left half: rows 1..bMid (passMessageTwoDec), strips left to right ===================
            for (int b0 = 0; b0 < bMid || b0 == 0; b0 += KS_RB) {
            const int b1 = min(bMid, b0 + KS_RB);
            for (int s = 0; s < NS; s++) {
                const int xs = s * KS_W, c0 = xs + lane * C;
                const bool first = (s == 0), last = (s == NS - 1);
                const bool isLane0 = first && (lane == 0), isLaneRO = last && (lane == lRO);
                double P0[C], P1[C], lpn[C], lpe[C];
                uint32_t am[C], amH[J + 1];
                int offL[C];
#pragma unroll
                for (int c = 0; c < C; c++) {
                    const int x = c0 + c;
                    const int code = hp[x <= RO ? x : RO];
                    am[c] = ks_accept(hap, x, H);
                    lpe[c] = s_hp_lpe[code];
                    lpn[c] = s_hp_lpn[code];
                    offL[c] = code * (J - 1);
                    if (b0 == 0) { // alpha[0] = 0 (cpp:335-338)
                        P0[c] = 0.0;
                        P1[c] = 0.0;
                    } else {       // the strip's row b0, parked at the end of the previous block
                        P0[c] = __ldcg(g_save + ((s * 2 + 0) * C + c) * 32 + lane);
                        P1[c] = __ldcg(g_save + ((s * 2 + 1) * C + c) * 32 + lane);
                    }
                }
#pragma unroll
                for (int j = 1; j <= J; j++) amH[j] = ks_accept(hap, c0 - j, H); // columns across the strip boundary (lanes 0..2)
                const double lpeLm1 = s_hp_lpe[hp[(c0 >= 1 && c0 - 1 <= RO) ? c0 - 1 : 0]];
                // edge rows: strip s writes row b at edgeW[b - b0]; strip s+1 reads row b-1 at edgeR[b - 1 - b0]
                double *edgeW = sm_edge + (size_t)(s & 1) * (KS_RB + 1) * KS_EDGE;
                const double *edgeR = sm_edge + (size_t)((s + 1) & 1) * (KS_RB + 1) * KS_EDGE;
                if (!last && lane >= 29) { // row b0 of this strip's right edge
#pragma unroll
                    for (int c = 0; c < C; c++) edgeW[(lane - 29) * C + c] = P0[c];
                    if (lane == 31) edgeW[12] = P1[C - 1];
                }
                __syncwarp();

                // ONE copy of the row loop per half: RO's slot is tested at run time (a warp-uniform branch per column
                // slot).  Per-slot instantiations (as k1_tiled.cu has them) multiply the code the warps of an SM -- each
                // in another strip, half and slot at any time -- pull through the instruction cache: with five copies
                // per half ncu showed 8.8 warps waiting for instructions per issue (profiles/r2_k1_strip.md).
                {
                for (int b = b0 + 1; b <= b1; b++) {
                    const double eq = sm_eq[b - 1], uq = sm_uq[b - 1];
                    const uint32_t rbit = 1u << (sm_ch[b - 1] & 31);
                    // previous row's values of the 11 columns left of the strip (+ plane 1 of the nearest one)
                    const double *er = edgeR + (size_t)(b - 1 - b0) * KS_EDGE;
                    double hv[J + 1];
                    double hp1 = 0.0;
                    if (!first && lane < 3) {
#pragma unroll
                        for (int j = 1; j <= J; j++) {
                            const int rel = 12 + lane * C - j; // boundary block = columns xs-12 .. xs-1
                            hv[j] = (rel < 12) ? er[rel < 12 ? rel : 0] : 0.0;
                        }
                        if (lane == 0) hp1 = er[12];
                    } else {
#pragma unroll
                        for (int j = 1; j <= J; j++) hv[j] = K1_NEG_INF; // a source left of LO can never win
                    }
                    double WO[C + J], WP[C + J], u[C], t1w[C + 1];
#pragma unroll
                    for (int c = 0; c < C; c++) {
                        WO[J + c] = (am[c] & rbit) ? eq : uq;
                        WP[J + c] = P0[c];
                        u[c] = WO[J + c] + P0[c];
                        t1w[c + 1] = eq + P1[c];
                    }
#pragma unroll
                    for (int j = 1; j <= J; j++) {
                        const int d = (j + C - 1) / C;
                        const int ci = d * C - j;
                        const double wp = ks_up(WP[J + ci], d), wo = ks_up(WO[J + ci], d);
                        WP[J - j] = (lane < d) ? hv[j] : wp;
                        WO[J - j] = (lane < d) ? ((amH[j] & rbit) ? eq : uq) : wo;
                    }
                    t1w[0] = ks_up(t1w[C], 1);
                    if (lane == 0) t1w[0] = eq + hp1; // strip 0: only LO would use it (overridden)

                    double best[C], a1[C];
                    uint32_t code[C], bit1[C];
                    // jumps y = 1..J from column x-y (cpp:1784-1795): "take iff v >= best" picks the LAST y attaining
                    // the maximum; reduced pairwise (later operand preferred on ties), as in k1_tiled.cu
#pragma unroll
                    for (int y = 1; y <= J; y += 2) {
#pragma unroll
                        for (int c = 0; c < C; c++) {
                            const double va = (y == 1) ? ((WO[J + c - 1] + lpn[c]) + WP[J + c - 1]) + lpn[c]
                                                       : ((WO[J + c - y] + s_tl[offL[c] + ((y >= 2) ? y - 2 : 0)]) + WP[J + c - y]) + lpn[c];
                            double pm = va;
                            uint32_t pc = (uint32_t)(y - 1);
                            if (y + 1 <= J) {
                                const double vb = ((WO[J + c - (y + 1)] + s_tl[offL[c] + y - 1]) + WP[J + c - (y + 1)]) + lpn[c];
                                const bool pb = (vb >= va);
                                pm = pb ? vb : va;
                                pc = pb ? (uint32_t)y : (uint32_t)(y - 1);
                            }
                            if (y == 1) {
                                best[c] = pm;
                                code[c] = pc;
                            } else {
                                const bool p = (pm >= best[c]);
                                best[c] = p ? pm : best[c];
                                code[c] = p ? pc : code[c];
                            }
                        }
                    }
                    // from plane-1 x-1 (cpp:1807-1811)
#pragma unroll
                    for (int c = 0; c < C; c++) {
                        const double vi = t1w[c] + lpe[c];
                        const bool p = (vi > best[c] + K1_EPS);
                        best[c] = p ? vi : best[c];
                        code[c] = p ? (uint32_t)J : code[c];
                    }
                    // plane 1 (cpp:1816-1826)
#pragma unroll
                    for (int c = 0; c < C; c++) {
                        const double stay = t1w[c + 1] + lInsgIns;
                        const double f = u[c] + lNoInsgIns;
                        const bool p = (f >= stay);
                        a1[c] = p ? f : stay;
                        bit1[c] = p ? 1u : 0u;
                        if (c == 0) { // LO (cpp:1798-1799); plane 1 only stays (loop from x=1)
                            const double lo = u[0] + lNoInsgNoIns;
                            best[0] = isLane0 ? lo : best[0];
                            code[0] = isLane0 ? 0u : code[0];
                            a1[0] = isLane0 ? stay : a1[0];
                            bit1[0] = isLane0 ? 0u : bit1[0];
                        }
                    }
                    // RO (cpp:1780-1782, 1804-1805)
#pragma unroll
                    for (int c = 0; c < C; c++)
                    if (last && c == cRO) {
                        const double uH = WO[J + c - 1] + WP[J + c - 1];
                        double rv = (u[c] + lLOgLO) + lNoInsgNoIns;
                        uint32_t rcode = 0;
                        const double v2 = (uH + lFirstgLO) + lNoInsgNoIns;
                        if (v2 >= rv) { rv = v2; rcode = 1; }
                        const double v3 = (t1w[c + 1] + lLOgLO) + lpe[c];
                        if (v3 > rv + K1_EPS) { rv = v3; rcode = 2; }
                        const double lpeH = (c >= 1) ? lpe[(c >= 1) ? c - 1 : 0] : lpeLm1;
                        const double v4 = (t1w[c] + lFirstgLO) + lpeH;
                        const bool take = (rcode == 2) ? (v4 >= rv) : (v4 > rv + K1_EPS);
                        if (take) { rv = v4; rcode = 3; }
                        best[c] = isLaneRO ? rv : best[c];
                        code[c] = isLaneRO ? rcode : code[c];
                    }
#pragma unroll
                    for (int c = 0; c < C; c++) {
                        P0[c] = best[c];
                        P1[c] = a1[c];
                    }
                    // back-pointers of this strip-row; this row's right edge for the next strip (slot b - s + 2, the one
                    // just read: every lane of the warp is past its loads of it)
                    g_codes[((size_t)s * btRows + b) * 32 + lane] = (uint16_t)(code[0] | (code[1] << 4) | (code[2] << 8) | (code[3] << 12));
#pragma unroll
                    for (int c = 0; c < C; c++) {
                        const unsigned m = __ballot_sync(FULLMASK, bit1[c] != 0);
                        if (lane == c) g_bits[((size_t)s * btRows + b) * 4 + c] = m;
                    }
                    if (!last && lane >= 29) {
                        double *ew = edgeW + (size_t)(b - b0) * KS_EDGE;
#pragma unroll
                        for (int c = 0; c < C; c++) ew[(lane - 29) * C + c] = P0[c];
                        if (lane == 31) ew[12] = P1[C - 1];
                    }
                }
                } // rows of the block
                if (b1 >= bMid) {
                    // alpha[bMid] + obs[bMid] (first add of cpp:1098) -> shared, by state column
                    const double eq = sm_eq[bMid], uq = sm_uq[bMid];
                    const uint32_t rbit = 1u << (sm_ch[bMid] & 31);
#pragma unroll
                    for (int c = 0; c < C; c++) {
                        sm_ao[c0 + c] = P0[c] + ((am[c] & rbit) ? eq : uq);
                        sm_ao[AO1 + c0 + c] = P1[c] + eq;
                    }
                } else {
#pragma unroll
                    for (int c = 0; c < C; c++) { // park row b1 until the next block
                        __stcg(g_save + ((s * 2 + 0) * C + c) * 32 + lane, P0[c]);
                        __stcg(g_save + ((s * 2 + 1) * C + c) * 32 + lane, P1[c]);
                    }
                }
                __syncwarp();
            }
            }
            // =================== right half: rows R-2..bMid (passMessageTwoInc), strips right to left ===================
            const int nRight = R - 1 - bMid; // rows t = 1 .. nRight (row t holds beta[R-1-t])
            for (int t0 = 0; t0 < nRight || t0 == 0; t0 += KS_RB) {
            const int t1 = min(nRight, t0 + KS_RB);
            for (int s = NS - 1; s >= 0; s--) {
                const int xs = s * KS_W, c0 = xs + lane * C;
                const int k = NS - 1 - s; // order of processing
                const bool first = (s == 0), last = (s == NS - 1);
                const bool isLane0 = first && (lane == 0), isLaneRO = last && (lane == lRO);
                double P0[C], P1[C], lpeN[C];
                uint32_t am[C], amH[J + 1];
                int offR[C + J];
#pragma unroll
                for (int q = 1; q < C + J; q++) offR[q] = (int)hp[(c0 + q <= RO) ? c0 + q : RO] * J;
#pragma unroll
                for (int c = 0; c < C; c++) {
                    const int x = c0 + c;
                    am[c] = ks_accept(hap, x, H);
                    lpeN[c] = s_hp_lpe[hp[(x + 1 <= RO) ? x + 1 : RO]]; // logProbError[x+1] (cpp:1748)
                    if (t0 == 0) {
                        P0[c] = (x <= RO) ? 0.0 : K1_NEG_INF; // beta[R-1] = 0; columns beyond RO never win
                        P1[c] = (x <= RO) ? 0.0 : K1_NEG_INF;
                    } else {
                        P0[c] = __ldcg(g_save + ((s * 2 + 0) * C + c) * 32 + lane);
                        P1[c] = __ldcg(g_save + ((s * 2 + 1) * C + c) * 32 + lane);
                    }
                }
#pragma unroll
                for (int j = 1; j <= J; j++) amH[j] = ks_accept(hap, c0 + C - 1 + j, H); // columns right of the strip (lanes 29..31)
                const double lpnRO = s_hp_lpn[hp[RO]];
                // edge rows: the strip of order k writes row t at edgeW[t - t0]; the next one reads row t at edgeR[t - t0]
                double *edgeW = sm_edge + (size_t)(k & 1) * (KS_RB + 1) * KS_EDGE;
                const double *edgeR = sm_edge + (size_t)((k + 1) & 1) * (KS_RB + 1) * KS_EDGE;
                if (!first && lane < 3) {
#pragma unroll
                    for (int c = 0; c < C; c++) edgeW[lane * C + c] = P0[c];
                }
                __syncwarp();

                {
                for (int t = t0; t < t1; t++) {
                    const int b = R - 1 - t; // beta[b] (row t) -> beta[b-1] (row t+1)
                    const double eq = sm_eq[b], uq = sm_uq[b];
                    const uint32_t rbit = 1u << (sm_ch[b] & 31);
                    // beta[b] of the 11 columns right of the strip
                    const double *er = edgeR + (size_t)(t - t0) * KS_EDGE;
                    double hv[J + 1];
                    if (!last && lane >= 29) {
#pragma unroll
                        for (int j = 1; j <= J; j++) {
                            const int rel = lane * C + C - 1 + j - KS_W; // column xs+128+rel
                            hv[j] = (rel >= 0) ? er[rel >= 0 ? rel : 0] : 0.0;
                        }
                    } else {
#pragma unroll
                        for (int j = 1; j <= J; j++) hv[j] = K1_NEG_INF; // beyond the last strip: never wins
                    }
                    double WO[C + J], WP[C + J], uw[C + 1], t1[C];
#pragma unroll
                    for (int c = 0; c < C; c++) {
                        WO[c] = (am[c] & rbit) ? eq : uq;
                        WP[c] = P0[c];
                        uw[c] = WO[c] + P0[c];
                        t1[c] = eq + P1[c];
                    }
#pragma unroll
                    for (int j = 1; j <= J; j++) {
                        const int d = (C - 1 + j) / C;
                        const int ci = (C - 1 + j) % C;
                        const double wp = ks_down(WP[ci], d), wo = ks_down(WO[ci], d);
                        WP[C - 1 + j] = (lane + d > 31) ? hv[j] : wp;
                        WO[C - 1 + j] = (lane + d > 31) ? ((amH[j] & rbit) ? eq : uq) : wo;
                    }
                    uw[C] = WO[C] + WP[C]; // = the neighbour's uw[0], operand for operand (last strip, lane 31: RO or padding, overridden)

                    double best[C], a1[C];
                    uint32_t code[C], bit1[C];
                    // jumps y = 1..J to column x+y (cpp:1724-1737)
#pragma unroll
                    for (int c = 0; c < C; c++) {
                        best[c] = (s_tr[offR[c + 1]] + WP[c + 1]) + WO[c + 1];
                        code[c] = 0;
                    }
#pragma unroll
                    for (int y = 2; y <= J; y++) {
#pragma unroll
                        for (int c = 0; c < C; c++) {
                            const double v = (s_tr[offR[c + y] + y - 1] + WP[c + y]) + WO[c + y];
                            const bool p = (v > best[c] + K1_EPS);
                            best[c] = p ? v : best[c];
                            code[c] = p ? (uint32_t)(y - 1) : code[c];
                        }
                    }
                    // from plane-1 x (cpp:1746-1749); LO's candidates (cpp:1720-1722, 1748)
                    double vi0 = 0.0;
#pragma unroll
                    for (int c = 0; c < C; c++) {
                        const double vi = t1[c] + lpeN[c];
                        if (c == 0) vi0 = vi;
                        const bool p = (vi > best[c] + K1_EPS);
                        best[c] = p ? vi : best[c];
                        code[c] = p ? (uint32_t)J : code[c];
                    }
                    double f[C];
#pragma unroll
                    for (int c = 0; c < C; c++) f[c] = uw[c + 1] + lNoInsgIns;
                    { // LO (cpp:1720-1722, 1748, 1762)
                        double rv = (uw[0] + lLOgLO) + lNoInsgNoIns;
                        uint32_t rcode = 0;
                        const double v2 = (uw[1] + lFirstgLO) + lNoInsgNoIns;
                        if (v2 > rv + K1_EPS) { rv = v2; rcode = 1; }
                        if (vi0 > rv + K1_EPS) { rv = vi0; rcode = J; }
                        const double f0 = uw[0] + lNoInsgIns;
                        best[0] = isLane0 ? rv : best[0];
                        code[0] = isLane0 ? rcode : code[0];
                        f[0] = isLane0 ? f0 : f[0];
                    }
#pragma unroll
                    for (int c = 0; c < C; c++)
                    if (last && c == cRO) { // RO (cpp:1741-1742, 1750, 1764)
                        double rv = uw[c] + lpnRO;
                        uint32_t rcode = 0;
                        if (t1[c] > rv + K1_EPS) { rv = t1[c]; rcode = J; }
                        const double fr = uw[c] + lNoInsgIns;
                        best[c] = isLaneRO ? rv : best[c];
                        code[c] = isLaneRO ? rcode : code[c];
                        f[c] = isLaneRO ? fr : f[c];
                    }
                    // plane 1 (cpp:1754-1767)
#pragma unroll
                    for (int c = 0; c < C; c++) {
                        const double stay = t1[c] + lInsgIns;
                        const bool p = (f[c] >= stay);
                        a1[c] = p ? f[c] : stay;
                        bit1[c] = p ? 1u : 0u;
                    }
#pragma unroll
                    for (int c = 0; c < C; c++) {
                        P0[c] = best[c];
                        P1[c] = a1[c];
                    }
                    g_codes[((size_t)s * btRows + b) * 32 + lane] = (uint16_t)(code[0] | (code[1] << 4) | (code[2] << 8) | (code[3] << 12)); // btb[b-1] at row b
#pragma unroll
                    for (int c = 0; c < C; c++) {
                        const unsigned m = __ballot_sync(FULLMASK, bit1[c] != 0);
                        if (lane == c) g_bits[((size_t)s * btRows + b) * 4 + c] = m;
                    }
                    if (!first && lane < 3) { // beta[b-1] of this strip's left edge
                        double *ew = edgeW + (size_t)(t + 1 - t0) * KS_EDGE;
#pragma unroll
                        for (int c = 0; c < C; c++) ew[lane * C + c] = P0[c];
                    }
                }
                } // rows of the block
                if (t1 >= nRight) {
                    // meeting-row sums by state column: (alpha[bMid] + obs[bMid]) + beta[bMid]  (cpp:1098), in place
#pragma unroll
                    for (int c = 0; c < C; c++) {
                        sm_ao[c0 + c] = sm_ao[c0 + c] + P0[c];
                        sm_ao[AO1 + c0 + c] = sm_ao[AO1 + c0 + c] + P1[c];
                    }
                } else {
#pragma unroll
                    for (int c = 0; c < C; c++) {
                        __stcg(g_save + ((s * 2 + 0) * C + c) * 32 + lane, P0[c]);
                        __stcg(g_save + ((s * 2 + 1) * C + c) * 32 + lane, P1[c]);
                    }
                }
                __syncwarp();
            }
            }
            // =================== meeting row (cpp:1075-1144) ===================
            KsPrior Q;
            Q.pLO0 = L->priorLO[0][mapq]; Q.pLO1 = L->priorLO[1][mapq]; Q.pOn0 = L->priorOn[0][mapq]; Q.pOn1 = L->priorOn[1][mapq];
            Q.hLO0 = L->priorLO_H[0]; Q.hLO1 = L->priorLO_H[1]; Q.hOn0 = L->priorOn_H[0]; Q.hOn1 = L->priorOn_H[1];
            Q.p1m = L->prior1m[mapq]; Q.p1mH = L->prior1m_H; Q.pIns0 = L->priorIns[0]; Q.pIns1 = L->priorIns[1];
            Q.insMode = DINDEL_INS_NONE; Q.insAnchor = 0; Q.insMax = 1; Q.hapStart = (int)hapStart; Q.R = R; Q.bMid = bMid;
            Q.insLogp = nullptr; Q.ins0 = 0.0;
            if constexpr (INS) {
                Q.insMode = (int)B.read_ins_mode[r];
                if (Q.insMode != DINDEL_INS_NONE) {
                    const int lib = B.read_lib[r];
                    Q.insAnchor = B.read_ins_anchor[r];
                    Q.insLogp = B.lib_logp + B.lib_off[lib];
                    Q.insMax = (int)(B.lib_off[lib + 1] - B.lib_off[lib]);
                    Q.ins0 = B.lib_log95[lib];
                }
            }
            double ll, llH, llOn;
            int msR, msH;
            ks_scan<INS, false>(sm_ao, sm_ao + AO1, Q, S, RO, lane, ll, msR, &llOn);
            ks_scan<INS, true>(sm_ao, sm_ao + AO1, Q, S, RO, lane, llH, msH, nullptr);
            double llOff;
            {
                double a, b, c, d;
                ks_prior<INS>(Q, 0, RO, a, b, c, d);
                const double v0 = sm_ao[0] + a, v1 = sm_ao[AO1] + b;
                llOff = (v0 > v1) ? v0 : v1; // states 0 and S, plain '>' (cpp:1104-1105)
            }
            const int xr = msR >= S ? msR - S : msR, xh = msH >= S ? msH - S : msH;
            const int offHap = (xr == 0 || xr == RO), offHapHMQ = (xh == 0 || xh == RO);

            const uint16_t *codes = g_codes;
            const uint32_t *bits = g_bits;
            auto btf = [codes, bits, btRows](int row, int x) -> int {
                const int st = x >> 7, xl = x & (KS_W - 1), ln = xl >> 2, c = xl & 3;
                const size_t rr = (size_t)st * btRows + row;
                const uint32_t cd = (codes[rr * 32 + ln] >> (4 * c)) & 15u;
                const uint32_t b1 = (bits[rr * 4 + c] >> ln) & 1u;
                return (int)(cd | (b1 << 7));
            };
            __syncwarp();
            k1_tail(btf, lane, R, H, J, bMid, msH, hap, sm_ch, sm_ph, sm_ms, sm_eq, L, ll, llOn, llOff,
                    offHap, offHapHMQ, out, B.hpos ? B.hpos + pair * (long long)B.hpos_stride : nullptr);
        }
    }
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
extern "C" size_t k1_strip_slot_bytes(int maxR) { return ks_slot_bytes(maxR); }
extern "C" size_t k1_strip_slots(int num_sms) { return (size_t)num_sms * K1_STRIP_MINB * KS_WPC; }
extern "C" int k1_strip_supported(int maxS, int maxR)
{
    return maxS <= KS_MAX_STRIPS * KS_W && ks_smem_per_warp(maxR) * KS_WPC <= 150 * 1024;
}

template <int J, bool INS>
static cudaError_t strip_launch_ins(const K1Batch *B, int num_sms, int cls, cudaStream_t st)
{
    const size_t smem = ks_smem_per_warp(B->max_R) * KS_WPC;
    cudaError_t e = cudaFuncSetAttribute(k1_strip_kernel<J, K1_STRIP_MINB, INS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    int per_sm = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k1_strip_kernel<J, K1_STRIP_MINB, INS>, KS_WPC * 32, smem);
    if (e != cudaSuccess) return e;
    if (per_sm < 1) return cudaErrorInvalidConfiguration;
    if (per_sm > K1_STRIP_MINB) per_sm = K1_STRIP_MINB; // scratch slots are sized for this
    long long grid = (long long)num_sms * per_sm;
    long long need = (B->unit_end - B->unit_begin + KS_WPC - 1) / KS_WPC;
    if (grid > need) grid = need;
    if (grid < 1) grid = 1;
    k1_strip_kernel<J, K1_STRIP_MINB, INS><<<(int)grid, KS_WPC * 32, smem, st>>>(*B, cls);
    return cudaGetLastError();
}

extern "C" cudaError_t k1_strip_launch_class(const K1Batch *B, int num_sms, int cls, int J, cudaStream_t st)
{
    if (J != 11) return cudaErrorInvalidValue;
    return B->read_ins_mode ? strip_launch_ins<11, true>(B, num_sms, cls, st) : strip_launch_ins<11, false>(B, num_sms, cls, st);
}
