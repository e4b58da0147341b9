// k1_tiled.cu — K1, register-tiled variant: the production kernel for the common shapes.
//
// One warp owns one (read, haplotype) pair.  The 2*(H+2) states of a row live in registers:
// lane l holds C consecutive state columns x = l*C .. l*C+C-1 of both planes (plane 0 "not in
// insertion", plane 1 "in insertion"); row b depends only on row b-1, so a row is data-parallel
// and the J-wide look-back (left half) / look-ahead (right half) window is exchanged with warp
// shuffles.  Per-haplotype transition tables sit in registers for the duration of a half;
// back-pointers go to shared memory as 4-bit (J <= 7) or 8-bit codes; the meeting-row scan, the
// back-trace and the alignment summary follow (k1_tail.cuh).
//
// Bit-exactness: the reference's updateMax (ObservationModelFB.cpp:877-888) is order dependent.
// Inside each state the candidates are offered in the reference's source order and every sum is
// associated as written there; the two-branch rule is specialised per call site using only facts
// about the candidate index order (derivations in DESIGN.md "updateMax specialisation"):
//   * offered index below every index seen so far  ->  take iff new >= dest
//   * offered index above every index seen so far  ->  take iff new >  dest + 1e-10
//   * clamped duplicates of LO / RO can never be taken -> offered as -inf
// Requires: sequence bytes in [A-Z=] (one-hot accept masks), forceReadOnHaplotype off; other
// batches take the generic kernel (k1_generic.cu).
#include <stdlib.h>
#include <type_traits>
#include "k1_common.cuh"
#include "k1_tail.cuh"

#define K1_EPS 1e-10
#ifndef K1_TILED_MINB
#define K1_TILED_MINB 4
#endif
#ifndef K1_TILED_WPC
#define K1_TILED_WPC 4 // warps per CTA
#endif
#ifndef K1_TILED_CRO_MAXC
#define K1_TILED_CRO_MAXC 5
#endif
#ifndef K1_LEFT_TREE
#define K1_LEFT_TREE 1
#endif
#ifndef K1_FMAX
#define K1_FMAX 0 // bit mask: 'take iff v >= best' value selects as DMNMX (one FP64-pipe instruction instead of two ALU selects):
                  // 1 left-half pair leaves, 2 left-half tree nodes, 4 plane 1 of the left half, 8 plane 1 of the right half
#endif
#ifndef K1_LEFT_ADD
// shared-table classes: the left half's jump entries logProbError + (y-1)*lInsgIns (cpp:1791) are rebuilt by one DADD
// from the column's lpe and the compile-time constant (y-1) * -0.5 (lInsgIns is the constant -.5: api.cu, cpp:1664;
// the product is exact, so the sum is the table's) instead of a shared load per candidate: +2.9 % at J = 11
#define K1_LEFT_ADD 1
#endif
#ifndef K1_DIAG
#define K1_DIAG 0 // timing diagnostics only (WRONG results): bit 0 skips the left-half rows, bit 1 the right-half rows, bit 2 the back-trace + summary
#endif
#ifndef K1_ROW_UNROLL
#define K1_ROW_UNROLL 1 // J = 6 classes: row loops unrolled by this factor (experiment (b) of profiles/r2_k1_j6_experiments.md)
#endif
#ifndef K1_TILED_J11_MINB
#define K1_TILED_J11_MINB 3
#endif
#ifndef K1_TILED_TBL_MINC
#define K1_TILED_TBL_MINC 4 // J = 6 classes wider than this also take the shared-memory jump tables
#endif
#define K1_TILED_BTG_MAX_CTAS K1_TILED_J11_MINB // resident CTAs per SM of the classes that keep back-pointers in global memory
#define FULLMASK 0xffffffffu

// call f(integral_constant<int, cro>) for the run-time value cro in [0, C)
template <int C, int I = 0, class F>
__device__ __forceinline__ void static_for_cro(int cro, F &f)
{
    if constexpr (I < C) {
        if (cro == I) f(std::integral_constant<int, I>{});
        else static_for_cro<C, I + 1>(cro, f);
    }
}

template <int C, int J>
struct TiledCfg {
    static constexpr bool NIB = (J <= 7);
    static constexpr int U16 = NIB ? (C + 3) / 4 : (C + 1) / 2; // 16-bit back-pointer units per lane-row
};

__host__ __device__ inline size_t tiled_smem_per_warp(int C, int U16, int maxR, bool bt_global = false)
{
    size_t b = (bt_global ? 0 : (size_t)maxR * U16 * 64) // back-pointers
               + (size_t)maxR * 16       // eq / uq
               + (size_t)2 * 32 * C * 8  // alpha[bMid]+obs[bMid]
               + (size_t)maxR * 2        // map state
               + (size_t)maxR * 2;       // read chars + phreds
    return (b + 15) & ~(size_t)15;
}

__device__ __forceinline__ double shfl_up_d(double v, int d) { return __shfl_up_sync(FULLMASK, v, d); }
__device__ __forceinline__ double shfl_down_d(double v, int d) { return __shfl_down_sync(FULLMASK, v, d); }

// Opaque copy: keeps a loop-invariant table entry in its register (the optimiser otherwise sinks
// the entry's construction -- an add and a select -- into the row loop to save registers).
__device__ __forceinline__ double pin_d(double v)
{
    asm volatile("" : "+d"(v));
    return v;
}

__device__ __forceinline__ double warp_max_d(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        double w = __shfl_xor_sync(FULLMASK, v, o);
        v = (w > v) ? w : v;
    }
    return v;
}

// accept mask of a state column: which read symbols (bit = char & 31) emit "match"
// (ObservationModelFB.cpp:237-250: LO, RO and hap 'N' always match).
__device__ __forceinline__ uint32_t accept_mask(const uint8_t *hap, int x, int H)
{
    if (x < 1 || x > H) return 0xffffffffu;
    uint8_t c = hap[x - 1];
    return (c == 'N') ? 0xffffffffu : (1u << (c & 31));
}

// Sequential "v > best + EPS, first wins" scan over all 2S states (cpp:1096-1116), done exactly
// without serialising over the states: only states within 1e-9 of the maximum can ever be taken
// once the first of them has been seen (and that one is taken, provided nothing lies in the
// guard band [M-2e-9, M-1e-9)); those few are then replayed in state order.  If the guard band
// is populated every state is replayed.
template <int C>
__device__ __forceinline__ void scan_first_wins(const double (&v0)[C], const double (&v1)[C], int lane,
                                                int S, double &best, int &bestState)
{
    double m = K1_NEG_INF;
#pragma unroll
    for (int c = 0; c < C; c++) {
        m = (v0[c] > m) ? v0[c] : m;
        m = (v1[c] > m) ? v1[c] : m;
    }
    m = warp_max_d(m);
    const double tNear = m - 1e-9, tGuard = m - 2e-9;
    uint32_t near = 0, guard = 0, valid = 0;
#pragma unroll
    for (int c = 0; c < C; c++) {
        const bool ok = (lane * C + c) < S;
        if (ok) valid |= (1u << c) | (1u << (C + c));
        if (ok && v0[c] >= tNear) near |= 1u << c;
        if (ok && v1[c] >= tNear) near |= 1u << (C + c);
        if (ok && v0[c] < tNear && v0[c] >= tGuard) guard |= 1u << c;
        if (ok && v1[c] < tNear && v1[c] >= tGuard) guard |= 1u << (C + c);
    }
    if (__any_sync(FULLMASK, guard != 0)) near = valid;
    best = K1_NEG_INF;
    bestState = 0;
    for (;;) {
        int bit = near ? (__ffs(near) - 1) : -1;
        int key = 0x7fffffff;
        if (bit >= 0) key = (bit < C) ? (lane * C + bit) : (S + lane * C + (bit - C));
        const int kmin = __reduce_min_sync(FULLMASK, key);
        if (kmin == 0x7fffffff) break;
        double vv = 0.0;
#pragma unroll
        for (int c = 0; c < C; c++) {
            if (bit == c) vv = v0[c];
            if (bit == C + c) vv = v1[c];
        }
        const int owner = __ffs(__ballot_sync(FULLMASK, key == kmin)) - 1;
        vv = __shfl_sync(FULLMASK, vv, owner);
        if (vv > best + K1_EPS) {
            best = vv;
            bestState = kmin;
        }
        if (lane == owner) near &= near - 1;
    }
}

// MINB: resident CTAs per SM the register budget is set for; BTG: back-pointer rows in a per-warp slot of
// global memory (L2 resident) instead of shared memory (long reads x wide classes: 250 x 352 columns = 96 KB).
// INS: some read of the batch carries an insert-size prior (--libFile); kept out of the common instantiation
// (its presence in the meeting row costs the row loops 4 % through register allocation).
template <int C, int J, int MINB, bool BTG, bool INS>
__global__ void __launch_bounds__(32 * K1_TILED_WPC, MINB) k1_tiled_kernel(K1Batch B, int cls)
{
    using Cfg = TiledCfg<C, J>;
    constexpr int U16 = Cfg::U16;
    // TBL (J > 7): the per-column jump tables lp(x, y) do not fit the register file next to the row windows
    // (C*(2J-1) doubles); they are kept once per CTA in shared memory, indexed by the homopolymer code of the
    // column that owns the entry, and the "-inf outside the haplotype" entries become -inf halo columns instead.
    constexpr bool TBL = (J > 7) || (C > K1_TILED_TBL_MINC);
    constexpr int kRowUnroll = (J <= 7) ? K1_ROW_UNROLL : 1; // the J = 11 loops are 20 KB each as they are
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ double s_hp_lpe[K1_HP_LUT], s_hp_lpn[K1_HP_LUT];
    __shared__ double s_tl[(TBL && !K1_LEFT_ADD) ? K1_HP_LUT * (J - 1) : 1]; // [code][y-2] = logProbError + (y-1)*lInsgIns (cpp:1791)
    // Row pitch of s_tr: EVEN, so that entries (y, y+1) of a row form a 16-byte aligned pair.  Cell c at jump length y+1
    // uses the neighbour of what cell c+1 uses at length y, and with aligned rows the compiler fetches both with one
    // LDS.128 (60 of them in <10,11>): J = 11 with pitch 11 -> 12 alone took configs[3] from 153.2 to 161.9 GCUPS.
    constexpr int JP = (J + 1) & ~1;
    __shared__ __align__(16) double s_tr[TBL ? K1_HP_LUT * JP : 1]; // [code][y-1] = lp(y) + logProbNoError (cpp:1732-1735)
    const int lane = threadIdx.x & 31;
    const int wib = threadIdx.x >> 5;
    const K1Luts *__restrict__ L = B.luts;
    const int maxR = B.max_R;
    for (int i = threadIdx.x; i < K1_HP_LUT; i += blockDim.x) {
        s_hp_lpe[i] = L->hp_lpe[i];
        s_hp_lpn[i] = L->hp_lpn[i];
    }
    if constexpr (TBL) {
        for (int i = threadIdx.x; !K1_LEFT_ADD && i < K1_HP_LUT * (J - 1); i += blockDim.x) {
            const int code = i / (J - 1), y = i % (J - 1) + 2;
            s_tl[i] = L->hp_lpe[code] + (double)(y - 1) * L->lInsgIns;
        }
        for (int i = threadIdx.x; i < K1_HP_LUT * JP; i += blockDim.x) {
            const int code = i / JP, y = i % JP + 1;
            const double lpn_s = L->hp_lpn[code], lpe_s = L->hp_lpe[code];
            const double lp = (y == 1) ? lpn_s : (lpe_s + (double)(y - 1) * L->lInsgIns);
            s_tr[i] = (y <= J) ? lp + lpn_s : 0.0;
        }
    }
    __syncthreads();
    unsigned char *base = smem_raw + tiled_smem_per_warp(C, U16, maxR, BTG) * wib;
    uint16_t *sm_bt = BTG ? reinterpret_cast<uint16_t *>(B.bt_scratch + ((long long)blockIdx.x * K1_TILED_WPC + wib) * B.bt_slot_bytes)
                          : reinterpret_cast<uint16_t *>(base);
    double *sm_eq = reinterpret_cast<double *>(base + (BTG ? 0 : (size_t)maxR * U16 * 64));
    double *sm_uq = sm_eq + maxR;
    double *sm_ao = sm_uq + maxR; // [2][32*C]
    int16_t *sm_ms = reinterpret_cast<int16_t *>(sm_ao + 2 * 32 * C);
    uint8_t *sm_ch = reinterpret_cast<uint8_t *>(sm_ms + maxR);
    uint8_t *sm_ph = sm_ch + maxR;

    const double lLOgLO = L->lLOgLO, lFirstgLO = L->lFirstgLO, lInsgIns = L->lInsgIns,
                 lNoInsgIns = L->lNoInsgIns, lNoInsgNoIns = L->lNoInsgNoIns;
    const int c0 = lane * C;

    // Units of this class, sorted by RO slot: the warps of the GPU walk the list together, so at
    // any time (nearly) all of them run the same instantiation of the row loops (I-cache).
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
        const uint8_t *hap = B.hap_seq + hoff;
        const uint8_t *hp = B.hap_hp + hoff + 2LL * (h0 + h);
        const uint32_t hapStart = B.hap_start[w];
        const bool hapsize_err = L->maxLengthDel > H;
        const int rb = r0 + chunk * K1_READ_CHUNK;
        const int re = min(rb + K1_READ_CHUNK, r0 + nr);

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

            double P0[C], P1[C];
            uint32_t am[C];
#pragma unroll
            for (int c = 0; c < C; c++) am[c] = accept_mask(hap, c0 + c, H);

            // Column RO = H+1 sits at a warp-uniform register slot cRO of lane lRO; its special
            // transitions are evaluated once per row under a warp-uniform test.
            const int lRO = RO / C, cRO = RO - lRO * C;
            const bool isLane0 = (lane == 0), isLaneRO = (lane == lRO);

            // The row loops are instantiated once per RO slot (K1_TILED_CRO_MAXC: classes up to that
            // many columns per lane), so RO's transitions are straight-line code inside the cells'
            // basic block; wider classes test the slot at run time (CRO = -1).
            // =================== left half: rows 1..bMid (passMessageTwoDec) ===================
            // Per row: a window W*[k] of the previous row over columns c0-J .. c0+C-1 (own columns at
            // k = J+c, the J columns to the left fetched from the neighbour lanes), then the C cells'
            // candidate chains advanced in lock step (y outer, c inner) so that the chains of the C
            // cells interleave in the instruction stream.
            {
                double lpn[C], lpe[C], lpD[TBL ? 1 : C][TBL ? 1 : J - 1];
                int offL[TBL ? C : 1];
#pragma unroll
                for (int c = 0; c < C; c++) {
                    const int x = c0 + c;
                    const int code = hp[x <= RO ? x : RO];
                    lpe[c] = s_hp_lpe[code];
                    lpn[c] = s_hp_lpn[code];
                    if constexpr (TBL) {
                        offL[c] = code * (J - 1);
                    } else {
#pragma unroll
                        for (int y = 2; y <= J; y++) // cpp:1791; a source left of LO can never win -> -inf
                            lpD[c][y - 2] = pin_d((x - y >= 0) ? lpe[c] + (double)(y - 1) * lInsgIns : K1_NEG_INF);
                    }
                    P0[c] = 0.0; // alpha[0] = 0 (cpp:335-338)
                    P1[c] = 0.0;
                }
                // lpe of the column left of this lane's first (RO's "from S+H" candidate, cpp:1805)
                const double lpeLm1 = s_hp_lpe[hp[(c0 >= 1 && c0 - 1 <= RO) ? c0 - 1 : 0]];

                auto LPD = [&](int c, int k) -> double {
#if K1_LEFT_ADD
                    if constexpr (TBL) return lpe[c] + (-0.5 * (double)(k + 1));
#else
                    if constexpr (TBL) return s_tl[offL[c] + k];
#endif
                    else return lpD[c][k];
                };
                auto left_rows = [&](auto cro_tag) {
                constexpr int CRO = decltype(cro_tag)::value;
#pragma unroll kRowUnroll
                for (int b = 1; b <= ((K1_DIAG & 1) ? 0 : bMid); b++) {
                    const double eq = sm_eq[b - 1], uq = sm_uq[b - 1];
                    const uint32_t rbit = 1u << (sm_ch[b - 1] & 31);
                    double WO[C + J], WP[C + J], u[C], t1w[C + 1];
#pragma unroll
                    for (int c = 0; c < C; c++) {
                        WO[J + c] = (am[c] & rbit) ? eq : uq;
                        WP[J + c] = P0[c];
                        u[c] = WO[J + c] + P0[c];
                        t1w[c + 1] = eq + P1[c];
                    }
#pragma unroll
                    for (int j = 1; j <= J; j++) { // lanes without a left neighbour: table entry is -inf
                        const int d = (j + C - 1) / C;
                        const int ci = d * C - j;
                        WP[J - j] = shfl_up_d(WP[J + ci], d);
                        WO[J - j] = shfl_up_d(WO[J + ci], d);
                        if constexpr (TBL) WP[J - j] = (lane < d) ? K1_NEG_INF : WP[J - j]; // column c0-j < 0
                    }
                    t1w[0] = shfl_up_d(t1w[C], 1); // lane 0: only LO would use it (overridden)

                    double best[C], a1[C];
                    uint32_t code[C], bit1[C];
                    // jumps y = 1..J from column x-y (cpp:1784-1795)
#if K1_LEFT_TREE
                    // "take iff v >= best" over y = 1..J picks the LAST y attaining the maximum; the same winner
                    // comes out of any reduction tree whose nodes prefer the later operand on ties, so the
                    // candidates are reduced pairwise (depth 1 + J/2 instead of J dependent compare-selects).
#pragma unroll
                    for (int y = 1; y <= J; y += 2) {
#pragma unroll
                        for (int c = 0; c < C; c++) {
                            const double va = (y == 1) ? ((WO[J + c - 1] + lpn[c]) + WP[J + c - 1]) + lpn[c]
                                                       : ((WO[J + c - y] + LPD(c, (y >= 2) ? y - 2 : 0)) + WP[J + c - y]) + lpn[c];
                            double pm = va;
                            uint32_t pc = (uint32_t)(y - 1);
                            if (y + 1 <= J) {
                                const double vb = ((WO[J + c - (y + 1)] + LPD(c, y - 1)) + WP[J + c - (y + 1)]) + lpn[c];
                                const bool pb = (vb >= va);
#if K1_FMAX & 1
                                pm = fmax(va, vb); // == pb ? vb : va bit for bit (no NaN, no zero of either sign among these sums)
#else
                                pm = pb ? vb : va;
#endif
                                pc = pb ? (uint32_t)y : (uint32_t)(y - 1);
                            }
                            if (y == 1) {
                                best[c] = pm;
                                code[c] = pc;
                            } else {
                                const bool p = (pm >= best[c]);
#if K1_FMAX & 2
                                best[c] = fmax(best[c], pm);
#else
                                best[c] = p ? pm : best[c];
#endif
                                code[c] = p ? pc : code[c];
                            }
                        }
                    }
#else
#pragma unroll
                    for (int c = 0; c < C; c++) {
                        best[c] = ((WO[J + c - 1] + lpn[c]) + WP[J + c - 1]) + lpn[c];
                        code[c] = 0;
                    }
#pragma unroll
                    for (int y = 2; y <= J; y++) {
#pragma unroll
                        for (int c = 0; c < C; c++) {
                            const double v = ((WO[J + c - y] + LPD(c, y - 2)) + WP[J + c - y]) + lpn[c];
                            const bool p = (v >= best[c]);
                            best[c] = p ? v : best[c];
                            code[c] = p ? (uint32_t)(y - 1) : code[c];
                        }
                    }
#endif
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
#if K1_FMAX & 4
                        a1[c] = fmax(stay, f);
#else
                        a1[c] = p ? f : stay;
#endif
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
                    for (int c = 0; c < C; c++) {
                        if (CRO < 0 ? (c == cRO) : (c == CRO)) {
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
                    }
#pragma unroll
                    for (int c = 0; c < C; c++) {
                        P0[c] = best[c];
                        P1[c] = a1[c];
                    }
                    uint16_t *row = sm_bt + (size_t)b * U16 * 32;
#pragma unroll
                    for (int k = 0; k < U16; k++) {
                        uint32_t v = 0;
                        if (Cfg::NIB) {
#pragma unroll
                            for (int q = 0; q < 4; q++)
                                if (4 * k + q < C) {
                                    const int cc = (4 * k + q < C) ? 4 * k + q : 0;
                                    v += (code[cc] + 8u * bit1[cc]) << (4 * q);
                                }
                        } else {
#pragma unroll
                            for (int q = 0; q < 2; q++)
                                if (2 * k + q < C) {
                                    const int cc = (2 * k + q < C) ? 2 * k + q : 0;
                                    v += (code[cc] + 128u * bit1[cc]) << (8 * q);
                                }
                        }
                        row[k * 32 + lane] = (uint16_t)v;
                    }
                }
                }; // left_rows
                if (C <= K1_TILED_CRO_MAXC) static_for_cro<C>(cRO, left_rows);
                else left_rows(std::integral_constant<int, -1>{});
            }
            // alpha[bMid] + obs[bMid] (first add of cpp:1098) -> shared, by state column
            {
                const double eq = sm_eq[bMid], uq = sm_uq[bMid];
                const uint32_t rbit = 1u << (sm_ch[bMid] & 31);
#pragma unroll
                for (int c = 0; c < C; c++) {
                    sm_ao[c0 + c] = P0[c] + ((am[c] & rbit) ? eq : uq);
                    sm_ao[32 * C + c0 + c] = P1[c] + eq;
                }
            }
            // =================== right half: rows R-2..bMid (passMessageTwoInc) ===================
            // Window W*[k] over columns c0 .. c0+C-1+J (own at k = c, the J columns to the right
            // from the neighbour lanes).
            {
                double cI[TBL ? 1 : C][TBL ? 1 : J], lpeN[C];
                int offR[TBL ? C + J : 1]; // [k]: table row of column c0+k (clamped to RO), k = 1 .. C+J-1
                if constexpr (TBL) {
#pragma unroll
                    for (int k = 1; k < C + J; k++) offR[k] = (int)hp[(c0 + k <= RO) ? c0 + k : RO] * JP;
                }
#pragma unroll
                for (int c = 0; c < C; c++) {
                    const int x = c0 + c;
                    if constexpr (TBL) {
                        lpeN[c] = s_hp_lpe[hp[(x + 1 <= RO) ? x + 1 : RO]]; // logProbError[x+1] (cpp:1748)
                    } else {
#pragma unroll
                        for (int y = 1; y <= J; y++) {
                            const int nx = (x + y <= RO) ? x + y : RO; // cpp:1728-1729
                            const int code = hp[nx];
                            const double lpn_s = s_hp_lpn[code], lpe_s = s_hp_lpe[code];
                            const double lp = (y == 1) ? lpn_s : (lpe_s + (double)(y - 1) * lInsgIns); // cpp:1732
                            cI[c][y - 1] = pin_d((x + y <= RO) ? lp + lpn_s : K1_NEG_INF); // duplicates of RO never win
                            if (y == 1) lpeN[c] = lpe_s; // logProbError[x+1] (cpp:1748)
                        }
                    }
                    P0[c] = (x <= RO) ? 0.0 : K1_NEG_INF; // beta[R-1] = 0; columns beyond RO never win
                    P1[c] = (x <= RO) ? 0.0 : K1_NEG_INF;
                }
                const double lpnRO = s_hp_lpn[hp[RO]];

                auto CI = [&](int c, int k) -> double { // k = y-1; target column x+y = c0 + c + k + 1
                    if constexpr (TBL) return s_tr[offR[c + k + 1] + k];
                    else return cI[c][k];
                };
                auto right_rows = [&](auto cro_tag) {
                constexpr int CRO = decltype(cro_tag)::value;
#pragma unroll kRowUnroll
                for (int b = ((K1_DIAG & 2) ? bMid : R - 1); b > bMid; b--) {
                    const double eq = sm_eq[b], uq = sm_uq[b];
                    const uint32_t rbit = 1u << (sm_ch[b] & 31);
                    double WO[C + J], WP[C + J], uw[C + 1], t1[C];
#pragma unroll
                    for (int c = 0; c < C; c++) {
                        WO[c] = (am[c] & rbit) ? eq : uq;
                        WP[c] = P0[c];
                        uw[c] = WO[c] + P0[c];
                        t1[c] = eq + P1[c];
                    }
#pragma unroll
                    for (int j = 1; j <= J; j++) { // beyond lane 31: table entry is -inf
                        const int d = (C - 1 + j) / C;
                        const int ci = (C - 1 + j) % C;
                        WP[C - 1 + j] = shfl_down_d(WP[ci], d);
                        WO[C - 1 + j] = shfl_down_d(WO[ci], d);
                        if constexpr (TBL) WP[C - 1 + j] = (lane + d > 31) ? K1_NEG_INF : WP[C - 1 + j]; // beyond lane 31
                    }
                    uw[C] = shfl_down_d(uw[0], 1); // lane 31: column is RO or padding (overridden)

                    double best[C], a1[C];
                    uint32_t code[C], bit1[C];
                    // jumps y = 1..J to column x+y (cpp:1724-1737)
#pragma unroll
                    for (int c = 0; c < C; c++) {
                        best[c] = (CI(c, 0) + WP[c + 1]) + WO[c + 1];
                        code[c] = 0;
                    }
#pragma unroll
                    for (int y = 2; y <= J; y++) {
#pragma unroll
                        for (int c = 0; c < C; c++) {
                            const double v = (CI(c, y - 1) + WP[c + y]) + WO[c + y];
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
                    for (int c = 0; c < C; c++) { // RO (cpp:1741-1742, 1750, 1764)
                        if (CRO < 0 ? (c == cRO) : (c == CRO)) {
                            double rv = uw[c] + lpnRO;
                            uint32_t rcode = 0;
                            if (t1[c] > rv + K1_EPS) { rv = t1[c]; rcode = J; }
                            const double fr = uw[c] + lNoInsgIns;
                            best[c] = isLaneRO ? rv : best[c];
                            code[c] = isLaneRO ? rcode : code[c];
                            f[c] = isLaneRO ? fr : f[c];
                        }
                    }
                    // plane 1 (cpp:1754-1767)
#pragma unroll
                    for (int c = 0; c < C; c++) {
                        const double stay = t1[c] + lInsgIns;
                        const bool p = (f[c] >= stay);
#if K1_FMAX & 8
                        a1[c] = fmax(stay, f[c]);
#else
                        a1[c] = p ? f[c] : stay;
#endif
                        bit1[c] = p ? 1u : 0u;
                    }
#pragma unroll
                    for (int c = 0; c < C; c++) {
                        P0[c] = best[c];
                        P1[c] = a1[c];
                    }
                    uint16_t *row = sm_bt + (size_t)b * U16 * 32; // btb[b-1] stored at row b
#pragma unroll
                    for (int k = 0; k < U16; k++) {
                        uint32_t v = 0;
                        if (Cfg::NIB) {
#pragma unroll
                            for (int q = 0; q < 4; q++)
                                if (4 * k + q < C) {
                                    const int cc = (4 * k + q < C) ? 4 * k + q : 0;
                                    v += (code[cc] + 8u * bit1[cc]) << (4 * q);
                                }
                        } else {
#pragma unroll
                            for (int q = 0; q < 2; q++)
                                if (2 * k + q < C) {
                                    const int cc = (2 * k + q < C) ? 2 * k + q : 0;
                                    v += (code[cc] + 128u * bit1[cc]) << (8 * q);
                                }
                        }
                        row[k * 32 + lane] = (uint16_t)v;
                    }
                }
                }; // right_rows
                if (C <= K1_TILED_CRO_MAXC) static_for_cro<C>(cRO, right_rows);
                else right_rows(std::integral_constant<int, -1>{});
            }
            __syncwarp();
            // =================== meeting row (cpp:1075-1144) ===================
            double vR0[C], vR1[C], vH0[C], vH1[C];
            double onMax = K1_NEG_INF;
            {
                const double pLO0 = L->priorLO[0][mapq], pLO1 = L->priorLO[1][mapq];
                const double pOn0 = L->priorOn[0][mapq], pOn1 = L->priorOn[1][mapq];
                const double hLO0 = L->priorLO_H[0], hLO1 = L->priorLO_H[1];
                const double hOn0 = L->priorOn_H[0], hOn1 = L->priorOn_H[1];
                // insert-size prior of this read (cpp:279-293): warp-uniform, rare (only with --libFile)
                int insMode = DINDEL_INS_NONE, insAnchor = 0, insMax = 1;
                const double *insLogp = nullptr;
                double ins0 = 0.0;
                if constexpr (INS) {
                    insMode = (int)B.read_ins_mode[r];
                    if (insMode != DINDEL_INS_NONE) {
                        const int lib = B.read_lib[r];
                        insAnchor = B.read_ins_anchor[r];
                        insLogp = B.lib_logp + B.lib_off[lib];
                        insMax = (int)(B.lib_off[lib + 1] - B.lib_off[lib]);
                        ins0 = B.lib_log95[lib];
                    }
                }
#pragma unroll
                for (int c = 0; c < C; c++) {
                    const int x = c0 + c;
                    const double t0 = sm_ao[x] + P0[c];
                    const double t1v = sm_ao[32 * C + x] + P1[c];
                    const bool lo = (x == 0), ro = (x == RO), pad = (x > RO);
                    double qR0 = lo ? pLO0 : pOn0, qR1 = lo ? pLO1 : pOn1, qH0 = lo ? hLO0 : hOn0, qH1 = lo ? hLO1 : hOn1;
                    if (INS && insMode != DINDEL_INS_NONE && !ro && !pad) {
                        if (lo) { // (log(pOff) + logpIns) + pinsert[0]  (cpp:298)
                            qR0 = pLO0 + ins0; qR1 = pLO1 + ins0; qH0 = hLO0 + ins0; qH1 = hLO1 + ins0;
                        } else {  // (pinsert[x] + log(1 - pOff)) + logpIns  (cpp:301)
                            const double pin = k1_pinsert(insLogp, insMax, insMode, insAnchor, (int)hapStart, x, R, bMid);
                            const double aR = pin + L->prior1m[mapq], aH = pin + L->prior1m_H;
                            qR0 = aR + L->priorIns[0]; qR1 = aR + L->priorIns[1];
                            qH0 = aH + L->priorIns[0]; qH1 = aH + L->priorIns[1];
                        }
                    }
                    vR0[c] = pad ? K1_NEG_INF : t0 + (ro ? -100.0 : qR0);
                    vR1[c] = pad ? K1_NEG_INF : t1v + (ro ? -100.0 : qR1);
                    vH0[c] = pad ? K1_NEG_INF : t0 + (ro ? -100.0 : qH0);
                    vH1[c] = pad ? K1_NEG_INF : t1v + (ro ? -100.0 : qH1);
                    if (x >= 1 && x <= H) {
                        onMax = (vR0[c] > onMax) ? vR0[c] : onMax;
                        onMax = (vR1[c] > onMax) ? vR1[c] : onMax;
                    }
                }
            }
            double ll, llH;
            int msR, msH;
            scan_first_wins<C>(vR0, vR1, lane, S, ll, msR);
            scan_first_wins<C>(vH0, vH1, lane, S, llH, msH);
            const double llOn = warp_max_d(onMax);
            double llOff = (vR0[0] > vR1[0]) ? vR0[0] : vR1[0]; // states 0 and S, plain '>' (cpp:1104-1105)
            llOff = __shfl_sync(FULLMASK, llOff, 0);
            const int xr = msR >= S ? msR - S : msR, xh = msH >= S ? msH - S : msH;
            const int offHap = (xr == 0 || xr == RO), offHapHMQ = (xh == 0 || xh == RO);

            const uint16_t *btp = sm_bt;
            auto btf = [btp](int row, int x) -> int {
                if constexpr (BTG) { // rows in global memory: 64-bit row offsets
                    const int ln = x / C, c = x - ln * C;
                    if (Cfg::NIB) {
                        const uint32_t u16 = btp[((size_t)row * U16 + (c >> 2)) * 32 + ln];
                        const uint32_t v = (u16 >> (4 * (c & 3))) & 15u;
                        return (int)((v & 7u) | ((v >> 3) << 7));
                    } else {
                        const uint32_t u16 = btp[((size_t)row * U16 + (c >> 1)) * 32 + ln];
                        return (int)((u16 >> (8 * (c & 1))) & 255u);
                    }
                } else {             // rows in shared memory: unsigned 32-bit index arithmetic (x >= 0)
                    const unsigned ln = (unsigned)x / (unsigned)C, c = (unsigned)x - ln * (unsigned)C;
                    if (Cfg::NIB) {
                        const uint32_t u16 = btp[((unsigned)row * U16 + (c >> 2)) * 32 + ln];
                        const uint32_t v = (u16 >> (4 * (c & 3))) & 15u;
                        return (int)((v & 7u) | ((v >> 3) << 7));
                    } else {
                        const uint32_t u16 = btp[((unsigned)row * U16 + (c >> 1)) * 32 + ln];
                        return (int)((u16 >> (8 * (c & 1))) & 255u);
                    }
                }
            };
            int16_t *hrow = B.hpos ? B.hpos + pair * (long long)B.hpos_stride : nullptr;
            if constexpr ((K1_DIAG & 4) != 0) { // diagnostics: no back-trace / summary
                if (lane == 0) { out->ll = ll; out->llOn = llOn; out->llOff = llOff; out->mapState0 = (uint16_t)msH; out->offHap = (uint8_t)offHap; out->offHapHMQ = (uint8_t)offHapHMQ; }
            } else
            if constexpr (BTG) { // rows in global memory: prefetch down the diagonal (k1_tail.cuh)
                auto pff = [btp](int row, int x) {
                    const int ln = x / C, c = x - ln * C;
                    const uint16_t *a = btp + ((size_t)row * U16 + (Cfg::NIB ? (c >> 2) : (c >> 1))) * 32 + ln;
                    asm volatile("prefetch.global.L1 [%0];" ::"l"(a));
                };
                k1_tail(btf, pff, lane, R, H, J, bMid, msH, hap, sm_ch, sm_ph, sm_ms, sm_eq, L, ll, llOn, llOff, offHap, offHapHMQ, out, hrow);
            } else {
                k1_tail(btf, lane, R, H, J, bMid, msH, hap, sm_ch, sm_ph, sm_ms, sm_eq, L, ll, llOn, llOff, offHap, offHapHMQ, out, hrow);
            }
        }
    }
}

// ------------------------------------------------------------------------------------------
// host side: shape classes and launches
// ------------------------------------------------------------------------------------------
// Shape classes = state columns per lane.  J = 6 (CLI default maxLengthIndel 5): six classes up to S = 256.
// J = 11 (maxLengthDel 10, configs[3]): three wide classes up to S = 352, back-pointers in global memory.
static const int kClassC6[6] = {2, 3, 4, 5, 6, 8};
static const int kClassC11[6] = {8, 10, 11, 0, 0, 0};
// J = 11 through the strip kernel (k1_strip.cu): classes = 1, 2, 3 strips of 128 columns (4 per lane and strip)
static const int kClassStrip11[6] = {4, 8, 12, 0, 0, 0};
extern "C" size_t k1_strip_slot_bytes(int maxR);
extern "C" size_t k1_strip_slots(int num_sms);
extern "C" int k1_strip_supported(int maxS, int maxR);
extern "C" cudaError_t k1_strip_launch_class(const K1Batch *B, int num_sms, int cls, int J, cudaStream_t st);
// J = 11 runs on the 8/10/11-columns-per-lane classes of this file; $DINDEL_B200_J11_STRIP=1 (read at every call, so a
// test can switch between two engines of one process) takes the strip kernel instead.  Measured on configs[3]
// (profiles/r2_k1_strip.md): strip 106.5 GCUPS against 133.6 here -- its strips pad S ~ 302 to 384 columns and pay
// the boundary selects every row, 6.9 warp instructions per cell against 4.7.
static bool use_strip(int J)
{
    if (J != 11) return false;
    const char *e = getenv("DINDEL_B200_J11_STRIP");
    return e && e[0] == '1';
}
static const int *class_table(int J) { return J == 6 ? kClassC6 : (J == 11 ? (use_strip(J) ? kClassStrip11 : kClassC11) : nullptr); }
extern "C" void k1_tiled_class_table(int J, int32_t out[6])
{
    const int *t = class_table(J);
    for (int i = 0; i < 6; i++) out[i] = t ? t[i] : 0;
}

__device__ __forceinline__ int tiled_class_of_dev(const K1Batch &B, int S)
{
#pragma unroll
    for (int i = 0; i < 6; i++)
        if (B.class_c[i] > 0 && S <= 32 * B.class_c[i]) return i;
    return -1;
}

// Work-unit lists: count (pass 0) and fill (pass 1), one thread per window; in between one thread
// turns the counts into bucket offsets.  A haplotype outside every class (S > 256) is never listed:
// such batches take the generic kernel.
__global__ void k1_tiled_units(K1Batch B, int pass)
{
    const int w = B.w_begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= B.w_end) return;
    const int h0 = B.win_hap[w], nh = B.win_hap[w + 1] - h0;
    const int nr = B.win_read[w + 1] - B.win_read[w];
    const unsigned nchunk = (unsigned)((nr + K1_READ_CHUNK - 1) / K1_READ_CHUNK);
    if (nchunk == 0) return;
    for (int h = 0; h < nh; h++) {
        const int H = (int)(B.hap_off[h0 + h + 1] - B.hap_off[h0 + h]);
        const int cls = tiled_class_of_dev(B, H + 2);
        if (cls < 0) continue;
        const int bkt = cls * K1_TILED_SLOTS + (H + 1) % B.class_c[cls];
        if (pass == 0) {
            atomicAdd(&B.bucket_work[bkt], nchunk);
        } else {
            const unsigned at = B.bucket_start[bkt] + atomicAdd(&B.bucket_work[K1_TILED_BUCKETS + bkt], nchunk);
            for (unsigned ch = 0; ch < nchunk; ch++) B.units[at + ch] = make_uint2((unsigned)w, ((unsigned)h << 16) | ch);
        }
    }
}
__global__ void k1_tiled_units_scan(K1Batch B)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    unsigned acc = 0;
    for (int b = 0; b < K1_TILED_BUCKETS; b++) {
        B.bucket_start[b] = acc;
        acc += B.bucket_work[b];
        B.bucket_work[K1_TILED_BUCKETS + b] = 0;
    }
    B.bucket_start[K1_TILED_BUCKETS] = acc;
}

// Builds the bucketed unit lists of the resident batch (once per upload).  3 launches.
extern "C" cudaError_t k1_tiled_prepare(const K1Batch *B, cudaStream_t st, int *launches)
{
    if (launches) *launches = 0;
    if (B->w_end <= B->w_begin) return cudaSuccess;
    cudaError_t e = cudaMemsetAsync(B->bucket_work, 0, sizeof(unsigned) * 2 * K1_TILED_BUCKETS, st);
    if (e != cudaSuccess) return e;
    const int threads = 128, blocks = (B->w_end - B->w_begin + threads - 1) / threads;
    k1_tiled_units<<<blocks, threads, 0, st>>>(*B, 0);
    k1_tiled_units_scan<<<1, 32, 0, st>>>(*B);
    k1_tiled_units<<<blocks, threads, 0, st>>>(*B, 1);
    if (launches) *launches = 3;
    return cudaGetLastError();
}

template <int C, int J, int MINB, bool BTG, bool INS>
static cudaError_t launch_one_ins(const K1Batch *B, int num_sms, int cls, cudaStream_t st);

template <int C, int J, int MINB, bool BTG>
static cudaError_t launch_one(const K1Batch *B, int num_sms, int cls, cudaStream_t st)
{
    return B->read_ins_mode ? launch_one_ins<C, J, MINB, BTG, true>(B, num_sms, cls, st)
                            : launch_one_ins<C, J, MINB, BTG, false>(B, num_sms, cls, st);
}

template <int C, int J, int MINB, bool BTG, bool INS>
static cudaError_t launch_one_ins(const K1Batch *B, int num_sms, int cls, cudaStream_t st)
{
    constexpr int U16 = TiledCfg<C, J>::U16;
    const int warps = K1_TILED_WPC;
    const size_t smem = tiled_smem_per_warp(C, U16, B->max_R, BTG) * warps;
    cudaError_t e = cudaFuncSetAttribute(k1_tiled_kernel<C, J, MINB, BTG, INS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    int per_sm = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k1_tiled_kernel<C, J, MINB, BTG, INS>, warps * 32, smem);
    if (e != cudaSuccess) return e;
    if (per_sm < 1) return cudaErrorInvalidConfiguration;
    if (BTG && per_sm > K1_TILED_BTG_MAX_CTAS) per_sm = K1_TILED_BTG_MAX_CTAS; // scratch slots are sized for this
    long long grid = (long long)num_sms * per_sm;
    long long need = (B->unit_end - B->unit_begin + warps - 1) / warps;
    if (grid > need) grid = need;
    if (grid < 1) grid = 1;
    k1_tiled_kernel<C, J, MINB, BTG, INS><<<(int)grid, warps * 32, smem, st>>>(*B, cls);
    return cudaGetLastError();
}

// global back-pointer scratch the J > 7 classes need: one slot per resident warp of a full grid
extern "C" size_t k1_tiled_bt_slot_bytes(int maxR, int J)
{
    if (J <= 7) return 0;
    if (use_strip(J)) return k1_strip_slot_bytes(maxR);
    const int *t = class_table(J);
    int C = 0;
    for (int i = 0; t && i < 6; i++) if (t[i] > C) C = t[i];
    return (((size_t)maxR + 1) * ((C + 1) / 2) * 64 + 255) & ~(size_t)255;
}
extern "C" size_t k1_tiled_bt_slots(int num_sms)
{
    const size_t a = (size_t)num_sms * K1_TILED_BTG_MAX_CTAS * K1_TILED_WPC, b = k1_strip_slots(num_sms);
    return a > b ? a : b;
}

extern "C" int k1_tiled_supported(int maxS, int maxR, int J)
{
    const int *t = class_table(J);
    if (!t) return 0;
    if (use_strip(J)) return k1_strip_supported(maxS, maxR);
    int C = 0;
    for (int i = 0; i < 6; i++)
        if (t[i] > 0 && maxS <= 32 * t[i]) { C = t[i]; break; }
    if (C == 0) return 0;
    if (J > 7) return tiled_smem_per_warp(C, (C + 1) / 2, maxR, true) * K1_TILED_WPC <= 100 * 1024;
    // back-pointers of the largest class present must fit one CTA of 4 warps in shared memory
    return tiled_smem_per_warp(C, (C + 3) / 4, maxR) * K1_TILED_WPC <= 200 * 1024;
}

// One launch per shape class present (class = state columns per lane).  `class_mask` bit i set
// when some haplotype of the batch falls in class i.
extern "C" cudaError_t k1_tiled_launch(const K1Batch *B, int num_sms, cudaStream_t st, int *launches,
                                       unsigned class_mask, int J)
{
    int n = 0;
    cudaError_t e = cudaSuccess;
    const int *t = class_table(J);
    for (int i = 0; t && i < 6 && e == cudaSuccess; i++) {
        if (!(class_mask & (1u << i)) || t[i] == 0) continue;
        if (n == 0) {
            e = cudaMemsetAsync(B->counter, 0, 8 * sizeof(unsigned long long), st);
            if (e != cudaSuccess) break;
        }
        if (J == 6) {
            switch (t[i]) {
            case 2: e = launch_one<2, 6, K1_TILED_MINB, false>(B, num_sms, i, st); break;
            case 3: e = launch_one<3, 6, K1_TILED_MINB, false>(B, num_sms, i, st); break;
            case 4: e = launch_one<4, 6, K1_TILED_MINB, false>(B, num_sms, i, st); break;
            case 5: e = launch_one<5, 6, K1_TILED_MINB, false>(B, num_sms, i, st); break;
            case 6: e = launch_one<6, 6, K1_TILED_MINB, false>(B, num_sms, i, st); break;
            case 8: e = launch_one<8, 6, K1_TILED_MINB, false>(B, num_sms, i, st); break;
            }
        } else if (use_strip(J)) {
            e = k1_strip_launch_class(B, num_sms, i, J, st);
        } else {
            switch (t[i]) {
            case 8: e = launch_one<8, 11, K1_TILED_J11_MINB, true>(B, num_sms, i, st); break;
            case 10: e = launch_one<10, 11, K1_TILED_J11_MINB, true>(B, num_sms, i, st); break;
            case 11: e = launch_one<11, 11, K1_TILED_J11_MINB, true>(B, num_sms, i, st); break;
            }
        }
        n++;
    }
    if (launches) *launches = n;
    return e;
}

extern "C" unsigned k1_tiled_class_of(int S, int J)
{
    const int *t = class_table(J);
    for (int i = 0; t && i < 6; i++)
        if (t[i] > 0 && S <= 32 * t[i]) return (unsigned)i;
    return 0xffffffffu;
}
