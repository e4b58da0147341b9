// k1_generic.cu — K1, shape-generic variant: one warp per (read, haplotype) pair, rows of the
// two-sided Viterbi kept in shared memory, lanes striding over state columns.  Handles any
// haplotype length / read length / maxLengthDel the batch brings (the register-tiled kernel in
// k1_tiled.cu covers the common shapes faster).  Arithmetic follows the reference statement by
// statement so that every log-likelihood and back-pointer is bit-identical:
//   passMessageTwoDec  ObservationModelFB.cpp:1775-1829     (left half, rows 1..bMid)
//   passMessageTwoInc  ObservationModelFB.cpp:1715-1773     (right half, rows R-1..bMid+1)
//   updateMax          ObservationModelFB.cpp:877-888
//   calcLikelihoodFromLastSlice  cpp:1075-1144,  computeBMidPrior cpp:268-305 (via host LUT)
// Device arithmetic is FP64 add / compare only; all log() values come from K1Luts (host glibc).
#include "k1_common.cuh"
#include "k1_tail.cuh"

#define UPD(dv, di, nvexpr, niexpr)                                              \
    do {                                                                         \
        double _nv = (nvexpr);                                                   \
        int _ni = (niexpr);                                                      \
        if (_nv > dv + 1e-10) { dv = _nv; di = _ni; }                            \
        else if (_nv >= dv && _nv <= dv + 1e-5 && di > _ni) { dv = _nv; di = _ni; } \
    } while (0)

struct GenSmem {
    double *P[2];   // [2*S] each: plane 0 then plane 1
    double *O0;     // [S]  plane-0 emission of the source row
    double *lpe, *lpn; // [S]
    double *ao;     // [2*S] alpha[bMid]+obs[bMid]
    double *eqs, *uqs; // [R]
    uint8_t *hc;    // [S] hap char per column ('N' for LO/RO)
    uint8_t *ch, *ph; // [R]
    int16_t *ms;    // [R]
    uint8_t *bt;    // [R*S] (shared or global)
};

__host__ __device__ inline size_t gen_smem_doubles(int maxS, int maxR) { return (size_t)9 * maxS + 2 * maxR; }
__host__ __device__ inline size_t gen_smem_bytes_nobt(int maxS, int maxR)
{
    size_t b = gen_smem_doubles(maxS, maxR) * 8 + (size_t)maxS + 2 * (size_t)maxR + 2 * (size_t)maxR;
    return (b + 15) & ~(size_t)15;
}

__host__ __device__ inline size_t gen_smem_per_warp(int maxS, int maxR, int bt_in_smem)
{
    size_t b = gen_smem_bytes_nobt(maxS, maxR);
    if (bt_in_smem) b += (((size_t)maxS * maxR) + 15) & ~(size_t)15;
    return b;
}
extern "C" size_t k1_generic_smem_per_warp(int maxS, int maxR, int bt_in_smem)
{
    return gen_smem_per_warp(maxS, maxR, bt_in_smem);
}

__global__ void __launch_bounds__(128) k1_generic_kernel(K1Batch B, int bt_in_smem)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31;
    const int wib = threadIdx.x >> 5;
    const int warps_per_cta = blockDim.x >> 5;
    const K1Luts *__restrict__ L = B.luts;
    const int maxS = B.max_S, maxR = B.max_R;
    const size_t per_warp = gen_smem_per_warp(maxS, maxR, bt_in_smem);
    unsigned char *base = smem_raw + per_warp * wib;
    GenSmem sm;
    {
        double *d = reinterpret_cast<double *>(base);
        sm.P[0] = d; d += 2 * maxS;
        sm.P[1] = d; d += 2 * maxS;
        sm.O0 = d; d += maxS;
        sm.lpe = d; d += maxS;
        sm.lpn = d; d += maxS;
        sm.ao = d; d += 2 * maxS;
        sm.eqs = d; d += maxR;
        sm.uqs = d; d += maxR;
        uint8_t *u = reinterpret_cast<uint8_t *>(d);
        sm.ms = reinterpret_cast<int16_t *>(u); u += 2 * maxR;
        sm.hc = u; u += maxS;
        sm.ch = u; u += maxR;
        sm.ph = u; u += maxR;
        if (bt_in_smem) {
            sm.bt = base + gen_smem_bytes_nobt(maxS, maxR);
        } else {
            long long gw = (long long)blockIdx.x * warps_per_cta + wib;
            sm.bt = B.bt_scratch + gw * B.bt_slot_bytes;
        }
    }
    const double lLOgLO = L->lLOgLO, lFirstgLO = L->lFirstgLO, lInsgIns = L->lInsgIns,
                 lNoInsgIns = L->lNoInsgIns, lNoInsgNoIns = L->lNoInsgNoIns;
    const int T = L->numT, J = T - 1;
    const bool force = L->force != 0;

    for (;;) {
        long long unit = 0;
        if (lane == 0) unit = B.unit_begin + (long long)atomicAdd(B.counter, 1ULL);
        unit = __shfl_sync(0xffffffffu, unit, 0);
        if (unit >= B.unit_end) break;
        // search only this launch's window range: in a pipelined submit the CSR arrays of later chunks are
        // not uploaded yet when this chunk's kernel runs
        const int w = B.w_begin + k1_find_window(B.win_unit + B.w_begin, B.w_end - B.w_begin, unit);
        const int h0 = B.win_hap[w], nh = B.win_hap[w + 1] - h0;
        const int r0 = B.win_read[w], nr = B.win_read[w + 1] - r0;
        const int nchunk = (nr + K1_READ_CHUNK - 1) / K1_READ_CHUNK;
        const int local = (int)(unit - B.win_unit[w]);
        const int h = local / nchunk, chunk = local - h * nchunk;
        const long long hoff = B.hap_off[h0 + h];
        const int H = (int)(B.hap_off[h0 + h + 1] - hoff);
        const int S = H + 2, RO = H + 1;
        const uint8_t *hap = B.hap_seq + hoff;
        const uint8_t *hp = B.hap_hp + hoff + 2LL * (h0 + h);
        const uint32_t hapStart = B.hap_start[w];
        const bool hapsize_err = L->maxLengthDel > H;

        __syncwarp();
        for (int x = lane; x < S; x += 32) {
            int code = hp[x];
            sm.lpe[x] = L->hp_lpe[code];
            sm.lpn[x] = L->hp_lpn[code];
            sm.hc[x] = (x >= 1 && x <= H) ? hap[x - 1] : (uint8_t)'N';
        }
        __syncwarp();

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
            if (R <= 0) { // degenerate read: the reference would index read[-1]; report NaN skip
                if (lane == 0) {
                    k1_write_hapsize_error(out, 0);
                    out->status = DINDEL_WIN_NAN;
                }
                continue;
            }
            const int mapq = B.read_mapq[r];
            const bool unmapped = (B.read_flags[r] & DINDEL_READ_UNMAPPED) != 0;
            const int bMid = k1_bmid(hapStart, H, R, B.read_start[r], unmapped, L->bMidOverride);
            for (int b = lane; b < R; b += 32) {
                int ph = B.read_qual[roff + b];
                sm.ch[b] = B.read_seq[roff + b];
                sm.ph[b] = (uint8_t)ph;
                sm.eqs[b] = L->eq[ph];
                sm.uqs[b] = L->uq[ph];
            }
            __syncwarp();

            // ---------------- left half: alpha rows 1..bMid (cpp:1573-1575) ----------------
            int cur = 0;
            for (int s = lane; s < 2 * S; s += 32) sm.P[0][s] = 0.0; // cpp:335-338
            __syncwarp();
            for (int b = 1; b <= bMid; b++) {
                const double *P0 = sm.P[cur], *P1 = sm.P[cur] + S;
                double *A0 = sm.P[cur ^ 1], *A1 = sm.P[cur ^ 1] + S;
                const double eq = sm.eqs[b - 1], uq = sm.uqs[b - 1];
                const double eqLR = force ? -1000.0 : eq; // cpp:307-316
                const uint8_t rc = sm.ch[b - 1];
                for (int x = lane; x < S; x += 32) {
                    uint8_t c = sm.hc[x];
                    sm.O0[x] = (x == 0 || x == RO) ? eqLR : ((c == 'N' || c == rc) ? eq : uq);
                }
                __syncwarp();
                const double *O0 = sm.O0;
                uint8_t *btrow = sm.bt + (size_t)b * S;
                for (int x = lane; x < S; x += 32) {
                    const double O1x = (x == 0 || x == RO) ? eqLR : eq;
                    double a = K1_NEG_INF;
                    int idx = 0x7fffffff, code0;
                    if (x == RO) {
                        UPD(a, idx, O0[RO] + P0[RO] + lLOgLO + lNoInsgNoIns, RO);
                        UPD(a, idx, O0[H] + P0[H] + lFirstgLO + lNoInsgNoIns, H);
                        UPD(a, idx, eqLR + P1[RO] + lLOgLO + sm.lpe[RO], S + RO);
                        UPD(a, idx, eq + P1[H] + lFirstgLO + sm.lpe[H], S + H);
                        code0 = (idx == RO) ? 0 : (idx == H) ? 1 : (idx == S + RO) ? 2 : 3;
                    } else if (x == 0) {
                        a = O0[0] + P0[0] + lNoInsgNoIns;
                        code0 = 0;
                    } else {
                        const double lpt = sm.lpe[x], lpn = sm.lpn[x];
                        for (int y = 1; y < T; y++) {
                            int nx = x - y;
                            if (nx < 0) nx = 0;
                            double lp = (y == 1) ? lpn : (lpt + (double)(y - 1) * lInsgIns);
                            UPD(a, idx, O0[nx] + lp + P0[nx] + lpn, nx);
                        }
                        const double O1s = (x - 1 == 0) ? eqLR : eq;
                        UPD(a, idx, O1s + P1[x - 1] + lpt, S + x - 1);
                        code0 = (idx >= S) ? J : (x - idx - 1);
                    }
                    double a1 = O1x + P1[x] + lInsgIns;
                    int idx1 = S + x;
                    if (x >= 1) UPD(a1, idx1, O0[x] + P0[x] + lNoInsgIns, x);
                    A0[x] = a;
                    A1[x] = a1;
                    btrow[x] = (uint8_t)(code0 | ((idx1 < S) ? 128 : 0));
                }
                __syncwarp();
                cur ^= 1;
            }
            { // stash alpha[bMid] + obs[bMid]  (first add of cpp:1098)
                const double eq = sm.eqs[bMid], uq = sm.uqs[bMid];
                const double eqLR = force ? -1000.0 : eq;
                const uint8_t rc = sm.ch[bMid];
                const double *P0 = sm.P[cur], *P1 = sm.P[cur] + S;
                for (int x = lane; x < S; x += 32) {
                    uint8_t c = sm.hc[x];
                    bool edge = (x == 0 || x == RO);
                    double o0 = edge ? eqLR : ((c == 'N' || c == rc) ? eq : uq);
                    double o1 = edge ? eqLR : eq;
                    sm.ao[x] = P0[x] + o0;
                    sm.ao[S + x] = P1[x] + o1;
                }
            }
            __syncwarp();
            // ---------------- right half: beta rows R-2..bMid (cpp:1576-1578) ----------------
            cur = 0;
            for (int s = lane; s < 2 * S; s += 32) sm.P[0][s] = 0.0;
            __syncwarp();
            for (int b = R - 1; b > bMid; b--) {
                const double *P0 = sm.P[cur], *P1 = sm.P[cur] + S;
                double *B0 = sm.P[cur ^ 1], *B1 = sm.P[cur ^ 1] + S;
                const double eq = sm.eqs[b], uq = sm.uqs[b];
                const double eqLR = force ? -1000.0 : eq;
                const uint8_t rc = sm.ch[b];
                for (int x = lane; x < S; x += 32) {
                    uint8_t c = sm.hc[x];
                    sm.O0[x] = (x == 0 || x == RO) ? eqLR : ((c == 'N' || c == rc) ? eq : uq);
                }
                __syncwarp();
                const double *O0 = sm.O0;
                uint8_t *btrow = sm.bt + (size_t)b * S; // btb[b-1] stored at row b
                for (int x = lane; x < S; x += 32) {
                    const double O1x = (x == 0 || x == RO) ? eqLR : eq;
                    double a = K1_NEG_INF;
                    int idx = 0x7fffffff, code0;
                    if (x == 0) {
                        UPD(a, idx, O0[0] + P0[0] + lLOgLO + lNoInsgNoIns, 0);
                        UPD(a, idx, O0[1] + P0[1] + lFirstgLO + lNoInsgNoIns, 1);
                        UPD(a, idx, O1x + P1[0] + sm.lpe[1], S);
                        code0 = (idx == 0) ? 0 : (idx == 1) ? 1 : J;
                    } else if (x == RO) {
                        UPD(a, idx, O0[RO] + P0[RO] + sm.lpn[RO], RO);
                        UPD(a, idx, O1x + P1[RO], S + RO);
                        code0 = (idx == RO) ? 0 : J;
                    } else {
                        for (int y = 1; y < T; y++) {
                            int nx = x + y;
                            if (nx > H) nx = RO;
                            double lpn = sm.lpn[nx], lpt = sm.lpe[nx];
                            double lp = (y == 1) ? lpn : (lpt + (double)(y - 1) * lInsgIns);
                            UPD(a, idx, lp + lpn + P0[nx] + O0[nx], nx);
                        }
                        UPD(a, idx, O1x + P1[x] + sm.lpe[x + 1], S + x);
                        code0 = (idx >= S) ? J : (idx - x - 1);
                    }
                    double a1 = O1x + P1[x] + lInsgIns;
                    int idx1 = S + x;
                    {
                        int nx = (x == 0) ? 0 : ((x + 1 > RO) ? RO : x + 1);
                        UPD(a1, idx1, O0[nx] + P0[nx] + lNoInsgIns, nx);
                    }
                    B0[x] = a;
                    B1[x] = a1;
                    btrow[x] = (uint8_t)(code0 | ((idx1 < S) ? 128 : 0));
                }
                __syncwarp();
                cur ^= 1;
            }
            // ---------------- meeting row, cpp:1075-1144 ----------------
            double ll = K1_NEG_INF, llH = K1_NEG_INF, llOff = K1_NEG_INF, llOn = K1_NEG_INF;
            int msR = 0, msH = 0;
            if (lane == 0) {
                const double *Bm = sm.P[cur];
                // insert-size prior of this read (cpp:279-293; only with --libFile)
                const int insMode = B.read_ins_mode ? (int)B.read_ins_mode[r] : DINDEL_INS_NONE;
                int insAnchor = 0, insMax = 1;
                const double *insLogp = nullptr;
                double ins0 = 0.0;
                if (insMode != DINDEL_INS_NONE) {
                    const int lib = B.read_lib[r];
                    insAnchor = B.read_ins_anchor[r];
                    insLogp = B.lib_logp + B.lib_off[lib];
                    insMax = (int)(B.lib_off[lib + 1] - B.lib_off[lib]);
                    ins0 = B.lib_log95[lib];
                }
                for (int s = 0; s < 2 * S; s++) {
                    int pl = s >= S;
                    int x = pl ? s - S : s;
                    double t = sm.ao[s] + Bm[s];
                    double pr = (x == 0) ? L->priorLO[pl][mapq] : (x == RO) ? -100.0 : L->priorOn[pl][mapq];
                    double prH = (x == 0) ? L->priorLO_H[pl] : (x == RO) ? -100.0 : L->priorOn_H[pl];
                    if (insMode != DINDEL_INS_NONE && x != RO) {
                        if (x == 0) { pr = pr + ins0; prH = prH + ins0; } // cpp:298
                        else {                                            // cpp:301
                            const double pin = k1_pinsert(insLogp, insMax, insMode, insAnchor, (int)hapStart, x, R, bMid);
                            pr = (pin + L->prior1m[mapq]) + L->priorIns[pl];
                            prH = (pin + L->prior1m_H) + L->priorIns[pl];
                        }
                    }
                    double v = t + pr;
                    if (v > ll + 1e-10) { ll = v; msR = s; }
                    if (x == 0) { if (v > llOff) llOff = v; }
                    else if (x != RO) { if (v > llOn) llOn = v; }
                    v = t + prH;
                    if (v > llH + 1e-10) { llH = v; msH = s; }
                }
            }
            ll = __shfl_sync(0xffffffffu, ll, 0);
            llOn = __shfl_sync(0xffffffffu, llOn, 0);
            llOff = __shfl_sync(0xffffffffu, llOff, 0);
            msR = __shfl_sync(0xffffffffu, msR, 0);
            msH = __shfl_sync(0xffffffffu, msH, 0);
            const int xr = msR % S, xh = msH % S;
            const int offHap = (xr == 0 || xr == RO), offHapHMQ = (xh == 0 || xh == RO);
            const uint8_t *btp = sm.bt;
            auto btf = [btp, S](int row, int x) -> int { return btp[(size_t)row * S + x]; };
            k1_tail(btf, lane, R, H, J, bMid, msH, hap, sm.ch, sm.ph, sm.ms, sm.eqs, L, ll, llOn,
                    llOff, offHap, offHapHMQ, out,
                    B.hpos ? B.hpos + pair * (long long)B.hpos_stride : nullptr);
        }
    }
}

extern "C" cudaError_t k1_generic_launch(const K1Batch *B, int grid, int warps_per_cta,
                                         int bt_in_smem, size_t smem_bytes, cudaStream_t stream)
{
    cudaError_t e = cudaFuncSetAttribute(k1_generic_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)smem_bytes);
    if (e != cudaSuccess) return e;
    k1_generic_kernel<<<grid, warps_per_cta * 32, smem_bytes, stream>>>(*B, bt_in_smem);
    return cudaGetLastError();
}
