// k6_read_stats.cu — K6: the per-read aggregation loops that follow the arg-max in DetInDel::diploidGLF
// (DInDel.cpp:3126-3132 + :3148-3170 "dip.map" sites, :3396-3404 + :3504-3554 per variant position) and in
// DetInDel::estimateHaplotypeFrequenciesBayesEM (:2697-2742, per variant and pool), over the resident K1 records.
//
// One CTA per job (one site of one window with its candidate haplotypes).  The reads are walked in tiles of
// K6_TILE: every thread evaluates whole reads (record fields of the candidate haplotypes, the coverage test
// firstBase + pad <= startRead && lastBase - pad >= endRead of Variant.hpp:125-128, strand / unmapped bits, mq by
// phred byte) and leaves the read's three FP64 contributions (mq^2 when covered, mq^2, mLogBQ of the chosen record;
// +0.0 where the reference adds nothing) in shared memory; three threads of three different warps then add the tile
// to the running sums IN READ ORDER, which keeps msq / allmsq / mLogBQ bit-identical to the reference's loops (an
// FP64 sum is order dependent; adding +0.0 changes no partial sum).  The integer counters are reduced with warp
// shuffles + shared atomics.  Integer / byte work on 64-byte records already in HBM/L2: no floating-point
// throughput, no tensor cores; the kernel is bound by the record reads (64 B x candidates per read).
#include "k6.cuh"

#define K6_THREADS 128
#define K6_TILE 1024
#define FULLMASK 0xffffffffu

__device__ __forceinline__ bool k6_covered(const dindel_pair_out &p, int pad, int start, int end)
{
    return p.firstBase + pad <= start && p.lastBase - pad >= end; // AlignedVariant::isCovered
}

__global__ void __launch_bounds__(K6_THREADS) k6_read_stats_kernel(K6Args a)
{
    __shared__ double s_msq[K6_TILE], s_all[K6_TILE], s_bq[K6_TILE];
    __shared__ double s_sum[3];
    __shared__ int s_cnt[11];
    const int j = blockIdx.x;
    if (j >= a.n_jobs) return;
    const int tid = threadIdx.x;
    const int w = a.job_window[j], mode = a.job_mode[j], grp = a.job_group[j];
    const long long c0 = a.job_cand[j];
    const int nc = (int)(a.job_cand[j + 1] - c0);
    const int nh = a.win_hap[w + 1] - a.win_hap[w];
    const int r0 = a.win_read[w], nr = a.win_read[w + 1] - r0;
    const dindel_pair_out *P = a.pairs + a.win_pair[w];
    const bool grouped = grp >= 0 && a.read_group != nullptr;
    if (tid < 3) s_sum[tid] = 0.0;
    if (tid < 11) s_cnt[tid] = 0;
    __syncthreads();
    // counters: 0 n, 1 nf, 2 nr, 3 numOffBoth, 4 numIndels, 5 nBQT, 6 nmmBQT, 7 nMMLeft, 8 nMMRight, 9 numUnmappedRealigned, 10 n_reads
    int cnt[11];
#pragma unroll
    for (int k = 0; k < 11; k++) cnt[k] = 0;

    // MAP walks the reads once per candidate (candidate-major FP order, DInDel.cpp:3152-3170); SITE / POOL once
    const int npass = (mode == DINDEL_STATS_MAP) ? (nc < 2 ? nc : 2) : 1;
    for (int pass = 0; pass < npass; pass++) {
        for (int t0 = 0; t0 < nr; t0 += K6_TILE) {
            const int tn = min(K6_TILE, nr - t0);
            for (int i = tid; i < tn; i += K6_THREADS) {
                const int r = t0 + i;
                double vm = 0.0, va = 0.0, vb = 0.0;
                const bool in = !grouped || a.read_group[r0 + r] == grp;
                if (in) {
                    const double mq = a.mq[a.read_mapq[r0 + r]];
                    const unsigned fl = a.read_flags[r0 + r];
                    const bool rev = (fl & DINDEL_READ_REVERSE) != 0, unm = (fl & DINDEL_READ_UNMAPPED) != 0;
                    const dindel_pair_out *row = P + (long long)r * nh;
                    if (mode == DINDEL_STATS_SITE) {
                        const dindel_pair_out &p1 = row[a.cand_hap[c0]], &p2 = row[a.cand_hap[c0 + 1]];
                        cnt[10]++;
                        va = mq * mq;                                                  // :3505-3506
                        if (p1.offHap && p2.offHap) cnt[3]++;                          // :3512
                        if (unm && (!p1.offHap || !p2.offHap)) cnt[9]++;               // :3397-3403
                        const int k = (p1.ll >= p2.ll) ? 0 : 1;                        // :3514-3520
                        const dindel_pair_out &p = k ? p2 : p1;
                        cnt[4] += p.numIndels; cnt[5] += p.nBQT; cnt[6] += p.nmmBQT;   // :3529-3531
                        vb = p.mLogBQ;                                                 // :3532
                        if (p.nMMLeft >= 2) cnt[7]++;                                  // :3533
                        if (p.nMMRight >= 2) cnt[8]++;                                 // :3534
                        if ((a.cand_flags[c0 + k] & DINDEL_CAND_PRESENT) &&
                            k6_covered(p, a.pad, a.cand_start[c0 + k], a.cand_end[c0 + k])) { // :3537-3550
                            if (rev) cnt[2]++; else cnt[1]++;                          // one haplotype per read: nft / nrt
                            vm = mq * mq;
                            cnt[0]++;
                        }
                    } else if (mode == DINDEL_STATS_MAP) {
                        if (pass == 0) {
                            cnt[10]++;
                            if (nc >= 2) {
                                const dindel_pair_out &p1 = row[a.cand_hap[c0]], &p2 = row[a.cand_hap[c0 + 1]];
                                if (unm && (!p1.offHap || !p2.offHap)) cnt[9]++;       // :3126-3132
                            }
                        }
                        if (a.cand_flags[c0 + pass] & DINDEL_CAND_PRESENT) {           // :3155-3156
                            const dindel_pair_out &p = row[a.cand_hap[c0 + pass]];
                            if (k6_covered(p, a.pad, a.cand_start[c0 + pass], a.cand_end[c0 + pass])) { // :3161-3168
                                if (rev) cnt[2]++; else cnt[1]++;
                                vm = mq * mq;
                                cnt[0]++;
                            }
                        }
                    } else { // DINDEL_STATS_POOL
                        cnt[10]++;
                        double ml = __longlong_as_double(0xfff0000000000000LL);
                        for (int c = 0; c < nc; c++) {                                 // :2702-2707
                            const double l = row[a.cand_hap[c0 + c]].ll;
                            if (l >= ml) ml = l;
                        }
                        const double thr = ml - 1e-7;                                  // :2709
                        bool hit = false;
                        for (int c = 0; c < nc; c++) {
                            const dindel_pair_out &p = row[a.cand_hap[c0 + c]];
                            const unsigned cf = a.cand_flags[c0 + c];
                            if (p.ll >= thr && (cf & DINDEL_CAND_PRESENT) && (cf & DINDEL_CAND_HAS_VAR) &&
                                k6_covered(p, a.pad, a.cand_start[c0 + c], a.cand_end[c0 + c])) hit = true; // :2716-2733
                        }
                        if (hit) { if (rev) cnt[2]++; else cnt[1]++; }                 // :2739-2740
                        vm = mq * mq;                                                  // :2736-2737
                        cnt[0]++;
                    }
                }
                s_msq[i] = vm; s_all[i] = va; s_bq[i] = vb;
            }
            __syncthreads();
            if ((tid & 31) == 0 && tid < 96) { // three warps, one sum each, in read order
                const double *src = tid == 0 ? s_msq : (tid == 32 ? s_all : s_bq);
                double acc = s_sum[tid >> 5];
                for (int i = 0; i < tn; i++) acc += src[i];
                s_sum[tid >> 5] = acc;
            }
            __syncthreads();
        }
    }
#pragma unroll
    for (int k = 0; k < 11; k++) {
        const int v = __reduce_add_sync(FULLMASK, cnt[k]);
        if ((tid & 31) == 0 && v) atomicAdd(&s_cnt[k], v);
    }
    __syncthreads();
    if (tid == 0) {
        dindel_read_stats o;
        o.msq = s_sum[0]; o.allmsq = s_sum[1]; o.mLogBQ = s_sum[2];
        o.n = s_cnt[0]; o.nf = s_cnt[1]; o.nr = s_cnt[2]; o.numOffBoth = s_cnt[3]; o.numIndels = s_cnt[4];
        o.nBQT = s_cnt[5]; o.nmmBQT = s_cnt[6]; o.nMMLeft = s_cnt[7]; o.nMMRight = s_cnt[8];
        o.numUnmappedRealigned = s_cnt[9]; o.n_reads = s_cnt[10]; o.reserved = 0;
        a.out[j] = o;
    }
}

extern "C" cudaError_t k6_launch(const K6Args *a, cudaStream_t st)
{
    if (a->n_jobs <= 0) return cudaSuccess;
    k6_read_stats_kernel<<<a->n_jobs, K6_THREADS, 0, st>>>(*a);
    return cudaGetLastError();
}
