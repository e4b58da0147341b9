// k23_inference.cu — the per-window inference kernels that consume the read x haplotype
// log-likelihood matrix IN PLACE (rl[r*nh+h] = pairs[win_pair[w] + r*nh + h].ll, the resident K1
// results; DInDel.cpp:2949-2960 / :2125-2139).
//
//   K2  pair sums   acc <- start; for r in window (in order): acc += log(.5) + addLogs(rl[r,h1], rl[r,h2])
//                   for every h1 <= h2: the inner loop of DetInDel::diploidGLF (DInDel.cpp:3085-3113,
//                   :3349-3390) and of the pooled genotype likelihoods (DInDel.cpp:2683-2694, reads
//                   restricted to one pool member).  One CTA per job.  The window's ll column block is staged
//                   once per read tile from the 64-byte pair records into shared memory (one coalesced sweep,
//                   no thread re-reads a record); one thread per haplotype pair then sums the tile's reads in
//                   the CPU's order, so the only arithmetic difference to the reference is exp/log (CUDA libm
//                   vs glibc, ulp level).
//   K3  Bayes EM    the variational-Bayes EM over haplotype frequencies for one set of active variants
//                   (DetInDel::estimateHaplotypeFrequenciesBayesEM, DInDel.cpp:2431-2545).  One CTA per set;
//                   the ll matrix lives in shared memory (odd pitch: conflict-free column walks) for all <= 27
//                   passes, one thread per read with its responsibilities in registers, warp-shuffle + fixed
//                   cross-warp reductions (deterministic), digammas one per lane.  Responsibilities use the
//                   max-shifted log-sum-exp (nh exp + 1 log per read instead of the reference's chain of nh
//                   addLogs + nh exp): mathematically the same numbers, a few ulp apart.
// Neither kernel can be bit-identical to glibc; both report what the host needs to keep the TEXT output identical:
// K2's error bound is a closed form of (reads, magnitude) (dindel_b200_sum_error_bound); K3 returns per set the
// margin min_pass ||dE| - tol| of the stopping rule (DInDel.cpp:2515) so that a knife-edge set is re-run on the host.
#include "k1_common.cuh"
#include "k23.cuh"

// Utils.hpp:29-38 — note log(1+exp()), not log1p
__device__ __forceinline__ double addlogs(double l1, double l2)
{
    if (l1 > l2) {
        double diff = l2 - l1;
        return l1 + log(1.0 + exp(diff));
    } else {
        double diff = l1 - l2;
        return l2 + log(1.0 + exp(diff));
    }
}

// Mirror of oracle ora_digamma (Boost's digamma is third-party and unpinned: "parity unpinned").
__device__ __forceinline__ double digamma_pos(double x)
{
    double r = 0.0;
    while (x < 10.0) {
        r -= 1.0 / x;
        x += 1.0;
    }
    double f = 1.0 / (x * x);
    double t = f * (-1.0 / 12.0 + f * (1.0 / 120.0 + f * (-1.0 / 252.0 + f * (1.0 / 240.0 + f * (-1.0 / 132.0
               + f * (691.0 / 32760.0 + f * (-1.0 / 12.0)))))));
    return r + log(x) - 0.5 / x + t;
}

// ------------------------------------------------------------------------------------------------------------
// K2
// ------------------------------------------------------------------------------------------------------------
#define K2_TILE_BYTES 32768

__global__ void k2_pair_sums(K2Args A)
{
    extern __shared__ double s_rl[]; // [tile_r][nh] then uint8 keep[tile_r]
    const int j = blockIdx.x;
    if (j >= A.n_jobs) return;
    const int w = A.job_window[j], grp = A.job_group[j];
    const int nh = A.win_hap[w + 1] - A.win_hap[w];
    const int r0 = A.win_read[w], nr = A.win_read[w + 1] - r0;
    const int npairs = nh * (nh + 1) / 2;
    if (nh <= 0) return;
    const dindel_pair_out *p = A.pairs + A.win_pair[w];
    const long long off = A.job_off[j];
    int tile_r = K2_TILE_BYTES / (nh * 9);
    if (tile_r > nr) tile_r = nr;
    if (tile_r < 1) tile_r = 1;
    uint8_t *s_keep = reinterpret_cast<uint8_t *>(s_rl + (size_t)tile_r * nh);
    const bool grouped = grp >= 0 && A.read_group != nullptr;

    // this thread's haplotype pairs: q -> (h1 <= h2), row-major upper triangle; up to 4 per thread in registers
    int h1v[4], h2v[4];
    double acc[4];
    int nq = 0;
    for (int q = threadIdx.x; q < npairs && nq < 4; q += blockDim.x, nq++) {
        int h1 = 0, rem = q;
        while (rem >= nh - h1) { rem -= nh - h1; h1++; }
        h1v[nq] = h1; h2v[nq] = h1 + rem;
        acc[nq] = A.start ? A.start[off + q] : 0.0;
    }
    for (int t0 = 0; t0 < nr; t0 += tile_r) {
        const int tn = min(tile_r, nr - t0);
        __syncthreads();
        for (int i = threadIdx.x; i < tn * nh; i += blockDim.x) s_rl[i] = p[(long long)t0 * nh + i].ll;
        for (int i = threadIdx.x; i < tn; i += blockDim.x) s_keep[i] = grouped ? (A.read_group[r0 + t0 + i] == grp) : 1;
        __syncthreads();
#pragma unroll
        for (int k = 0; k < 4; k++) {
            if (k < nq) {
                const int h1 = h1v[k], h2 = h2v[k];
                double a = acc[k];
                for (int r = 0; r < tn; r++) {
                    if (!s_keep[r]) continue;
                    a += A.log_half + addlogs(s_rl[r * nh + h1], s_rl[r * nh + h2]);
                }
                acc[k] = a;
            }
        }
    }
#pragma unroll
    for (int k = 0; k < 4; k++)
        if (k < nq) A.out[off + threadIdx.x + k * blockDim.x] = acc[k];
    // more than 4 pairs per thread (nh > ~44 at 256 threads): the remaining pairs straight from the records
    for (int q = threadIdx.x + 4 * blockDim.x; q < npairs; q += blockDim.x) {
        int h1 = 0, rem = q;
        while (rem >= nh - h1) { rem -= nh - h1; h1++; }
        const int h2 = h1 + rem;
        double a = A.start ? A.start[off + q] : 0.0;
        for (int r = 0; r < nr; r++) {
            if (grouped && A.read_group[r0 + r] != grp) continue;
            a += A.log_half + addlogs(p[(long long)r * nh + h1].ll, p[(long long)r * nh + h2].ll);
        }
        A.out[off + q] = a;
    }
}

// ------------------------------------------------------------------------------------------------------------
// K3
// ------------------------------------------------------------------------------------------------------------
#define K3_MAXH 64
#define K3_MAX_WARPS 16

// NHP: haplotype slots unrolled per thread (nh <= NHP).  STAGE: the ll matrix fits shared memory.
template <int NHP, bool STAGE>
__global__ void __launch_bounds__(32 * K3_MAX_WARPS) k3_bayes_em(K3Args A)
{
    extern __shared__ double s_mem[];
    __shared__ double s_lpi[K3_MAXH], s_pi[K3_MAXH], s_nk[K3_MAXH];
    __shared__ double s_wpart[K3_MAX_WARPS][K3_MAXH + 2];
    __shared__ double s_scal[4]; // loglik, B, eNew
    const int s = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, nwarps = blockDim.x >> 5;
    if (s >= A.n_sets) return;
    const int w = A.set_window[s];
    const int nh = A.win_hap[w + 1] - A.win_hap[w];
    const int nr = A.win_read[w + 1] - A.win_read[w];
    const dindel_pair_out *p = A.pairs + A.win_pair[w];
    const uint8_t *compat = A.compatible + A.set_off[s];
    const int pitch = nh | 1; // odd number of doubles: the 32 lanes of a warp walk 32 rows without bank conflicts
    if (STAGE) {
        for (int i = tid; i < nr * nh; i += blockDim.x) {
            const int r = i / nh, h = i - r * nh;
            s_mem[r * pitch + h] = p[i].ll;
        }
    }
    int numah = 0;
    for (int h = 0; h < nh; h++) numah += compat[h] ? 1 : 0;
    if (tid < nh) s_lpi[tid] = compat[tid] ? log(1.0 / (double)numah) : -100.0; // DInDel.cpp:2418
    __syncthreads();
    double eOld = K1_NEG_INF, margin = __longlong_as_double(0x7ff0000000000000LL), loglik = 0.0;
    int iter = 0;
    for (;;) {
        // ---- E-step over this thread's reads (DInDel.cpp:2439-2457) ----
        double nkp[NHP];
#pragma unroll
        for (int h = 0; h < NHP; h++) nkp[h] = 0.0;
        double ll = 0.0, Bp = 0.0;
        for (int r = tid; r < nr; r += blockDim.x) {
            double e[NHP];
            double m = K1_NEG_INF;
#pragma unroll
            for (int h = 0; h < NHP; h++) {
                if (h < nh) {
                    e[h] = s_lpi[h] + (STAGE ? s_mem[r * pitch + h] : p[(long long)r * nh + h].ll);
                    m = fmax(m, e[h]);
                }
            }
            double sum = 0.0;
#pragma unroll
            for (int h = 0; h < NHP; h++) {
                if (h < nh) {
                    e[h] = exp(e[h] - m);
                    sum += e[h];
                }
            }
            const double inv = 1.0 / sum;
#pragma unroll
            for (int h = 0; h < NHP; h++) {
                if (h < nh) {
                    const double zz = e[h] * inv;
                    nkp[h] += zz;
                    Bp += zz * (STAGE ? s_mem[r * pitch + h] : p[(long long)r * nh + h].ll);
                }
            }
            ll += m + log(sum);
        }
        // ---- deterministic reductions: xor tree inside the warp, then the warps in order ----
#pragma unroll
        for (int h = 0; h < NHP; h++) {
            if (h < nh) {
                double v = nkp[h];
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                if (lane == 0) s_wpart[wid][h] = v;
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            ll += __shfl_xor_sync(0xffffffffu, ll, o);
            Bp += __shfl_xor_sync(0xffffffffu, Bp, o);
        }
        if (lane == 0) { s_wpart[wid][K3_MAXH] = ll; s_wpart[wid][K3_MAXH + 1] = Bp; }
        __syncthreads();
        if (tid < nh) {
            double v = 0.0;
            for (int k = 0; k < nwarps; k++) v += s_wpart[k][tid];
            s_nk[tid] = v;
        }
        if (tid == 0) {
            double v0 = 0.0, v1 = 0.0;
            for (int k = 0; k < nwarps; k++) { v0 += s_wpart[k][K3_MAXH]; v1 += s_wpart[k][K3_MAXH + 1]; }
            s_scal[0] = v0; s_scal[1] = v1;
        }
        __syncthreads();
        // ---- M-step (DInDel.cpp:2461-2480), one haplotype per thread ----
        if (tid < nh) {
            double ahat = 0.0;
            for (int h = 0; h < nh; h++)
                if (compat[h]) ahat += s_nk[h] + A.a0;
            if (compat[tid]) {
                s_lpi[tid] = digamma_pos(s_nk[tid] + A.a0) - digamma_pos(ahat);
                s_pi[tid] = log((A.a0 + s_nk[tid]) / ((double)numah * A.a0 + (double)nr));
            } else {
                s_lpi[tid] = -100.0;
                s_pi[tid] = -100.0;
            }
        }
        __syncthreads();
        // ---- energy (DInDel.cpp:2484-2492: sum z*(pi+rl) = sum z*rl + sum_h pi[h]*nk[h]) and stopping rule (:2515);
        //      every thread evaluates the same expression on the same shared values ----
        double eNew = s_scal[1];
        for (int h = 0; h < nh; h++) eNew += s_pi[h] * s_nk[h];
        loglik = s_scal[0];
        const double dE = fabs(eOld - eNew);
        if (iter <= 25) margin = fmin(margin, fabs(dE - A.tol));
        const bool done = (dE < A.tol) || (iter > 25);
        eOld = eNew;
        iter++;
        if (done) break;
        __syncthreads(); // s_scal / s_wpart are rewritten by the next pass
    }
    if (tid == 0) {
        double zs = 0.0;
        for (int h = 0; h < nh; h++) zs += exp(s_pi[h]); // :2528-2531
        const long long off = A.set_off[s];
        for (int h = 0; h < nh; h++) {
            A.pi[off + h] = s_pi[h];
            A.freqs[off + h] = exp(s_pi[h]) / zs; // :2543
        }
        A.loglik[s] = loglik;
        A.iters[s] = iter;
        if (A.margin) A.margin[s] = margin;
        if (A.energy) A.energy[s] = eOld;
    }
}

// K3 for windows with more than K3_MAXH haplotypes (deep pooled data: 70-90 candidate haplotypes per window, thousands
// of reads).  A read's responsibilities no longer fit one thread's registers, so the roles turn: a WARP owns a read, its
// lanes hold the haplotypes h = lane, lane + 32, ... (K3W_KH slots); max / sum of the log-sum-exp by warp shuffles, the
// counts nk[h] accumulate per lane over the warp's reads, then the warps are summed in a fixed order (deterministic).
// The ll values come straight from the pair records (L2) every pass: 2 500 x 92 doubles do not fit shared memory.
// Same arithmetic per read as k3_bayes_em, another reduction order: inside the same error bound.
#define K3W_MAXH 256
#define K3W_KH (K3W_MAXH / 32)
__global__ void __launch_bounds__(32 * K3_MAX_WARPS) k3_bayes_em_wide(K3Args A)
{
    __shared__ double s_lpi[K3W_MAXH], s_pi[K3W_MAXH], s_nk[K3W_MAXH];
    __shared__ double s_wpart[K3_MAX_WARPS][K3W_MAXH + 2];
    __shared__ double s_scal[4];
    const int s = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, nwarps = blockDim.x >> 5;
    if (s >= A.n_sets) return;
    const int w = A.set_window[s];
    const int nh = A.win_hap[w + 1] - A.win_hap[w];
    const int nr = A.win_read[w + 1] - A.win_read[w];
    const dindel_pair_out *p = A.pairs + A.win_pair[w];
    const uint8_t *compat = A.compatible + A.set_off[s];
    int numah = 0;
    for (int h = 0; h < nh; h++) numah += compat[h] ? 1 : 0;
    for (int h = tid; h < nh; h += blockDim.x) s_lpi[h] = compat[h] ? log(1.0 / (double)numah) : -100.0; // DInDel.cpp:2418
    __syncthreads();
    double eOld = K1_NEG_INF, margin = __longlong_as_double(0x7ff0000000000000LL), loglik = 0.0;
    int iter = 0;
    for (;;) {
        // ---- E-step: this warp's reads (DInDel.cpp:2439-2457) ----
        double nkp[K3W_KH];
#pragma unroll
        for (int k = 0; k < K3W_KH; k++) nkp[k] = 0.0;
        double ll = 0.0, Bp = 0.0;
        for (int r = wid; r < nr; r += nwarps) {
            double e[K3W_KH], rl[K3W_KH];
            double m = K1_NEG_INF;
#pragma unroll
            for (int k = 0; k < K3W_KH; k++) {
                const int h = lane + 32 * k;
                if (h < nh) {
                    rl[k] = p[(long long)r * nh + h].ll;
                    e[k] = s_lpi[h] + rl[k];
                    m = fmax(m, e[k]);
                }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
            double sum = 0.0;
#pragma unroll
            for (int k = 0; k < K3W_KH; k++) {
                if (lane + 32 * k < nh) {
                    e[k] = exp(e[k] - m);
                    sum += e[k];
                }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
            const double inv = 1.0 / sum;
#pragma unroll
            for (int k = 0; k < K3W_KH; k++) {
                if (lane + 32 * k < nh) {
                    const double zz = e[k] * inv;
                    nkp[k] += zz;
                    Bp += zz * rl[k];
                }
            }
            if (lane == 0) ll += m + log(sum);
        }
        // ---- reductions: Bp over the lanes, then everything over the warps in order ----
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) Bp += __shfl_xor_sync(0xffffffffu, Bp, o);
#pragma unroll
        for (int k = 0; k < K3W_KH; k++)
            if (lane + 32 * k < nh) s_wpart[wid][lane + 32 * k] = nkp[k];
        if (lane == 0) { s_wpart[wid][K3W_MAXH] = ll; s_wpart[wid][K3W_MAXH + 1] = Bp; }
        __syncthreads();
        for (int h = tid; h < nh; h += blockDim.x) {
            double v = 0.0;
            for (int k = 0; k < nwarps; k++) v += s_wpart[k][h];
            s_nk[h] = v;
        }
        if (tid == 0) {
            double v0 = 0.0, v1 = 0.0;
            for (int k = 0; k < nwarps; k++) { v0 += s_wpart[k][K3W_MAXH]; v1 += s_wpart[k][K3W_MAXH + 1]; }
            s_scal[0] = v0; s_scal[1] = v1;
        }
        __syncthreads();
        // ---- M-step (DInDel.cpp:2461-2480) ----
        double ahat = 0.0;
        for (int h = 0; h < nh; h++)
            if (compat[h]) ahat += s_nk[h] + A.a0;
        __syncthreads(); // every thread has read s_nk before s_lpi / s_pi change (they are not read above, kept for clarity)
        for (int h = tid; h < nh; h += blockDim.x) {
            if (compat[h]) {
                s_lpi[h] = digamma_pos(s_nk[h] + A.a0) - digamma_pos(ahat);
                s_pi[h] = log((A.a0 + s_nk[h]) / ((double)numah * A.a0 + (double)nr));
            } else {
                s_lpi[h] = -100.0;
                s_pi[h] = -100.0;
            }
        }
        __syncthreads();
        // ---- energy and stopping rule (:2484-2492, :2515), every thread on the same shared values ----
        double eNew = s_scal[1];
        for (int h = 0; h < nh; h++) eNew += s_pi[h] * s_nk[h];
        loglik = s_scal[0];
        const double dE = fabs(eOld - eNew);
        if (iter <= 25) margin = fmin(margin, fabs(dE - A.tol));
        const bool done = (dE < A.tol) || (iter > 25);
        eOld = eNew;
        iter++;
        if (done) break;
        __syncthreads();
    }
    if (tid == 0) {
        double zs = 0.0;
        for (int h = 0; h < nh; h++) zs += exp(s_pi[h]);
        const long long off = A.set_off[s];
        for (int h = 0; h < nh; h++) {
            A.pi[off + h] = s_pi[h];
            A.freqs[off + h] = exp(s_pi[h]) / zs;
        }
        A.loglik[s] = loglik;
        A.iters[s] = iter;
        if (A.margin) A.margin[s] = margin;
        if (A.energy) A.energy[s] = eOld;
    }
}

extern "C" cudaError_t k2_launch(const K2Args *a, int max_npairs, cudaStream_t st)
{
    if (a->n_jobs <= 0) return cudaSuccess;
    int threads = (max_npairs + 31) & ~31;
    if (threads < 32) threads = 32;
    if (threads > 256) threads = 256;
    // tile of reads: at most K2_TILE_BYTES of ll values + 1 keep byte per read
    k2_pair_sums<<<a->n_jobs, threads, K2_TILE_BYTES + 1024, st>>>(*a);
    return cudaGetLastError();
}

template <int NHP>
static cudaError_t k3_launch_nhp(const K3Args *a, int threads, size_t smem, bool stage, cudaStream_t st)
{
    if (stage) {
        cudaError_t e = cudaFuncSetAttribute(k3_bayes_em<NHP, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        k3_bayes_em<NHP, true><<<a->n_sets, threads, smem, st>>>(*a);
    } else {
        k3_bayes_em<NHP, false><<<a->n_sets, threads, 0, st>>>(*a);
    }
    return cudaGetLastError();
}

// max_nh / max_nr: over the windows the sets refer to; smem_optin: the device's opt-in limit per CTA
extern "C" cudaError_t k3_launch(const K3Args *a, int max_nh, int max_nr, size_t smem_optin, cudaStream_t st)
{
    if (a->n_sets <= 0) return cudaSuccess;
    // one CTA per set and (with the ll matrix staged in shared memory) one CTA per SM: the CTA's warps are all the
    // latency hiding there is (ncu at 256 threads: 8 warps per SM, `wait` 2.3 stalls per issue, FP64 pipe 24 %)
    int threads = max_nr <= 64 ? 64 : (max_nr <= 128 ? 128 : (max_nr <= 256 ? 256 : 512));
    const size_t need = (size_t)max_nr * (size_t)(max_nh | 1) * sizeof(double);
    const bool stage = need + 8192 <= smem_optin;
    if (max_nh <= 8) return k3_launch_nhp<8>(a, threads, need, stage, st);
    if (max_nh <= 16) return k3_launch_nhp<16>(a, threads, need, stage, st);
    if (max_nh <= 32) return k3_launch_nhp<32>(a, threads, need, stage, st);
    if (max_nh <= K3_MAXH) return k3_launch_nhp<64>(a, threads, need, stage, st);
    k3_bayes_em_wide<<<a->n_sets, 32 * K3_MAX_WARPS, 0, st>>>(*a); // a warp per read
    return cudaGetLastError();
}
extern "C" int k3_max_haps(void) { return K3W_MAXH; }
