// k23_inference.cu — the per-window inference kernels that consume the read x haplotype
// log-likelihood matrix IN PLACE (rl[r*nh+h] = pairs[win_pair[w] + r*nh + h].ll, the resident K1
// results; DInDel.cpp:2949-2960 / :2125-2139).
//
//   K2  pair sums   acc <- start; for r in window (in order): acc += log(.5) + addLogs(rl[r,h1], rl[r,h2])
//                   for every h1 <= h2: the inner loop of DetInDel::diploidGLF (DInDel.cpp:3085-3113,
//                   :3349-3390, start = 0 resp. prior - sitePrior) and of the pooled genotype
//                   likelihoods (DInDel.cpp:2683-2694, reads restricted to one pool member).
//                   One CTA per job, one thread per haplotype pair, reads summed sequentially in r
//                   like the CPU (same association; only exp/log differ from glibc at the ulp level).
//   K3  Bayes EM    the variational-Bayes EM over haplotype frequencies for one set of active
//                   variants (DetInDel::estimateHaplotypeFrequenciesBayesEM, DInDel.cpp:2431-2545):
//                   E-step responsibilities by a chain of addLogs in h order per read, counts nk,
//                   lpi = digamma(nk+a0) - digamma(sum), energy, <= 27 passes.  One CTA per set.
// Tolerance (stated in tests/test_k23_gpu.py): 1e-9 relative against the CPU restatement; the
// only arithmetic differences are CUDA libm vs glibc and the reduction tree over reads in K3.
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

__global__ void __launch_bounds__(128) k2_pair_sums(K2Args A)
{
    const int j = blockIdx.x;
    if (j >= A.n_jobs) return;
    const int w = A.job_window[j], grp = A.job_group[j];
    const int nh = A.win_hap[w + 1] - A.win_hap[w];
    const int r0 = A.win_read[w], nr = A.win_read[w + 1] - r0;
    const int npairs = nh * (nh + 1) / 2;
    const dindel_pair_out *p = A.pairs + A.win_pair[w];
    const long long off = A.job_off[j];
    for (int q = threadIdx.x; q < npairs; q += blockDim.x) {
        // q -> (h1 <= h2), row-major upper triangle
        int h1 = 0, rem = q;
        while (rem >= nh - h1) { rem -= nh - h1; h1++; }
        const int h2 = h1 + rem;
        double acc = A.start[off + q];
        for (int r = 0; r < nr; r++) {
            if (grp >= 0 && A.read_group && A.read_group[r0 + r] != grp) continue;
            const double a = p[(long long)r * nh + h1].ll, b = p[(long long)r * nh + h2].ll;
            acc += A.log_half + addlogs(a, b);
        }
        A.out[off + q] = acc;
    }
}

#define K3_THREADS 128
#define K3_MAXH 64

__global__ void __launch_bounds__(K3_THREADS) k3_bayes_em(K3Args A)
{
    __shared__ double s_lpi[K3_MAXH], s_pi[K3_MAXH], s_nk[K3_MAXH];
    __shared__ double s_part[K3_THREADS];
    __shared__ double s_loglik, s_B;
    __shared__ int s_done;
    const int s = blockIdx.x, tid = threadIdx.x;
    if (s >= A.n_sets) return;
    const int w = A.set_window[s];
    const int nh = A.win_hap[w + 1] - A.win_hap[w];
    const int nr = A.win_read[w + 1] - A.win_read[w];
    const dindel_pair_out *p = A.pairs + A.win_pair[w];
    const uint8_t *compat = A.compatible + A.set_off[s];
    int numah = 0;
    for (int h = 0; h < nh; h++) numah += compat[h] ? 1 : 0;
    if (tid < nh) s_lpi[tid] = compat[tid] ? log(1.0 / (double)numah) : -100.0; // DInDel.cpp:2417
    if (tid == 0) s_done = 0;
    __syncthreads();
    double eOld = K1_NEG_INF;
    int iter = 0;
    double z[K3_MAXH];
    for (;;) {
        // E-step over this thread's reads (DInDel.cpp:2439-2457)
        double nkp[K3_MAXH];
        for (int h = 0; h < nh; h++) nkp[h] = 0.0;
        double ll = 0.0, Bp = 0.0;
        for (int r = tid; r < nr; r += K3_THREADS) {
            double lognorm = K1_NEG_INF;
            for (int h = 0; h < nh; h++) {
                z[h] = s_lpi[h] + p[(long long)r * nh + h].ll;
                lognorm = addlogs(lognorm, z[h]);
            }
            for (int h = 0; h < nh; h++) {
                const double zz = exp(z[h] - lognorm);
                nkp[h] += zz;
                Bp += zz * p[(long long)r * nh + h].ll;
            }
            ll += lognorm;
        }
        // deterministic block reductions (fixed tree)
        for (int h = 0; h < nh; h++) {
            s_part[tid] = nkp[h];
            __syncthreads();
            for (int o = K3_THREADS / 2; o > 0; o >>= 1) {
                if (tid < o) s_part[tid] += s_part[tid + o];
                __syncthreads();
            }
            if (tid == 0) s_nk[h] = s_part[0];
            __syncthreads();
        }
        s_part[tid] = ll;
        __syncthreads();
        for (int o = K3_THREADS / 2; o > 0; o >>= 1) {
            if (tid < o) s_part[tid] += s_part[tid + o];
            __syncthreads();
        }
        if (tid == 0) s_loglik = s_part[0];
        __syncthreads();
        s_part[tid] = Bp;
        __syncthreads();
        for (int o = K3_THREADS / 2; o > 0; o >>= 1) {
            if (tid < o) s_part[tid] += s_part[tid + o];
            __syncthreads();
        }
        if (tid == 0) s_B = s_part[0];
        __syncthreads();
        // M-step (DInDel.cpp:2461-2480) and energy (:2484-2492), one thread
        if (tid == 0) {
            double ahat = 0.0;
            for (int h = 0; h < nh; h++)
                if (compat[h]) ahat += s_nk[h] + A.a0;
            const double dahat = digamma_pos(ahat);
            double eNew = s_B;
            for (int h = 0; h < nh; h++) {
                if (compat[h]) {
                    s_lpi[h] = digamma_pos(s_nk[h] + A.a0) - dahat;
                    s_pi[h] = log((A.a0 + s_nk[h]) / ((double)numah * A.a0 + (double)nr));
                } else {
                    s_lpi[h] = -100.0;
                    s_pi[h] = -100.0;
                }
                eNew += s_pi[h] * s_nk[h];
            }
            s_done = (fabs(eOld - eNew) < A.tol) || (iter > 25); // :2515
            s_part[0] = eNew;
        }
        __syncthreads();
        eOld = s_part[0];
        iter++;
        const int done = s_done;
        __syncthreads();
        if (done) break;
    }
    if (tid == 0) {
        double zs = 0.0;
        for (int h = 0; h < nh; h++) zs += exp(s_pi[h]); // :2524-2527
        const long long off = A.set_off[s];
        for (int h = 0; h < nh; h++) {
            A.pi[off + h] = s_pi[h];
            A.freqs[off + h] = exp(s_pi[h]) / zs; // :2539
        }
        A.loglik[s] = s_loglik;
        A.iters[s] = iter;
    }
}

extern "C" cudaError_t k2_launch(const K2Args *a, cudaStream_t st)
{
    if (a->n_jobs <= 0) return cudaSuccess;
    k2_pair_sums<<<a->n_jobs, 128, 0, st>>>(*a);
    return cudaGetLastError();
}
extern "C" cudaError_t k3_launch(const K3Args *a, cudaStream_t st)
{
    if (a->n_sets <= 0) return cudaSuccess;
    k3_bayes_em<<<a->n_sets, K3_THREADS, 0, st>>>(*a);
    return cudaGetLastError();
}
extern "C" int k3_max_haps(void) { return K3_MAXH; }
