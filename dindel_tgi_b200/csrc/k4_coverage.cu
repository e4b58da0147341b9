// k4_coverage.cu — K4: DetInDel::filterHaplotypes (DInDel.cpp:1932-2100) over the resident K1 results.
//
// One CTA per window.  Phase 1 (data parallel, one warp per (entry, read)): the read's hpos row is
// scanned by the 32 lanes (coalesced int16 loads); a base hits when its haplotype position lies in the
// entry's [left, right]; the read covers the entry when it hits every position with at most maxMismatch
// mismatches.  Only reads aligned on the haplotype without indels are tested (DInDel.cpp:1951), so the
// on-haplotype positions of a row are strictly increasing and "number of hits" equals the size of the
// reference's std::set of hit positions (:1998, :2038).  Phase 2 (one thread per haplotype): the
// reference's sequential walk over the haplotype's variants with its `break` at the first uncovered
// one (:2066-2071).  Phase 3: per key, union of the covering-read bit sets over the evaluated entries of
// unfiltered haplotypes, per strand (:2089-2097).  Integer / byte work only: HBM-bound on the hpos rows
// (R*2 bytes per tested (entry, read)); no tensor cores, no floating point.
#include "k4.cuh"

#define K4_THREADS 128
#define FULLMASK 0xffffffffu

__global__ void __launch_bounds__(K4_THREADS) k4_coverage_kernel(K4Args a)
{
    extern __shared__ uint32_t sm[];
    const int w = a.win_list ? a.win_list[blockIdx.x] : a.w_begin + blockIdx.x;
    const long long e0 = a.win_entry[w], e1 = a.win_entry[w + 1];
    const long long k0 = a.win_key[w], k1 = a.win_key[w + 1];
    const int ne = (int)(e1 - e0), nk = (int)(k1 - k0);
    const int h0 = a.win_hap[w], nh = a.win_hap[w + 1] - h0;
    const int r0 = a.win_read[w], nr = a.win_read[w + 1] - r0;
    const int words = (nr + 31) >> 5;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    // shared: cov[ne][2][words] | any[ne] | active[ne] | cnt[nk][2] | filt[nh]; cov in global memory for the windows
    // it would not fit (deep pooled windows: thousands of reads x hundreds of entries)
    const bool big = a.cov_global != nullptr;
    uint32_t *cov = big ? a.cov_global + a.cov_off[blockIdx.x] : sm;
    uint32_t *any = big ? sm : cov + (size_t)ne * 2 * words;
    uint32_t *active = any + ne;
    uint32_t *cnt = active + ne;
    uint32_t *filt = cnt + 2 * nk;
    for (int i = threadIdx.x; i < (big ? 0 : ne * 2 * words) + 2 * ne + 2 * nk + nh; i += blockDim.x) sm[i] = 0;
    __syncthreads();

    // ---- phase 1: which reads cover which entry ----
    const long long ntask = (long long)ne * nr;
    for (long long t = warp; t < ntask; t += nwarps) {
        const int e = (int)(t / nr), r = (int)(t - (long long)e * nr);
        const int type = a.entry_type[e0 + e];
        if (type != DINDEL_VAR_DEL && type != DINDEL_VAR_INS) continue;
        const int h = a.entry_hap[e0 + e];
        const long long pair = a.win_pair[w] + (long long)r * nh + h;
        const dindel_pair_out *po = a.pairs + pair;
        if (po->offHapHMQ || po->numIndels != 0 || po->status != DINDEL_WIN_OK) continue; // :1951
        const int left = a.entry_left[e0 + e], right = a.entry_right[e0 + e];
        const int len = right - left + 1;
        const long long roff = a.read_off[r0 + r];
        const int R = (int)(a.read_off[r0 + r + 1] - roff);
        const long long hoff = a.hap_off[h0 + h];
        const int H = (int)(a.hap_off[h0 + h + 1] - hoff);
        const int16_t *hp = a.hpos + pair * (long long)a.hpos_stride;
        int hits = 0, nmm = 0;
        for (int b0 = 0; b0 < R; b0 += 32) { // warp-uniform trip count (ballots below)
            const int b = b0 + lane;
            const int hb = (b < R) ? (int)hp[b] : -1;
            const bool in = hb >= 0 && hb >= left && hb <= right && hb < H;
            bool mm = false;
            if (in) {
                const uint8_t hc = a.hap_seq[hoff + hb], rc = a.read_seq[roff + b];
                mm = (hc != rc) && (type == DINDEL_VAR_INS || hc != 'N'); // :2000 vs :2041
            }
            hits += __popc(__ballot_sync(FULLMASK, in));
            nmm += __popc(__ballot_sync(FULLMASK, mm));
        }
        // DEL: |c| >= len && nmm <= max (:2004); INS: every position of [left,right] hit by this read and
        // nmm <= max (:2047-2060; for both lenins branches the read must hit all len positions)
        if (hits >= len && nmm <= a.maxMismatch && lane == 0) {
            const int strand = a.read_strand[r0 + r] ? 1 : 0;
            atomicOr(&cov[((size_t)e * 2 + strand) * words + (r >> 5)], 1u << (r & 31));
            any[e] = 1;
        }
    }
    __syncthreads();

    // ---- phase 2: per haplotype, walk its entries in map order; stop at the first uncovered indel ----
    for (int h = threadIdx.x; h < nh; h += blockDim.x) {
        bool allCovered = true;
        for (int e = 0; e < ne; e++) {
            if (a.entry_hap[e0 + e] != h) continue;
            if (!allCovered) break; // entries after the break are never evaluated (:2071)
            active[e] = 1;
            const int type = a.entry_type[e0 + e];
            if ((type == DINDEL_VAR_DEL || type == DINDEL_VAR_INS) && !any[e]) allCovered = false;
        }
        const uint8_t unc = allCovered ? 0 : 1;
        if (a.hap_uncovered) a.hap_uncovered[h0 + h] = unc;
        filt[h] = (a.doFilter && unc) ? 1u : 0u; // :2078-2083
        a.hap_filtered[h0 + h] = (uint8_t)filt[h];
    }
    __syncthreads();
    // ---- phase 3: per key and strand, distinct covering reads over unfiltered haplotypes ----
    for (int t = threadIdx.x; t < nk * 2 * words; t += blockDim.x) {
        const int k = t / (2 * words), s = (t / words) & 1, wd = t % words;
        uint32_t m = 0;
        for (int e = 0; e < ne; e++) {
            if (a.entry_key[e0 + e] != k || !active[e] || filt[a.entry_hap[e0 + e]]) continue;
            m |= cov[((size_t)e * 2 + s) * words + wd];
        }
        if (m) atomicAdd(&cnt[k * 2 + s], (uint32_t)__popc(m));
    }
    __syncthreads();
    for (int k = threadIdx.x; k < nk; k += blockDim.x) {
        a.key_nf[k0 + k] = (int32_t)cnt[k * 2 + 0];
        a.key_nr[k0 + k] = (int32_t)cnt[k * 2 + 1];
    }
}

extern "C" size_t k4_smem_bytes(int ne, int nk, int nr, int nh)
{
    const size_t words = (size_t)((nr + 31) >> 5);
    return ((size_t)ne * 2 * words + 2 * (size_t)ne + 2 * (size_t)nk + (size_t)nh) * sizeof(uint32_t);
}
// the same with the covering-read bit sets in global memory: what stays in shared memory, and the words of the sets
extern "C" size_t k4_smem_bytes_big(int ne, int nk, int nh) { return (2 * (size_t)ne + 2 * (size_t)nk + (size_t)nh) * sizeof(uint32_t); }
extern "C" size_t k4_cov_words(int ne, int nr) { return (size_t)ne * 2 * (size_t)((nr + 31) >> 5); }

extern "C" cudaError_t k4_launch(const K4Args *a, int n_windows, size_t smem, cudaStream_t st)
{
    if (n_windows <= 0) return cudaSuccess;
    cudaError_t e = cudaFuncSetAttribute(k4_coverage_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    k4_coverage_kernel<<<n_windows, K4_THREADS, smem, st>>>(*a);
    return cudaGetLastError();
}
