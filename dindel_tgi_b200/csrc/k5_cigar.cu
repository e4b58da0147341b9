// k5_cigar.cu — K5: realigned-read CIGARs (--outputRealignedBAM) over the resident K1 results.
//
// Replaces, per read, the haplotype choice at the two call sites of DetInDel::getCIGAR (DInDel.cpp:503-509 pooled,
// :597-606 diploid) and getCIGAR itself (DInDel.cpp:728-882).  One warp works on 32 consecutive reads:
//   choose  lane i resolves read i: its window, its row of pair records and the haplotype (two ll loads, or a strict
//           first-arg-max scan over the row): the 32 chains of dependent loads run side by side;
//   stage   for each of the 32 reads in turn the warp streams the chosen pair's hpos row out of HBM (coalesced int16
//           loads, the next read's row in flight meanwhile), maps every base through the haplotype's own reference
//           alignment hap_ref_pos (npos[b], :745-747; the gather hits L1 / L2, a haplotype's row is shared by all reads
//           of the window) and parks npos in the warp's shared-memory tile;
//   walk    lane i finds the first and the last base with npos >= 0 (:778-796) and runs the reference's sequential
//           CIGAR state machine (:806-862) on read i out of shared memory
//           (row pitch = an odd number of 32-bit words: the 32 lanes hit 32 different banks), writing BAM-encoded operations.
// Integer / byte work, HBM-bound on the hpos rows of the chosen pairs (2R bytes per read in, 4 * (3 + cap) out);
// no floating point beyond the ll comparison, no tensor cores.
#include "k5.cuh"

#define K5_WARPS 4
#define FULLMASK 0xffffffffu
#define OP_M 0u
#define OP_I 1u
#define OP_D 2u
#define OP_S 4u

__device__ __forceinline__ int k5_find_window(const int32_t *win_read, int nw, int r)
{
    int lo = 0, hi = nw; // win_read[lo] <= r < win_read[hi]
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (win_read[mid] <= r) lo = mid; else hi = mid;
    }
    return lo;
}

// first 128 bases of a read's hpos row: 4 independent coalesced int16 loads per lane (-5 = past the read / no row)
__device__ __forceinline__ void k5_load_row(const int16_t *hp, int R, int lane, int x[4])
{
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int b = 32 * k + lane;
        x[k] = (hp && b < R) ? (int)hp[b] : -5;
    }
}

__global__ void __launch_bounds__(K5_WARPS * 32) k5_cigar_kernel(K5Args a)
{
    extern __shared__ int32_t k5_sm[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int32_t *tile = k5_sm + (size_t)warp * 32 * a.pitch;
    const int ngroups = (a.n_reads + 31) >> 5;
    for (int g = blockIdx.x * K5_WARPS + warp; g < ngroups; g += gridDim.x * K5_WARPS) {
        const int r_base = a.r_begin + (g << 5);   // reads [r_begin, r_begin + n_reads) of the resident range
        const int r = r_base + lane;
        // ---- choose: lane i resolves read r_base + i (window, pair row, haplotype): 32 dependent-load chains at once ----
        int my_R = 0, my_hap = -1, my_ref0 = 0, my_H = 0;
        const int16_t *my_hp = nullptr;
        const int32_t *my_ref = nullptr;
        int w = a.w_begin + k5_find_window(a.win_read + a.w_begin, a.n_windows, r_base);   // same address in every lane: one search per warp
        if (r < a.r_begin + a.n_reads) {
            while (r >= a.win_read[w + 1]) w++;
            const int h0 = a.win_hap[w], nh = a.win_hap[w + 1] - h0;
            my_R = (int)(a.read_off[r + 1] - a.read_off[r]);
            my_ref0 = a.win_ref_start[w];
            const long long p0 = a.win_pair[w] + (long long)(r - a.win_read[w]) * nh;
            const dindel_pair_out *row = a.pairs + p0;
            int hx = -1;
            if (a.mode == DINDEL_CIGAR_PAIR) {
                const int h1 = a.win_h1[w], h2 = a.win_h2[w];
                if (h1 >= 0 && h1 < nh && h2 >= 0 && h2 < nh) {
                    const double l1 = row[h1].ll, l2 = row[h2].ll;
                    if (fabs(l1 - l2) < 1e-8) hx = a.hap_n_indels[h0 + h1] < a.hap_n_indels[h0 + h2] ? h1 : h2; // :598-599
                    else hx = l1 > l2 ? h1 : h2;                                                                  // :601-605
                }
            } else if (a.on_hap[r]) {                                                                             // :502
                double best = -HUGE_VAL;                                                                          // :503-508
                hx = 0;
                for (int h = 0; h < nh; h++) {
                    const double v = row[h].ll;
                    if (v > best) { best = v; hx = h; }
                }
                if (nh <= 0) hx = -1;
            }
            if (hx >= 0) {
                const long long ho = a.hap_off[h0 + hx];
                my_H = (int)(a.hap_off[h0 + hx + 1] - ho);
                my_ref = a.hap_ref_pos + ho;
                my_hp = a.hpos + (size_t)(p0 + hx) * a.hpos_stride;
                my_hap = hx;
            }
        }
        // ---- stage: the warp streams read i's hpos row (coalesced), maps it through the haplotype's reference alignment
        //      (npos[b], :745-747) and parks it in row i of its tile; read i + 1's row is in flight while read i is mapped ----
        int xn[4];
        {
            const int16_t *hp0 = (const int16_t *)__shfl_sync(FULLMASK, (unsigned long long)my_hp, 0);
            k5_load_row(hp0, __shfl_sync(FULLMASK, my_R, 0), lane, xn);
        }
        for (int i = 0; i < 32; i++) {
            const int16_t *hp = (const int16_t *)__shfl_sync(FULLMASK, (unsigned long long)my_hp, i);
            const int32_t *ref = (const int32_t *)__shfl_sync(FULLMASK, (unsigned long long)my_ref, i);
            const int R = __shfl_sync(FULLMASK, my_R, i), H = __shfl_sync(FULLMASK, my_H, i);
            int x[4];
#pragma unroll
            for (int k = 0; k < 4; k++) x[k] = xn[k];
            if (i < 31) {
                const int16_t *hpn = (const int16_t *)__shfl_sync(FULLMASK, (unsigned long long)my_hp, i + 1);
                k5_load_row(hpn, __shfl_sync(FULLMASK, my_R, i + 1), lane, xn);
            }
            if (!hp) continue;                                   // warp-uniform: skipped read / past the batch
            int32_t *trow = tile + i * a.pitch;
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const int b = 32 * k + lane;
                if (b < R) trow[b] = x[k] >= 0 ? (x[k] < H ? ref[x[k]] : -5) : x[k];
            }
            for (int b = 128 + lane; b < R; b += 32) {           // reads longer than 128 bases
                const int xb = hp[b];
                trow[b] = xb >= 0 ? (xb < H ? ref[xb] : -5) : xb;
            }
        }
        __syncwarp();
        // first / last base on the haplotype and on the reference (:778-796), by the lane that owns the read
        const int my_skip = my_hap < 0;
        int my_first = my_R, my_last = -1;
        if (!my_skip) {
            const int32_t *np = tile + lane * a.pitch;
            int b = my_R - 1;
            while (b >= 0 && np[b] < 0) b--;
            my_last = b;
            if (b >= 0) { b = 0; while (np[b] < 0) b++; my_first = b; }
        }
        // ---- walk: lane i owns read r_base + i ----
        if (r < a.r_begin + a.n_reads) {
            int n = 0, ref_pos = -1;
            uint32_t *ops = a.ops + (size_t)r * a.cap;
#define K5_PUSH(op, len) do { if (n < a.cap) ops[n] = ((uint32_t)(len) << 4) | (op); n++; } while (0)
            if (!my_skip) {
                const int32_t *np = tile + lane * a.pitch;
                if (my_last < 0) {
                    K5_PUSH(OP_S, my_R);                                         // :782-786
                } else {
                    int b = my_first;
                    if (b > 0) K5_PUSH(OP_S, b);                                 // :796
                    int prevponr = np[b];
                    ref_pos = my_ref0 + prevponr;                                // :798
                    unsigned curr = OP_M; int len = 1, err = 0;
                    int chp = prevponr;
                    while (b < my_last) {                                        // :806-862
                        const int nhp = np[b + 1];
                        if (nhp == DINDEL_HPOS_INS) {
                            if (chp == DINDEL_HPOS_INS) {
                                if (curr != OP_I) { err = -1; break; }
                                len++;
                            } else if (chp >= 0) {
                                if (curr != OP_M) { err = -2; break; }
                                K5_PUSH(OP_M, len);
                                len = 1; curr = OP_I; prevponr = chp;
                            } else { err = -5; break; }
                        } else if (chp >= 0 && nhp >= 0 && nhp - chp == 1) {
                            if (curr != OP_M) { err = -3; break; }
                            len++; prevponr = nhp;
                        } else if (chp >= 0 && nhp >= 0 && nhp - chp > 1) {
                            if (curr != OP_M) { err = -4; break; }
                            K5_PUSH(OP_M, len);
                            K5_PUSH(OP_D, nhp - chp - 1);
                            curr = OP_M; len = 1; prevponr = nhp;
                        } else if (chp == DINDEL_HPOS_INS && nhp - prevponr == 1) {
                            K5_PUSH(OP_I, len);
                            curr = OP_M; len = 1; prevponr = nhp;
                        } else if (chp == DINDEL_HPOS_INS && nhp - prevponr > 1) {
                            K5_PUSH(OP_I, len);
                            K5_PUSH(OP_D, nhp - prevponr - 1);
                            curr = OP_M; len = 1; prevponr = nhp;
                        }
                        chp = nhp;
                        b++;
                    }
                    if (err) {
                        n = err;
                    } else {
                        K5_PUSH(curr, len);                                      // :865
                        if (my_R - 1 - my_last > 0) K5_PUSH(OP_S, my_R - 1 - my_last);
                    }
                }
            }
#undef K5_PUSH
            a.read_hap[r] = my_hap;
            a.read_ref_pos[r] = ref_pos;
            a.read_n_ops[r] = n;
        }
        __syncwarp();
    }
}

extern "C" size_t k5_smem_bytes(int pitch) { return (size_t)K5_WARPS * 32 * pitch * sizeof(int32_t); }

// row pitch in 32-bit words: >= R and odd (bank-conflict-free column walks)
extern "C" int k5_pitch(int max_R) { return (max_R < 1 ? 1 : max_R) | 1; }

extern "C" cudaError_t k5_launch(const K5Args *a, int n_sms, size_t smem, cudaStream_t st)
{
    if (a->n_reads <= 0) return cudaSuccess;
    cudaError_t e = cudaFuncSetAttribute(k5_cigar_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    int per_sm = 1;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k5_cigar_kernel, K5_WARPS * 32, smem);
    if (e != cudaSuccess) return e;
    if (per_sm < 1) per_sm = 1;
    const int ngroups = (a->n_reads + 31) >> 5;
    int grid = n_sms * per_sm;
    const int need = (ngroups + K5_WARPS - 1) / K5_WARPS;
    if (grid > need) grid = need;
    k5_cigar_kernel<<<grid, K5_WARPS * 32, smem, st>>>(*a);
    return cudaGetLastError();
}
